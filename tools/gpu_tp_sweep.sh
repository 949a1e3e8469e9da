#!/bin/bash
# bash tools/gpu_tp_sweep.sh N "args1" "args2" ... : tp-only bench variants on N GPUs (run with gpurun --gpus N)
N=$1; shift
mkdir -p gpurun_out
i=0
for A in "$@"; do
  i=$((i+1))
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29520+i)) \
    bench.py --gpus $N --steps 5 --warmup 3 --parallel tp $A > gpurun_out/tp${N}_v${i}.json 2> gpurun_out/tp${N}_v${i}.err; echo "[$A] rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/tp${N}_v${i}.json').read().strip().splitlines()[-1])
    c=d['column_parallel']
    print('  tok/s',round(d['value']),'ms',round(d['ms_per_step'],2),'e2e',round(d['e2e']['value']),'| serial',c['serialised'],'| exposed',round(c['overlapped']['exposed_comm_ms'],2),'| lim',round(c['limiting_collective']['algbw_GBs']),'GB/s', c['limiting_collective']['bytes_received_per_rank'])
except Exception as e:
    print('ERR',e); print(open('gpurun_out/tp${N}_v${i}.err').read()[-1500:])
PY
done
