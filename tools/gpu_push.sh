#!/bin/bash
# 2-GPU check of the push / multicast peer modes with tight timeouts (a protocol bug shows up as a stream that never advances)
mkdir -p gpurun_out
timeout 120 python -m pytest tests/test_parallel.py -m gpu -q -x -k nccl 2>&1 | tail -4
if [ ${PIPESTATUS[0]} -ne 0 ]; then echo "TEST FAILED OR TIMED OUT"; exit 0; fi
for mode in push mcast pull; do
  timeout 100 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 \
    bench.py --gpus 2 --steps 3 --warmup 3 --layers 8 --parallel tp --transport peer --peer-mode $mode > gpurun_out/push_$mode.json 2> gpurun_out/push_$mode.err
  echo "[$mode] rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/push_$mode.json').read().strip().splitlines()[-1]); c=d['column_parallel']
    print('  tok/s',round(d['value']),'ms',round(d['ms_per_step'],2),'| serial',c['serialised'],'| exposed',round(c['overlapped']['exposed_comm_ms'],2),'| lim',round(c['limiting_collective']['algbw_GBs']),'GB/s')
except Exception as e:
    print('  ERR',e); print(open('gpurun_out/push_$mode.err').read()[-600:])
PY
done
