#!/bin/bash
# bash tools/gpu_tp.sh N [transport]: N-GPU column-parallel check + bench (run with gpurun --gpus N)
N=${1:-2}
T=${2:-peer}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_parallel.py -m gpu -q -x -k "nccl" 2>&1 | tail -25
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
  bench.py --gpus $N --steps 5 --warmup 3 --transport $T > gpurun_out/tp${N}_${T}_bench.json 2> gpurun_out/tp${N}_${T}_bench.err; echo "bench rc=$?"
python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/tp${N}_${T}_bench.json').read().strip().splitlines()[-1])
    print({k:d[k] for k in ('value','ms_per_step','n_gpus')}, 'e2e',d['e2e']['value'])
    print(json.dumps(d.get('column_parallel'),indent=1)); print(d.get('data_parallel'))
except Exception as e: print('ERR',e)
PY
tail -15 gpurun_out/tp${N}_${T}_bench.err
