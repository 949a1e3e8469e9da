#!/bin/bash
# bash tools/gpu_quick.sh [pytest -k expr]: GPU parity suite (optionally filtered) + one bench line
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q 2>&1 | tail -8      # alone, in a fresh process
if [ -n "$1" ]; then timeout 1200 python -m pytest tests -m gpu -q -k "$1" 2>&1 | tail -25; else timeout 2400 python -m pytest tests -m gpu -q 2>&1 | tail -25; fi
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/q_bench.json 2> gpurun_out/q_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
try:
    d=json.loads(open('gpurun_out/q_bench.json').read().strip().splitlines()[-1])
    print({k:d[k] for k in ('value','ms_per_step')}, 'e2e',d['e2e']['value'])
    print('gemm',d['roofline']['achieved'],d['roofline']['frac'],d['roofline']['share_of_step'],d['roofline']['fractions'])
    print('quant',d['roofline_quantiser']['achieved'],d['roofline_quantiser']['frac'],d['roofline_quantiser']['share_of_step'])
    print(d.get('gemm_tops')); print(d.get('gpu_reference')); print(d.get('fp16_cublas')); print(d['clocks'])
    print(d.get('decode'))
except Exception as e: print('ERR',e); print(open('gpurun_out/q_bench.err').read()[-2000:])
PY
