#!/bin/bash
# One GPU session: parity tests, GEMM timing ladder (incl. debug modes), quantiser bandwidth probe, bench.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
for dbg in 0 1 2 3; do
  python tools/gemm_time.py 32768 4096 4096 128 0.05 10 $dbg
done > gpurun_out/gemm_time.log 2>&1
python tools/gemm_time.py 32768 4096 4096 1024 0.05 10 0 >> gpurun_out/gemm_time.log 2>&1
python tools/gemm_time.py 32768 4096 4096 64 0.05 10 0 >> gpurun_out/gemm_time.log 2>&1
python tools/gemm_time.py 32768 11008 4096 128 0.05 10 0 >> gpurun_out/gemm_time.log 2>&1
python tools/gemm_time.py 32768 4096 4096 128 0.0 10 0 >> gpurun_out/gemm_time.log 2>&1
python tools/k1_probe.py > gpurun_out/k1_probe.log 2>&1
python bench.py --steps 3 --warmup 3 > gpurun_out/bench.log 2>&1
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/gemm_time.log; tail -30 gpurun_out/k1_probe.log; tail -2 gpurun_out/bench.log
