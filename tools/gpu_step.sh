#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/gemm_probe.py > gpurun_out/gemm_probe.log 2>&1
for dbg in 0 1 2; do
  timeout 120 python tools/gemm_time.py 32768 4096 4096 128 0.05 10 $dbg
done > gpurun_out/gemm_time3.log 2>&1
timeout 120 python tools/gemm_time.py 32768 4096 4096 64 0.05 10 0 >> gpurun_out/gemm_time3.log 2>&1
timeout 120 python tools/gemm_time.py 32768 4096 4096 1024 0.05 10 0 >> gpurun_out/gemm_time3.log 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
cat gpurun_out/gemm_probe.log gpurun_out/gemm_time3.log; tail -5 gpurun_out/pytest_gpu.log
