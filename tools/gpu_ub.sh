#!/bin/bash
mkdir -p gpurun_out
timeout 120 ./tools/ubench > gpurun_out/ubench.log 2>&1; echo "rc=$?" >> gpurun_out/ubench.log
for dbg in 4 6 5; do
  timeout 120 python tools/gemm_time.py 32768 4096 4096 128 0.05 10 $dbg
done > gpurun_out/gemm_time2.log 2>&1
cat gpurun_out/ubench.log gpurun_out/gemm_time2.log
