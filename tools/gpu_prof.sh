#!/bin/bash
mkdir -p gpurun_out
export MQ_PROFILE=1
for cfg in "128 0" "128 1" "128 2" "128 4" "128 6" "1024 0"; do
  set -- $cfg
  timeout 120 python tools/gemm_time.py 32768 4096 4096 $1 0.05 10 $2
done > gpurun_out/gemm_prof.log 2>&1
cat gpurun_out/gemm_prof.log
