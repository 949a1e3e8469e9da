#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/gemm_probe.py > gpurun_out/gemm_probe.log 2>&1
grep -v '"nbad": 0' gpurun_out/gemm_probe.log | head; echo "probe ok cases: $(grep -c '"nbad": 0' gpurun_out/gemm_probe.log)"
export MQ_PROFILE=1
for cfg in "128 0" "1024 0" "64 0"; do
  set -- $cfg
  timeout 120 python tools/gemm_time.py 32768 4096 4096 $1 0.05 10 $2
done > gpurun_out/gemm_prof.log 2>&1
timeout 120 python tools/gemm_time.py 32768 11008 4096 128 0.05 10 0 >> gpurun_out/gemm_prof.log 2>&1
timeout 120 python tools/gemm_time.py 32768 4096 11008 128 0.05 10 0 >> gpurun_out/gemm_prof.log 2>&1
cat gpurun_out/gemm_prof.log
