#!/bin/bash
mkdir -p gpurun_out
export MQ_PROFILE=1
for d in 0 1 6 12; do timeout 120 python tools/gemm_time.py 32768 4096 4096 128 0.05 10 $d; done > gpurun_out/gemm_dbg.log 2>&1
timeout 120 python tools/gemm_time.py 32768 4096 4096 1024 0.05 10 0 >> gpurun_out/gemm_dbg.log 2>&1
timeout 120 python tools/gemm_time.py 32768 4096 4096 64 0.05 10 0 >> gpurun_out/gemm_dbg.log 2>&1
grep -h "GEMM\|PROF\|rror" gpurun_out/gemm_dbg.log | cut -c1-330
