#!/bin/bash
# CTA-pair kernel vs single-CTA kernel at the 7B prefill shapes
mkdir -p gpurun_out
{
for pair in 2 1; do
  for cfg in "128 1.0" "1024 1.0" "256 1.0" "64 1.0" "64 0.0" "128 0.0" "128 0.7" "128 0.5"; do
    set -- $cfg
    MQ_PAIR=$pair MQ_INT_FRACTION=$2 timeout 60 python tools/gemm_time.py 32768 4096 4096 $1 0.05 10 0 2>&1 | grep GEMM | cut -c1-190
  done
  MQ_PAIR=$pair timeout 60 python tools/gemm_time.py 32768 11008 4096 128 0.05 10 0 2>&1 | grep GEMM | cut -c1-190
  MQ_PAIR=$pair timeout 60 python tools/gemm_time.py 32768 4096 11008 128 0.05 10 0 2>&1 | grep GEMM | cut -c1-190
  MQ_PAIR=$pair timeout 60 python tools/gemm_time.py 2048 4096 4096 128 0.05 20 0 2>&1 | grep GEMM | cut -c1-190
  MQ_PAIR=$pair timeout 60 python tools/gemm_time.py 512 4096 4096 128 0.05 20 0 2>&1 | grep GEMM | cut -c1-190
done
MQ_PAIR=1 MQ_PROFILE=1 timeout 60 python tools/gemm_time.py 32768 4096 4096 128 0.05 10 0 2>&1 | grep PROF
MQ_PAIR=1 MQ_PROFILE=1 timeout 60 python tools/gemm_time.py 32768 4096 4096 1024 0.05 10 0 2>&1 | grep PROF
} | tee gpurun_out/pair_sweep.txt
