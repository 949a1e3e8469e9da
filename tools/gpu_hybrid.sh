#!/bin/bash
# hybrid engine: parity tests, then GEMM time vs share of integer groups at the 7B prefill shape
mkdir -p gpurun_out
{
for g in 128 64 256; do
  for f in 1.0 0.9 0.8 0.7 0.6 0.5 0.4 0.3 0.0; do
    MQ_INT_FRACTION=$f timeout 60 python tools/gemm_time.py 32768 4096 4096 $g 0.05 10 0 2>&1 | grep GEMM | cut -c1-170
  done
done
for f in 1.0 0.7 0.6; do
  MQ_INT_FRACTION=$f timeout 60 python tools/gemm_time.py 32768 11008 4096 128 0.05 10 0 2>&1 | grep GEMM | cut -c1-170
  MQ_INT_FRACTION=$f timeout 60 python tools/gemm_time.py 32768 4096 11008 128 0.05 10 0 2>&1 | grep GEMM | cut -c1-170
done
} | tee gpurun_out/hybrid_sweep.txt
