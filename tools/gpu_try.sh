#!/bin/bash
# bash tools/gpu_try.sh VARIANT...: for each tools/libmq_VARIANT.so run the GPU parity suite and the GEMM timing ladder
cp smoothquant_mixedprecision_b200/libmq.so /tmp/libmq_orig.so
for v in "$@"; do
  cp tools/libmq_$v.so smoothquant_mixedprecision_b200/libmq.so
  echo "=== $v"
  timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
  for shape in "32768 4096 4096 128" "32768 11008 4096 128" "32768 4096 11008 128" "32768 4096 4096 1024" "32768 4096 4096 256" "2048 4096 4096 128"; do
    timeout 60 python tools/gemm_time.py $shape 0.05 10 0 2>&1 | grep GEMM | cut -c1-150
  done
done
cp /tmp/libmq_orig.so smoothquant_mixedprecision_b200/libmq.so
