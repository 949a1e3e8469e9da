#!/bin/bash
cp smoothquant_mixedprecision_b200/libmq.so /tmp/libmq_orig.so
echo "== default"; timeout 60 python tools/gemm_time.py 32768 4096 4096 128 0.05 10 0 2>&1 | grep GEMM | cut -c1-150
for f in tools/libmq_*.so; do
  cp $f smoothquant_mixedprecision_b200/libmq.so
  echo "== $f"; timeout 60 python tools/gemm_time.py 32768 4096 4096 128 0.05 10 0 2>&1 | grep GEMM | cut -c1-150
done
cp /tmp/libmq_orig.so smoothquant_mixedprecision_b200/libmq.so
