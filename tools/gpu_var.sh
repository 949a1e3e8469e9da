#!/bin/bash
cp smoothquant_mixedprecision_b200/libmq.so /tmp/libmq_orig.so
echo "== default"; timeout 60 python tools/k1_time.py 32768 4096 128 0.05 2>&1 | grep K1T
for f in tools/libmq_*.so; do
  cp $f smoothquant_mixedprecision_b200/libmq.so
  echo "== $f"; timeout 60 python tools/k1_time.py 32768 4096 128 0.05 2>&1 | grep K1T
done
cp /tmp/libmq_orig.so smoothquant_mixedprecision_b200/libmq.so
