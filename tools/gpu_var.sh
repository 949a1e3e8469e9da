#!/bin/bash
# default build: probe + timing; then prebuilt experimental variants of libmq.so (tools/libmq_*.so), timing only
timeout 100 python tools/gemm_probe.py > gpurun_out/gemm_probe.log 2>&1
echo "probe ok cases: $(grep -c '"nbad": 0' gpurun_out/gemm_probe.log) of $(grep -c CASE gpurun_out/gemm_probe.log)"
timeout 60 python tools/gemm_time.py 32768 4096 4096 128 0.05 10 0 2>&1 | grep GEMM | cut -c1-150
cp smoothquant_mixedprecision_b200/libmq.so /tmp/libmq_orig.so
for f in tools/libmq_*.so; do
  cp $f smoothquant_mixedprecision_b200/libmq.so
  echo "== $f"; timeout 60 python tools/gemm_time.py 32768 4096 4096 128 0.05 10 0 2>&1 | grep GEMM | cut -c1-150
done
cp /tmp/libmq_orig.so smoothquant_mixedprecision_b200/libmq.so
