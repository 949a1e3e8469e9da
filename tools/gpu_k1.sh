#!/bin/bash
# K1 parity (ops tests) + timing of the h2 kernel against the r01 kernel at the bench shapes
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_ops_gpu.py tests/test_layer_gpu.py -m gpu -q -x 2>&1 | tail -15
for args in "32768 4096 128 0.05" "32768 11008 128 0.05" "32768 8192 128 0.1" "32768 5120 128 0.05" "32768 4096 64 0.1" "32768 4096 1024 0.05"; do
  MQ_H2=2 timeout 60 python tools/k1_time.py $args 2>&1 | grep -E "K1T|rror"
  timeout 60 python tools/k1_time.py $args 2>&1 | grep -E "K1T|rror"
  timeout 60 python tools/k1_time.py $args 1 2>&1 | grep -E "K1T|rror"
done 2>&1 | tee gpurun_out/k1_times.log
