#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -k "norm_quant or fused_norm" 2>&1 | tail -25
for K in ; do
for N in "" rms layer unfused unfused_layer; do MQ_NORM=$N timeout 120 python tools/k1_time.py 32768 $K 128 0.05 2>&1 | grep K1T; done
done
