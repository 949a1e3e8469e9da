#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 300 python tools/k1_probe.py > gpurun_out/k1_probe.log 2>&1
grep -E "^K1 .*\"opt\": \{\}|K1deq|K0 |copy|fault" gpurun_out/k1_probe.log
