#!/bin/bash
mkdir -p gpurun_out
CMD="python tools/gemm_time.py 32768 4096 4096 128 0.05 2 0"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:mixed_gemm -s 2 -c 1 -f -o gpurun_out/prof_gemm $CMD > gpurun_out/ncu.log 2>&1
tail -5 gpurun_out/plain.log gpurun_out/ncu.log
ls -la gpurun_out/
