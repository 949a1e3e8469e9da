#!/bin/bash
# one --set full capture of the mixed GEMM at the 7B prefill shape: bash tools/gpu_ncu.sh <name> [g]
mkdir -p gpurun_out
NAME=${1:-gemm}; G=${2:-128}
CMD="python tools/gemm_time.py 32768 4096 4096 $G 0.05 2 0"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:mixed_gemm -s 2 -c 1 -f -o gpurun_out/$NAME $CMD > gpurun_out/ncu.log 2>&1
tail -2 gpurun_out/plain.log; tail -2 gpurun_out/ncu.log
