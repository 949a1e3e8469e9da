#!/bin/bash
# bash tools/gpu_vars.sh [SHAPE...]: time the default library and every tools/libmq_*.so variant (MQ_LIB override)
# on the given "M N K g" shapes (default: the 7B prefill q_proj shape), with the wait-cycle profile.
mkdir -p gpurun_out
SHAPES=("$@"); [ ${#SHAPES[@]} -eq 0 ] && SHAPES=("32768 4096 4096 128")
for lib in smoothquant_mixedprecision_b200/libmq.so tools/libmq_*.so; do
  echo "== $lib"
  for shape in "${SHAPES[@]}"; do
    MQ_LIB=$PWD/$lib MQ_PROFILE=1 timeout 120 python tools/gemm_time.py $shape 0.05 20 0 2>&1 | grep -E "GEMM|PROF|rror" | cut -c1-330
  done
done 2>&1 | tee gpurun_out/vars.log
