#!/bin/bash
# tools/build_variant.sh NAME [-DFLAG ...]: builds tools/libmq_NAME.so with extra nvcc flags on mixed_gemm.cu only
# (the other objects come from the regular in-tree build).  Timing experiments only; run them with tools/gpu_var.sh.
set -e
cd "$(dirname "$0")/.."
name=$1; shift
P=smoothquant_mixedprecision_b200
mkdir -p /tmp/mqvar
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr \
  -Wno-deprecated-gpu-targets -Xptxas -v "$@" -c $P/csrc/mixed_gemm.cu -o /tmp/mqvar/gemm_$name.o 2> /tmp/mqvar/ptxas_$name.log
nvcc -shared -o tools/libmq_$name.so $P/build/common.o $P/build/colstat.o $P/build/quant_rows.o /tmp/mqvar/gemm_$name.o \
  -Xcompiler -fPIC -Wno-deprecated-gpu-targets
echo "built tools/libmq_$name.so"
