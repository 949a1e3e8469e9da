#!/bin/bash
# tools/build_variant_quant.sh NAME [-DFLAG ...]: like build_variant.sh, for quant_rows.cu
set -e
cd "$(dirname "$0")/.."
name=$1; shift
P=smoothquant_mixedprecision_b200
mkdir -p /tmp/mqvar
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr \
  -Wno-deprecated-gpu-targets -Xptxas -v "$@" -c $P/csrc/quant_rows.cu -o /tmp/mqvar/quant_$name.o 2> /tmp/mqvar/ptxas_q$name.log
nvcc -shared -o tools/libmq_$name.so $P/build/common.o $P/build/colstat.o /tmp/mqvar/quant_$name.o $P/build/mixed_gemm.o \
  -Xcompiler -fPIC -Wno-deprecated-gpu-targets
echo "built tools/libmq_$name.so"
