// Host-side TMA plumbing shared by the GEMM kernels: cached cuTensorMapEncodeTiled descriptors and the one-time
// dynamic-shared-memory attribute of a kernel.
#pragma once
#include <cuda.h>
#include <stdint.h>

// 2-D row-major tensor [rows, cols] of `esize`-byte elements, leading dimension ld (elements).  box_cols == 0: operand
// tile box of 128 bytes x box_rows rows with the 128-byte swizzle; otherwise a plain (unswizzled) box of box_cols
// elements x box_rows rows.
int mq_make_map(CUtensorMap* map, const void* ptr, CUtensorMapDataType dt, int esize, int64_t rows, int64_t cols,
                int64_t ld, int box_rows, int box_cols = 0);
// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) once per (kernel, device)
int mq_ensure_smem_attr(const void* kern, int bytes);
