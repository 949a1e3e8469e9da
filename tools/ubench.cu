// Micro-benchmarks that size the mixed GEMM's design on a B200 (not part of libmq.so):
//   mma    back-to-back tcgen05.mma (i8 / f8f6f4 / f16, 128 x 256 x 32 B) on every SM, no loads, no epilogue
//   ldtm   TMEM -> register drain of a 128 x 256 accumulator with 4 or 8 warps (x16 double-buffered / 4 x x32 batched)
//   alu    the per-group rescale arithmetic of the epilogue on register data (i8 magic-number path vs f32 path)
//   exact  kind::f8f6f4 (E4M3 codes) vs kind::i8 on the same code bytes: are the fp32 sums exactly the int32 sums?
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/ubench tools/ubench.cu
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include <vector>

#include "../smoothquant_mixedprecision_b200/csrc/ptx.cuh"

using namespace mqptx;

#define CK(x)                                                                      \
  do {                                                                             \
    cudaError_t e = (x);                                                           \
    if (e != cudaSuccess) {                                                        \
      printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); \
      exit(1);                                                                     \
    }                                                                              \
  } while (0)

// ------------------------------------------------------------------------------------------------ mma
// kind: 0 = i8, 1 = f8f6f4 (E4M3), 2 = f16
__global__ void __launch_bounds__(128, 1) k_mma(int kind, int iters, long long* cycles, int fill) {
  extern __shared__ unsigned char smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  unsigned char* smem = smem_raw + (((raw + 1023u) & ~1023u) - raw);
  __shared__ uint64_t bar;
  __shared__ uint32_t tslot;
  for (int i = threadIdx.x; i < (16 + 32) * 1024 / 4; i += blockDim.x)
    reinterpret_cast<uint32_t*>(smem)[i] = fill ? 0x38383838u : 0u;  // E4M3 1.0 bytes / int8 56 / small f16
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    fence_mbar_init();
  }
  fence_proxy_async_smem();
  if (threadIdx.x < 32) tmem_alloc<512>(&tslot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = tslot;
  if (threadIdx.x == 0) {
    const uint32_t idesc = kind == 0 ? umma_idesc(2u, 1u, 128, 256) : kind == 1 ? umma_idesc(1u, 0u, 128, 256)
                                                                                 : umma_idesc(1u, 0u, 128, 256);
    const uint32_t sa = smem_u32(smem);
    const uint64_t da = umma_desc_k128(sa), db = umma_desc_k128(sa + 16 * 1024);
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
      const uint32_t d = tm + (i & 1) * 256;
#pragma unroll
      for (int s = 0; s < 4; ++s) {
        if (kind == 0)
          mma_i8_ss(d, da + 2 * s, db + 2 * s, idesc, 1);
        else if (kind == 1)
          mma_f8_ss(d, da + 2 * s, db + 2 * s, idesc, 1);
        else
          mma_f16_ss(d, da + 2 * s, db + 2 * s, idesc, 1);
      }
    }
    tc_commit(&bar);
    mbar_wait(&bar, 0, 0x900);
    const long long t1 = clock64();
    cycles[blockIdx.x] = t1 - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) {
    tc_fence_after();
    tmem_dealloc<512>(tm);
  }
}

// ------------------------------------------------------------------------------------------------ ldtm
// variant 0: x16 loads, double buffered, wait per load (the GEMM's current pattern); variant 1: 4 x x32 then one wait
template <int nwarps, int variant>
__global__ void __launch_bounds__(384, 1) k_ldtm(int iters, long long* cycles, uint32_t* sink) {
  __shared__ uint32_t tslot;
  const int warp = threadIdx.x >> 5;
  if (warp == 2) tmem_alloc<512>(&tslot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = tslot;
  uint32_t acc = 0;
  if (warp >= 4 && warp < 4 + nwarps) {
    const int ew = warp - 4, quarter = warp & 3;
    constexpr int cw = 256 / (nwarps / 4);  // columns per thread
    const int col0 = (ew >> 2) * cw;
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      const uint32_t taddr = tm + ((uint32_t)(quarter * 32) << 16) + (it & 1) * 256 + col0;
      if (variant == 0) {
        uint32_t v[2][16];
        tmem_ld_32x16(taddr, v[0]);
#pragma unroll
        for (int c = 0; c < cw / 16; ++c) {
          tmem_ld_wait();
          if (c + 1 < cw / 16) tmem_ld_32x16(taddr + (c + 1) * 16, v[(c + 1) & 1]);
#pragma unroll
          for (int j = 0; j < 16; ++j) acc ^= v[c & 1][j];
        }
      } else {
#pragma unroll
        for (int c = 0; c < cw / 128; ++c) {
          uint32_t v[4][32];
#pragma unroll
          for (int q = 0; q < 4; ++q) tmem_ld_32x32(taddr + c * 128 + q * 32, v[q]);
          tmem_ld_wait();
#pragma unroll
          for (int q = 0; q < 4; ++q)
#pragma unroll
            for (int j = 0; j < 32; ++j) acc ^= v[q][j];
        }
      }
    }
    const long long t1 = clock64();
    if ((threadIdx.x & 31) == 0 && warp == 4) cycles[blockIdx.x] = t1 - t0;
  }
  if (acc == 0x12345678u) sink[0] = acc;
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<512>(tm);
  }
}

// ------------------------------------------------------------------------------------------------ alu
// 8 warps, each thread rescales 128 values per iteration (= one K group of a 128 x 256 tile per CTA).
// path 0: int32 -> magic add -> f32x2 sub -> f32x2 mul (scale product) -> f32x2 fma     (kind::i8 accumulators)
// path 1: f32x2 mul (scale product) -> f32x2 fma                                         (fp32 accumulators)
// path 2: like 1 but the scale product is one scalar mul per pair hoisted: fma(v*as, ws, acc) -> mul2 + fma2 (same)
template <int path>
__global__ void __launch_bounds__(256, 1) k_alu(int iters, long long* cycles, float* sink, const float* ws_g) {
  float2 acc[64];
#pragma unroll
  for (int j = 0; j < 64; ++j) acc[j] = make_float2(0.f, 0.f);
  uint32_t v[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) v[j] = threadIdx.x * 131 + j;
  const float2 neg_magic = make_float2(-12582912.0f, -12582912.0f);
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    float as = __ldg(ws_g + (it & 1023));
    const float2 as2 = make_float2(as, as);
#pragma unroll
    for (int c = 0; c < 8; ++c) {
#pragma unroll
      for (int j = 0; j < 16; ++j) asm volatile("" : "+r"(v[j]));
#pragma unroll
      for (int j4 = 0; j4 < 4; ++j4) {
        const float4 ws = __ldg(reinterpret_cast<const float4*>(ws_g + ((it * 128 + c * 16 + j4 * 4) & 4095)));
        float2 f0, f1;
        if (path == 0) {
          f0 = make_float2(__uint_as_float(v[j4 * 4 + 0] + 0x4B400000u), __uint_as_float(v[j4 * 4 + 1] + 0x4B400000u));
          f1 = make_float2(__uint_as_float(v[j4 * 4 + 2] + 0x4B400000u), __uint_as_float(v[j4 * 4 + 3] + 0x4B400000u));
          f0 = __fadd2_rn(f0, neg_magic);
          f1 = __fadd2_rn(f1, neg_magic);
        } else {
          f0 = make_float2(__uint_as_float(v[j4 * 4 + 0]), __uint_as_float(v[j4 * 4 + 1]));
          f1 = make_float2(__uint_as_float(v[j4 * 4 + 2]), __uint_as_float(v[j4 * 4 + 3]));
        }
        const float2 s0 = __fmul2_rn(make_float2(ws.x, ws.y), as2);
        const float2 s1 = __fmul2_rn(make_float2(ws.z, ws.w), as2);
        acc[c * 8 + j4 * 2 + 0] = __ffma2_rn(f0, s0, acc[c * 8 + j4 * 2 + 0]);
        acc[c * 8 + j4 * 2 + 1] = __ffma2_rn(f1, s1, acc[c * 8 + j4 * 2 + 1]);
      }
    }
  }
  const long long t1 = clock64();
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < 64; ++j) s += acc[j].x + acc[j].y;
  if (s == 1.2345f) sink[0] = s;
}

// alu2: candidate epilogue formulations with the scale vector in shared memory (broadcast LDS.128), 8 warps.
//  form 0: packed   u2 = ffma2(mf2, ws2, cws2); acc2 = ffma2(u2, as2, acc2)           (ws + cws tables)
//  form 1: scalar   t = fmaf(mf, as, cas);      acc = fmaf(t, ws, acc)                  (ws table only; as/cas reused)
//  form 2: packed   t2 = ffma2(mf2, as2, cas2); acc2 = ffma2(t2, ws2, acc2)
//  form 3: scalar   f = mf - magic; t = f * as; acc = fmaf(t, ws, acc)
template <int form>
__global__ void __launch_bounds__(256, 1) k_alu2(int iters, long long* cycles, float* sink, const float* ws_g) {
  __shared__ __align__(16) float sws[512];
  __shared__ uint32_t tslot;
  for (int i = threadIdx.x; i < 512; i += blockDim.x) sws[i] = ws_g[i];
  if (threadIdx.x < 32) tmem_alloc<512>(&tslot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = tslot;
  const int warp = threadIdx.x >> 5;
  float acc[128];
#pragma unroll
  for (int j = 0; j < 128; ++j) acc[j] = 0.f;
  uint32_t vb[2][32];
  const int col0 = (threadIdx.x >> 7) * 128;
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    const uint32_t taddr = tm + ((uint32_t)((warp & 3) * 32) << 16) + (it & 1) * 256 + col0;
    tmem_ld_32x32(taddr, vb[0]);
    const float as = sws[(it + threadIdx.x) & 255];
    const float cas = -12582912.0f * as;
    const float2 as2 = make_float2(as, as), cas2 = make_float2(cas, cas);
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      tmem_ld_wait();
      if (c < 3) tmem_ld_32x32(taddr + (c + 1) * 32, vb[(c + 1) & 1]);
      const uint32_t* v = vb[c & 1];
      const float4* ws4 = reinterpret_cast<const float4*>(sws + ((col0 + c * 32) & 255));
      const float4* cws4 = reinterpret_cast<const float4*>(sws + 256 + ((col0 + c * 32) & 255));
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        const float4 ws = ws4[q];
        float mf[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) mf[e] = __uint_as_float(v[q * 4 + e] + 0x4B400000u);
        float* a = &acc[c * 32 + q * 4];
        if (form == 0) {
          const float4 cw = cws4[q];
          const float2 u0 = __ffma2_rn(make_float2(mf[0], mf[1]), make_float2(ws.x, ws.y), make_float2(cw.x, cw.y));
          const float2 u1 = __ffma2_rn(make_float2(mf[2], mf[3]), make_float2(ws.z, ws.w), make_float2(cw.z, cw.w));
          const float2 r0 = __ffma2_rn(u0, as2, make_float2(a[0], a[1]));
          const float2 r1 = __ffma2_rn(u1, as2, make_float2(a[2], a[3]));
          a[0] = r0.x; a[1] = r0.y; a[2] = r1.x; a[3] = r1.y;
        } else if (form == 1) {
          a[0] = fmaf(fmaf(mf[0], as, cas), ws.x, a[0]);
          a[1] = fmaf(fmaf(mf[1], as, cas), ws.y, a[1]);
          a[2] = fmaf(fmaf(mf[2], as, cas), ws.z, a[2]);
          a[3] = fmaf(fmaf(mf[3], as, cas), ws.w, a[3]);
        } else if (form == 2) {
          const float2 t0 = __ffma2_rn(make_float2(mf[0], mf[1]), as2, cas2);
          const float2 t1 = __ffma2_rn(make_float2(mf[2], mf[3]), as2, cas2);
          const float2 r0 = __ffma2_rn(t0, make_float2(ws.x, ws.y), make_float2(a[0], a[1]));
          const float2 r1 = __ffma2_rn(t1, make_float2(ws.z, ws.w), make_float2(a[2], a[3]));
          a[0] = r0.x; a[1] = r0.y; a[2] = r1.x; a[3] = r1.y;
        } else if (form == 4) {  // scalar t (lite/heavy pipes) + packed accumulate (heavy pipe)
          const float t0 = fmaf(mf[0], as, cas), t1 = fmaf(mf[1], as, cas), t2 = fmaf(mf[2], as, cas),
                      t3 = fmaf(mf[3], as, cas);
          const float2 r0 = __ffma2_rn(make_float2(t0, t1), make_float2(ws.x, ws.y), make_float2(a[0], a[1]));
          const float2 r1 = __ffma2_rn(make_float2(t2, t3), make_float2(ws.z, ws.w), make_float2(a[2], a[3]));
          a[0] = r0.x; a[1] = r0.y; a[2] = r1.x; a[3] = r1.y;
        } else if (form == 5) {  // fp32 accumulators (kind::f8f6f4): packed mul + packed fma, no integer op
          const float2 t0 = __fmul2_rn(make_float2(__uint_as_float(v[q * 4 + 0]), __uint_as_float(v[q * 4 + 1])), as2);
          const float2 t1 = __fmul2_rn(make_float2(__uint_as_float(v[q * 4 + 2]), __uint_as_float(v[q * 4 + 3])), as2);
          const float2 r0 = __ffma2_rn(t0, make_float2(ws.x, ws.y), make_float2(a[0], a[1]));
          const float2 r1 = __ffma2_rn(t1, make_float2(ws.z, ws.w), make_float2(a[2], a[3]));
          a[0] = r0.x; a[1] = r0.y; a[2] = r1.x; a[3] = r1.y;
        } else if (form == 6) {  // fp32 accumulators: scalar mul (both fma pipes) + packed fma
          const float t0 = __uint_as_float(v[q * 4 + 0]) * as, t1 = __uint_as_float(v[q * 4 + 1]) * as,
                      t2 = __uint_as_float(v[q * 4 + 2]) * as, t3 = __uint_as_float(v[q * 4 + 3]) * as;
          const float2 r0 = __ffma2_rn(make_float2(t0, t1), make_float2(ws.x, ws.y), make_float2(a[0], a[1]));
          const float2 r1 = __ffma2_rn(make_float2(t2, t3), make_float2(ws.z, ws.w), make_float2(a[2], a[3]));
          a[0] = r0.x; a[1] = r0.y; a[2] = r1.x; a[3] = r1.y;
        } else if (form == 7) {  // fp32 accumulators: all scalar
          a[0] = fmaf(__uint_as_float(v[q * 4 + 0]) * as, ws.x, a[0]);
          a[1] = fmaf(__uint_as_float(v[q * 4 + 1]) * as, ws.y, a[1]);
          a[2] = fmaf(__uint_as_float(v[q * 4 + 2]) * as, ws.z, a[2]);
          a[3] = fmaf(__uint_as_float(v[q * 4 + 3]) * as, ws.w, a[3]);
        } else {
          a[0] = fmaf((mf[0] - 12582912.0f) * as, ws.x, a[0]);
          a[1] = fmaf((mf[1] - 12582912.0f) * as, ws.y, a[1]);
          a[2] = fmaf((mf[2] - 12582912.0f) * as, ws.z, a[2]);
          a[3] = fmaf((mf[3] - 12582912.0f) * as, ws.w, a[3]);
        }
      }
    }
  }
  const long long t1 = clock64();
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
  float sum = 0.f;
#pragma unroll
  for (int j = 0; j < 128; ++j) sum += acc[j];
  if (sum == 1.2345f) sink[0] = sum;
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) {
    tc_fence_after();
    tmem_dealloc<512>(tm);
  }
}


// ------------------------------------------------------------------------------------------------ pipe
// Issue / pipe throughput of the epilogue's instruction kinds, 8 warps (2 per SMSP), 16 independent chains per thread.
//  mix 0: FFMA2 only   1: FFMA only   2: FFMA2 + 2 FFMA (same lane-ops on packed and scalar)   3: FFMA2 + VIADD
//  mix 4: FFMA + VIADD   5: FFMA2 + FFMA + VIADD   6: FFMA2 + FFMA
template <int mix>
__global__ void __launch_bounds__(256, 1) k_pipe(int iters, long long* cycles, float* sink, float a, float b) {
  float2 acc2[16];
  float acc[32];
  uint32_t iv[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) { acc2[j] = make_float2(threadIdx.x + j, j); iv[j] = threadIdx.x * 3 + j; }
#pragma unroll
  for (int j = 0; j < 32; ++j) acc[j] = threadIdx.x - j;
  const float2 a2 = make_float2(a, a), b2 = make_float2(b, b);
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      if (mix == 0 || mix == 2 || mix == 3 || mix == 5 || mix == 6) acc2[j] = __ffma2_rn(acc2[j], a2, b2);
      if (mix == 1 || mix == 4 || mix == 5 || mix == 6) acc[j] = fmaf(acc[j], a, b);
      if (mix == 2) { acc[2 * j] = fmaf(acc[2 * j], a, b); acc[2 * j + 1] = fmaf(acc[2 * j + 1], a, b); }
      if (mix == 3 || mix == 4 || mix == 5) iv[j] = iv[j] + 0x4B400000u;
    }
  }
  const long long t1 = clock64();
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
  float sum = 0.f;
#pragma unroll
  for (int j = 0; j < 16; ++j) sum += acc2[j].x + acc2[j].y + __uint_as_float(iv[j]);
#pragma unroll
  for (int j = 0; j < 32; ++j) sum += acc[j];
  if (sum == 1.2345f) sink[0] = sum;
}


// pipe3: FMAs whose three operands are all distinct registers (like acc = fma(t, ws, acc) in the epilogue)
//  mix 0: 16 FFMA2 (3 distinct 64-bit operands)   1: 32 FFMA (3 distinct 32-bit operands)   2: 16 FFMA2 with scalar a, b
//  mix 3: 8 FFMA2 3-op + 8 FFMA2 scalar-broadcast form (the epilogue's pair per 2 elements), + 16 VIADD
template <int mix>
__global__ void __launch_bounds__(256, 1) k_pipe3(int iters, long long* cycles, float* sink, float a, float b) {
  float2 acc2[16], x2[16], y2[16];
  uint32_t iv[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    acc2[j] = make_float2(threadIdx.x + j, j);
    x2[j] = make_float2(1.0f + 1e-6f * (threadIdx.x + j), 1.0f - 1e-6f * j);
    y2[j] = make_float2(a * j, b * (threadIdx.x & 3));
    iv[j] = threadIdx.x + j;
  }
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      if (mix == 0) acc2[j] = __ffma2_rn(x2[j], y2[(j + 5) & 15], acc2[j]);
      if (mix == 1) {
        acc2[j].x = fmaf(x2[j].x, y2[(j + 5) & 15].x, acc2[j].x);
        acc2[j].y = fmaf(x2[j].y, y2[(j + 5) & 15].y, acc2[j].y);
      }
      if (mix == 2) acc2[j] = __ffma2_rn(acc2[j], make_float2(a, a), make_float2(b, b));
      if (mix == 3) {
        if (j & 1) {
          acc2[j] = __ffma2_rn(x2[j], y2[(j + 5) & 15], acc2[j]);
        } else {
          x2[j + 1] = __ffma2_rn(make_float2(__uint_as_float(iv[j]), __uint_as_float(iv[j + 1])), make_float2(a, a), make_float2(b, b));
        }
        iv[j] = iv[j] + 0x4B400000u;
      }
    }
  }
  const long long t1 = clock64();
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
  float sum = 0.f;
#pragma unroll
  for (int j = 0; j < 16; ++j) sum += acc2[j].x + acc2[j].y + x2[j].x + y2[j].y + __uint_as_float(iv[j]);
  if (sum == 1.2345f) sink[0] = sum;
}

// ------------------------------------------------------------------------------------------------ exact
// One CTA: A (128 rows) and B (256 rows) tiles of nblk 128-byte K blocks with int8 codes in [-7, 7] (any layout:
// both kinds read the same bytes through the same descriptors).  i8 result -> out_i, f8 (bytes translated to E4M3)
// -> out_f, accumulating nblk*4 MMAs into one accumulator each.
__device__ __forceinline__ uint8_t code_to_e4m3(int c) {
  const uint8_t lut[8] = {0x00, 0x38, 0x40, 0x44, 0x48, 0x4A, 0x4C, 0x4E};
  return c < 0 ? (uint8_t)(0x80 | lut[-c]) : lut[c];
}
__global__ void __launch_bounds__(128, 1) k_exact(int nblk, int reps, int worst, int32_t* out_i, float* out_f) {
  extern __shared__ unsigned char smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  unsigned char* smem = smem_raw + (((raw + 1023u) & ~1023u) - raw);
  __shared__ uint64_t bar;
  __shared__ uint32_t tslot;
  const int blk_bytes = 48 * 1024;  // per K block: A 16 KB then B 32 KB
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    fence_mbar_init();
  }
  if (threadIdx.x < 32) tmem_alloc<512>(&tslot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = tslot;
  for (int pass = 0; pass < 2; ++pass) {
    uint32_t rng = 12345u;
    for (int i = threadIdx.x; i < nblk * blk_bytes; i += blockDim.x) {
      rng = (uint32_t)i * 2654435761u + 0x9E3779B9u;
      rng ^= rng >> 15;
      rng *= 2246822519u;
      rng ^= rng >> 13;
      int c = worst ? 7 : (int)(rng % 15u) - 7;
      smem[i] = pass == 0 ? (uint8_t)(int8_t)c : code_to_e4m3(c);
    }
    fence_proxy_async_smem();
    __syncthreads();
    if (threadIdx.x == 0) {
      const uint32_t idesc = pass == 0 ? umma_idesc(2u, 1u, 128, 256) : umma_idesc(1u, 0u, 128, 256);
      for (int bb = 0; bb < nblk * reps; ++bb) {
        const int b = bb % nblk;
        const uint32_t sa = smem_u32(smem + b * blk_bytes);
        const uint64_t da = umma_desc_k128(sa), db = umma_desc_k128(sa + 16 * 1024);
        for (int s = 0; s < 4; ++s) {
          if (pass == 0)
            mma_i8_ss(tm, da + 2 * s, db + 2 * s, idesc, (bb | s) ? 1 : 0);
          else
            mma_f8_ss(tm + 256, da + 2 * s, db + 2 * s, idesc, (bb | s) ? 1 : 0);
        }
      }
      tc_commit(&bar);
      mbar_wait(&bar, pass, 0x901);
    }
    __syncthreads();
    tc_fence_after();
  }
  {  // every warp reads its lane quarter, both accumulators
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int c = 0; c < 256; c += 16) {
      uint32_t vi[16], vf[16];
      tmem_ld_32x16(tm + ((uint32_t)(warp * 32) << 16) + c, vi);
      tmem_ld_32x16(tm + ((uint32_t)(warp * 32) << 16) + 256 + c, vf);
      tmem_ld_wait();
      for (int j = 0; j < 16; ++j) {
        out_i[(warp * 32 + lane) * 256 + c + j] = (int32_t)vi[j];
        out_f[(warp * 32 + lane) * 256 + c + j] = __uint_as_float(vf[j]);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) {
    tc_fence_after();
    tmem_dealloc<512>(tm);
  }
}

static double med(std::vector<long long>& v) {
  double s = 0;
  for (auto x : v) s += (double)x;
  return s / v.size();
}

int main() {
  int sms = 0;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  long long* d_cyc;
  uint32_t* d_sink;
  float* d_ws;
  CK(cudaMalloc(&d_cyc, sms * sizeof(long long)));
  CK(cudaMalloc(&d_sink, 64));
  CK(cudaMalloc(&d_ws, 8192 * 4));
  CK(cudaMemset(d_ws, 0, 8192 * 4));
  std::vector<long long> h(sms);
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));

  // ---- mma
  CK(cudaFuncSetAttribute(k_mma, cudaFuncAttributeMaxDynamicSharedMemorySize, 50 * 1024));
  const char* kn[3] = {"i8", "f8f6f4", "f16"};
  for (int fill = 0; fill < 2; ++fill)
    for (int kind = 0; kind < 3; ++kind) {
      const int iters = 20000;
      k_mma<<<sms, 128, 50 * 1024>>>(kind, 200, d_cyc, fill);
      CK(cudaEventRecord(e0));
      k_mma<<<sms, 128, 50 * 1024>>>(kind, iters, d_cyc, fill);
      CK(cudaEventRecord(e1));
      CK(cudaDeviceSynchronize());
      float ms;
      CK(cudaEventElapsedTime(&ms, e0, e1));
      CK(cudaMemcpy(h.data(), d_cyc, sms * sizeof(long long), cudaMemcpyDeviceToHost));
      const double cyc = med(h) / (iters * 4.0);
      const double macs = (double)sms * iters * 4.0 * 128 * 256 * (kind == 2 ? 16 : 32);
      printf("UB mma kind=%s fill=%d clk/mma=%.1f ms=%.3f eff_clock_GHz=%.3f chip_Tops=%.1f\n", kn[kind], fill, cyc, ms,
             med(h) / (ms * 1e6), 2.0 * macs / (ms * 1e-3) / 1e12);
    }


  // ---- pipe
  {
    const int iters = 2000;
    auto rep = [&](int mix, const char* what) {
      CK(cudaDeviceSynchronize());
      CK(cudaMemcpy(h.data(), d_cyc, sms * sizeof(long long), cudaMemcpyDeviceToHost));
      printf("UB pipe mix=%d (%s) clk per 16-chain step (2 warps/SMSP)=%.2f\n", mix, what, med(h) / iters);
    };
    float* fs = reinterpret_cast<float*>(d_sink);
    k_pipe<0><<<sms, 256>>>(iters, d_cyc, fs, 1.0001f, 0.5f); rep(0, "16 FFMA2");
    k_pipe<1><<<sms, 256>>>(iters, d_cyc, fs, 1.0001f, 0.5f); rep(1, "16 FFMA");
    k_pipe<2><<<sms, 256>>>(iters, d_cyc, fs, 1.0001f, 0.5f); rep(2, "16 FFMA2 + 32 FFMA");
    k_pipe<3><<<sms, 256>>>(iters, d_cyc, fs, 1.0001f, 0.5f); rep(3, "16 FFMA2 + 16 VIADD");
    k_pipe<4><<<sms, 256>>>(iters, d_cyc, fs, 1.0001f, 0.5f); rep(4, "16 FFMA + 16 VIADD");
    k_pipe<5><<<sms, 256>>>(iters, d_cyc, fs, 1.0001f, 0.5f); rep(5, "16 FFMA2 + 16 FFMA + 16 VIADD");
    k_pipe<6><<<sms, 256>>>(iters, d_cyc, fs, 1.0001f, 0.5f); rep(6, "16 FFMA2 + 16 FFMA");
  }


  {
    const int iters = 2000;
    auto rep = [&](int mix, const char* what) {
      CK(cudaDeviceSynchronize());
      CK(cudaMemcpy(h.data(), d_cyc, sms * sizeof(long long), cudaMemcpyDeviceToHost));
      printf("UB pipe3 mix=%d (%s) clk per step (2 warps/SMSP)=%.2f\n", mix, what, med(h) / iters);
    };
    float* fs = reinterpret_cast<float*>(d_sink);
    k_pipe3<0><<<sms, 256>>>(iters, d_cyc, fs, 1.0001f, 0.5f); rep(0, "16 FFMA2 3 distinct operands");
    k_pipe3<1><<<sms, 256>>>(iters, d_cyc, fs, 1.0001f, 0.5f); rep(1, "32 FFMA 3 distinct operands");
    k_pipe3<2><<<sms, 256>>>(iters, d_cyc, fs, 1.0001f, 0.5f); rep(2, "16 FFMA2 scalar a,b");
    k_pipe3<3><<<sms, 256>>>(iters, d_cyc, fs, 1.0001f, 0.5f); rep(3, "8 FFMA2 3-op + 8 FFMA2 scalar + 16 VIADD");
  }

  // ---- ldtm
  {
    const int iters = 4000;
    auto rep = [&](int nw, int var) {
      CK(cudaDeviceSynchronize());
      CK(cudaMemcpy(h.data(), d_cyc, sms * sizeof(long long), cudaMemcpyDeviceToHost));
      printf("UB ldtm warps=%d variant=%d clk per 128x256 drain=%.1f  (B/clk/SM=%.1f)\n", nw, var, med(h) / iters,
             131072.0 / (med(h) / iters));
    };
    k_ldtm<4, 0><<<sms, 384>>>(iters, d_cyc, d_sink); rep(4, 0);
    k_ldtm<4, 1><<<sms, 384>>>(iters, d_cyc, d_sink); rep(4, 1);
    k_ldtm<8, 0><<<sms, 384>>>(iters, d_cyc, d_sink); rep(8, 0);
    k_ldtm<8, 1><<<sms, 384>>>(iters, d_cyc, d_sink); rep(8, 1);
  }

  // ---- alu
  {
    const int iters = 4000;
    auto rep = [&](int path) {
      CK(cudaDeviceSynchronize());
      CK(cudaMemcpy(h.data(), d_cyc, sms * sizeof(long long), cudaMemcpyDeviceToHost));
      printf("UB alu path=%d clk per 128x256 group rescale=%.1f\n", path, med(h) / iters);
    };
    k_alu<0><<<sms, 256>>>(iters, d_cyc, reinterpret_cast<float*>(d_sink), d_ws); rep(0);
    k_alu<1><<<sms, 256>>>(iters, d_cyc, reinterpret_cast<float*>(d_sink), d_ws); rep(1);
  }

  // ---- alu2
  {
    const int iters = 4000;
    auto rep = [&](int form) {
      CK(cudaDeviceSynchronize());
      CK(cudaMemcpy(h.data(), d_cyc, sms * sizeof(long long), cudaMemcpyDeviceToHost));
      printf("UB alu2 form=%d clk per 128x256 group rescale=%.1f\n", form, med(h) / iters);
    };
    k_alu2<0><<<sms, 256>>>(iters, d_cyc, reinterpret_cast<float*>(d_sink), d_ws); rep(0);
    k_alu2<1><<<sms, 256>>>(iters, d_cyc, reinterpret_cast<float*>(d_sink), d_ws); rep(1);
    k_alu2<2><<<sms, 256>>>(iters, d_cyc, reinterpret_cast<float*>(d_sink), d_ws); rep(2);
    k_alu2<3><<<sms, 256>>>(iters, d_cyc, reinterpret_cast<float*>(d_sink), d_ws); rep(3);
    k_alu2<4><<<sms, 256>>>(iters, d_cyc, reinterpret_cast<float*>(d_sink), d_ws); rep(4);
    k_alu2<5><<<sms, 256>>>(iters, d_cyc, reinterpret_cast<float*>(d_sink), d_ws); rep(5);
    k_alu2<6><<<sms, 256>>>(iters, d_cyc, reinterpret_cast<float*>(d_sink), d_ws); rep(6);
    k_alu2<7><<<sms, 256>>>(iters, d_cyc, reinterpret_cast<float*>(d_sink), d_ws); rep(7);
  }

  // ---- exact
  CK(cudaFuncSetAttribute(k_exact, cudaFuncAttributeMaxDynamicSharedMemorySize, 4 * 48 * 1024 + 2048));
  int32_t* d_i;
  float* d_f;
  CK(cudaMalloc(&d_i, 128 * 256 * 4));
  CK(cudaMalloc(&d_f, 128 * 256 * 4));
  std::vector<int32_t> hi(128 * 256);
  std::vector<float> hf(128 * 256);
  for (int worst = 0; worst < 2; ++worst)
    for (int cfg = 0; cfg < 6; ++cfg) {
      const int nblk = cfg < 4 ? cfg + 1 : 4, reps = cfg < 4 ? 1 : (cfg == 4 ? 2 : 4);
      k_exact<<<1, 128, 4 * 48 * 1024 + 2048>>>(nblk, reps, worst, d_i, d_f);
      CK(cudaDeviceSynchronize());
      CK(cudaMemcpy(hi.data(), d_i, hi.size() * 4, cudaMemcpyDeviceToHost));
      CK(cudaMemcpy(hf.data(), d_f, hf.size() * 4, cudaMemcpyDeviceToHost));
      long bad = 0;
      int imax = 0;
      double maxdiff = 0;
      for (size_t k = 0; k < hi.size(); ++k) {
        if ((float)hi[k] != hf[k]) ++bad;
        if (abs(hi[k]) > imax) imax = abs(hi[k]);
        double d = fabs((double)hi[k] - (double)hf[k]);
        if (d > maxdiff) maxdiff = d;
      }
      printf("UB exact worst=%d kcols=%d mismatches=%ld max|sum|=%d maxdiff=%.1f sample i=%d f=%.1f\n", worst,
             nblk * 128 * reps, bad, imax, maxdiff, hi[1000], hf[1000]);
    }
  return 0;
}
