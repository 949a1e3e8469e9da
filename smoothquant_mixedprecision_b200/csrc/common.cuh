// Shared host/device helpers for libmq.so (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/mq.h"

// ---------------------------------------------------------------- host side: error convention of the C ABI
void mq_set_error(const char* fmt, ...);
int mq_get_option(const char* name, int dflt);
int mq_sm_count();
int mq_worker_sms();   // mq_sm_count() minus option sm_reserve

#define MQ_CHECK_ARG(cond, ...)            \
  do {                                     \
    if (!(cond)) {                         \
      mq_set_error(__VA_ARGS__);           \
      return MQ_ERR_ARG;                   \
    }                                      \
  } while (0)

#define MQ_CHECK_CUDA(expr)                                                                 \
  do {                                                                                      \
    cudaError_t e__ = (expr);                                                               \
    if (e__ != cudaSuccess) {                                                               \
      mq_set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, __LINE__); \
      return MQ_ERR_CUDA;                                                                   \
    }                                                                                       \
  } while (0)

// ---------------------------------------------------------------- device side: dtype-local rounding
// The reference performs every elementwise op in the tensor's own dtype: the fp32 result of each op is rounded to
// that dtype before the next op.  Num<T>::rnd() is that rounding.
template <typename T>
struct Num;

template <>
struct Num<__half> {
  static constexpr int kDtype = MQ_F16;
  __device__ __forceinline__ static float to_f(__half v) { return __half2float(v); }
  __device__ __forceinline__ static __half from_f(float v) { return __float2half_rn(v); }
  __device__ __forceinline__ static float rnd(float v) { return __half2float(__float2half_rn(v)); }
};
template <>
struct Num<__nv_bfloat16> {
  static constexpr int kDtype = MQ_BF16;
  __device__ __forceinline__ static float to_f(__nv_bfloat16 v) { return __bfloat162float(v); }
  __device__ __forceinline__ static __nv_bfloat16 from_f(float v) { return __float2bfloat16_rn(v); }
  __device__ __forceinline__ static float rnd(float v) { return __bfloat162float(__float2bfloat16_rn(v)); }
};
template <>
struct Num<float> {
  static constexpr int kDtype = MQ_F32;
  __device__ __forceinline__ static float to_f(float v) { return v; }
  __device__ __forceinline__ static float from_f(float v) { return v; }
  __device__ __forceinline__ static float rnd(float v) { return v; }
};

// scale = clamp(absmax, 1e-5).div(qmax), both ops rounded to T   (fake_quant.py:92-94, :137-139, :188-190)
template <typename T>
__device__ __forceinline__ float mq_group_scale(float absmax, float qmax) {
  float lo = Num<T>::rnd(1e-5f);
  float c = fmaxf(absmax, lo);
  return Num<T>::rnd(__fdiv_rn(c, qmax));
}
// Half-width of the band around k + 0.5 inside which rounding the quotient to T can change the code: half a T-ulp
// of the largest quotient (|q| < 2^(n_bits-1), T has `mant` explicit mantissa bits) plus slack for the reciprocal
// approximation.  Outside the band the code is rint(v * (1 / scale)).  Host side; passed to the kernels as
// thr = 0.5 - margin.
template <typename T> struct MantBits;
template <> struct MantBits<__half> { static constexpr int v = 10; };
template <> struct MantBits<__nv_bfloat16> { static constexpr int v = 7; };
template <> struct MantBits<float> { static constexpr int v = 23; };
template <typename T>
inline float mq_quot_threshold(int n_bits) {
  const float half_ulp = ldexpf(1.0f, (n_bits - 2) - MantBits<T>::v - 1);  // half ulp of values in [2^(b-2), 2^(b-1))
  const float margin = 1.25f * half_ulp + 4e-6f * (float)(1 << (n_bits - 1));
  return 0.5f - (margin < 0.25f ? margin : 0.25f);
}

// The same code as mq_code() without the IEEE division on the common path: t = rn(v * inv + 1.5 * 2^23) holds
// rint(v * inv) in its mantissa; only quotients farther than `thr` from the nearest integer (about 0.5 % of fp16
// values at 4 bits) take the exact divide -> round-to-T -> rint path, so the result is bit-identical to mq_code().
// Returns the code as a float; `inv` = 1 / scale (any good approximation).  Valid for |v / scale| < 2^22.
template <typename T>
__device__ __forceinline__ float mq_code_fast(float v, float scale, float inv, float thr) {
  constexpr float kM = 12582912.0f;
  float r = fmaf(v, inv, kM) - kM;
  const float d = fabsf(fmaf(v, inv, -r));
  if (!(d <= thr)) r = rintf(Num<T>::rnd(__fdiv_rn(v, scale)));
  return r;
}
// small integer-valued float -> int without F2I (SFU): bits(r + 1.5 * 2^23) - bits(1.5 * 2^23)
__device__ __forceinline__ int mq_small_int(float r) {
  return (int)(__float_as_uint(r + 12582912.0f) - 0x4B400000u);
}
// code = round_half_even(rnd_T(v / scale))                        (fake_quant.py:97, :142, :193)
template <typename T>
__device__ __forceinline__ float mq_code(float v, float scale) {
  return rintf(Num<T>::rnd(__fdiv_rn(v, scale)));
}
