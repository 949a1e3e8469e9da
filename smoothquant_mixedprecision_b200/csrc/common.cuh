// Shared host/device helpers for libmq.so (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/mq.h"

// ---------------------------------------------------------------- host side: error convention of the C ABI
void mq_set_error(const char* fmt, ...);
int mq_get_option(const char* name, int dflt);
int mq_sm_count();

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
// code = round_half_even(rnd_T(v / scale))                        (fake_quant.py:97, :142, :193)
template <typename T>
__device__ __forceinline__ float mq_code(float v, float scale) {
  return rintf(Num<T>::rnd(__fdiv_rn(v, scale)));
}
