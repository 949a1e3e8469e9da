// K1 / P0, 16-bit streaming path: the row quantiser of quant_rows.cu re-built around packed half2 / bfloat162
// arithmetic.  Included by quant_rows.cu inside its anonymous namespace (shares QuantArgs and the dispatch).
//
// Same contract as quant_rows_fast MODE 0 (codes int8 [R, ld_codes], scales fp32 [G, ld_scales] group-major, salient
// copy), bit for bit -- tests/test_ops_gpu.py requires the two kernels and the generic kernel to agree -- for fp16 /
// bf16 inputs, power-of-two groups of 32..1024 columns, no run-time smoothing divisor, no hybrid tail.
//
// What changed against quant_rows_fast (r01: 0.35 of the HBM roofline, 24 instructions per element, issue bound):
//   * two elements per arithmetic instruction.  The code of an element is rint(rn_T(x / s)) (fake_quant.py:97, :142).
//     rn_T(x / s) is the correctly rounded quotient IN T (the reference's fp32 intermediate is innocuous: a quotient of
//     two p-bit numbers is never within 2^-(2p+1) of a rounding boundary of T), and T itself can produce it without a
//     divide: q0 = rn(x * y), r = fma(-s, q0, x), q1 = fma(r, y, q0) with y = rn(1 / s) (Markstein), as HMUL2 / HFMA2 /
//     HFMA2 on pairs; the round-half-even to an integer is one more HADD2 of 1.5 * 2^10 (1.5 * 2^7 for bf16), whose low
//     byte is the two's-complement code.  For fp16 the group is first moved to a fixed binade -- s' = s * 2^j in
//     [8, 16), x' = x * 2^j, both exact -- so that the remainder never leaves the (denormal-bounded) fp16 grid and
//     1 / s' comes from a 1024-entry table of correctly rounded reciprocals built with integer division.  Verified
//     against the reference arithmetic for every (mantissa of s, x') pair (22 M cases fp16, all of bf16).
//     Groups whose scale is denormal / near the ends of the exponent range take the exact fp32 path of
//     quant_rows_fast per element (never on real activations).
//   * no per-element address arithmetic.  A thread owns the same 32 permuted positions of every row it touches, so
//     the 32 byte offsets of its gather live in registers for the whole kernel; a load is one LDS.U16 [R + UR].
//     With an identity permutation (perm == NULL: the producer already emits the consumer's order, see
//     fake_quant.fold_mlp_permutation) the gather disappears: four conflict-free LDS.128 per 32 elements.
//   * a dedicated producer warp keeps three stages of rows in flight with bulk TMA copies and full / empty mbarriers
//     (no block-wide barrier per stage) and copies the salient columns out while the 16 consumer warps quantise.
//   * blocked input (x_blocks > 1): a row may be spread over column blocks [b][r][blk_cols] -- the layout an
//     all-gather of column shards leaves behind (parallel.gather_columns_blocked) -- and is reassembled by the bulk
//     copies for free.
#pragma once

namespace h2 {

constexpr int kConsWarps = 16;
constexpr int kConsThreads = kConsWarps * 32;
constexpr int kThreadsH2 = kConsThreads + 32;     // + producer warp
constexpr int kStagesH2 = 3;
constexpr int kElemsPerLane = 32;                 // 4 pieces of 8 elements
constexpr int kRowPadH2 = 8;                      // zero elements behind every staged row

template <typename T> struct Pk;
template <> struct Pk<__half> {
  using V2 = __half2;
  static constexpr uint32_t kMagic2 = 0x66006600u;   // 1536.0 | 1536.0 : ulp 1 in [1024, 2048)
  static constexpr bool kScale = true;               // move the group to the binade [8, 16)
  static constexpr int kMant = 10;
  static constexpr int kMaxBits = 8;                 // |code| <= 127 < 512
  static constexpr uint32_t kBias = 0u;              // low byte of bits(1536 + code) is the code itself
  __device__ static __forceinline__ float to_f(uint32_t lo16) { return __half2float(__ushort_as_half((unsigned short)lo16)); }
  __device__ static __forceinline__ uint32_t from_f(float v) { return __half_as_ushort(__float2half_rn(v)); }
  __device__ static __forceinline__ float2 pair_f(uint32_t p) { return __half22float2(*reinterpret_cast<__half2*>(&p)); }
  __device__ static __forceinline__ uint32_t pack_rn(float a, float b) {
    const __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<const uint32_t*>(&h);
  }
};
template <> struct Pk<__nv_bfloat16> {
  using V2 = __nv_bfloat162;
  static constexpr uint32_t kMagic2 = 0x43404340u;   // 192.0 | 192.0 : ulp 1 in [128, 256)
  static constexpr bool kScale = false;
  static constexpr int kMant = 7;
  static constexpr int kMaxBits = 7;                 // |code| <= 63 < 64
  static constexpr uint32_t kBias = 0x40u;           // 7-bit mantissa: low byte of bits(192 + code) is 64 + code
  __device__ static __forceinline__ float to_f(uint32_t lo16) { return __uint_as_float(lo16 << 16); }
  __device__ static __forceinline__ uint32_t from_f(float v) { return __bfloat16_as_ushort(__float2bfloat16_rn(v)); }
  __device__ static __forceinline__ float2 pair_f(uint32_t p) {
    return make_float2(__uint_as_float(p << 16), __uint_as_float(p & 0xFFFF0000u));
  }
  __device__ static __forceinline__ uint32_t pack_rn(float a, float b) {
    const __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<const uint32_t*>(&h);
  }
};

template <typename V2> __device__ __forceinline__ V2 as_v2(uint32_t u) { return *reinterpret_cast<V2*>(&u); }
template <typename V2> __device__ __forceinline__ uint32_t as_u(V2 v) { return *reinterpret_cast<uint32_t*>(&v); }
template <typename V2> __device__ __forceinline__ uint32_t pmul(uint32_t a, uint32_t b) { return as_u(__hmul2(as_v2<V2>(a), as_v2<V2>(b))); }
template <typename V2> __device__ __forceinline__ uint32_t pfma(uint32_t a, uint32_t b, uint32_t c) { return as_u(__hfma2(as_v2<V2>(a), as_v2<V2>(b), as_v2<V2>(c))); }
template <typename V2> __device__ __forceinline__ uint32_t padd(uint32_t a, uint32_t b) { return as_u(__hadd2(as_v2<V2>(a), as_v2<V2>(b))); }
template <typename V2> __device__ __forceinline__ uint32_t pmax(uint32_t a, uint32_t b) { return as_u(__hmax2(as_v2<V2>(a), as_v2<V2>(b))); }
template <typename V2> __device__ __forceinline__ uint32_t pabsmax(uint32_t acc, uint32_t v) { return as_u(__hmax2(as_v2<V2>(acc), __habs2(as_v2<V2>(v)))); }

__device__ __forceinline__ uint32_t lds_u16(uint32_t addr) {
  uint32_t r;
  asm volatile("ld.shared.u16 %0, [%1];" : "=r"(r) : "r"(addr));
  return r;
}
__device__ __forceinline__ uint4 lds_u128(uint32_t addr) {
  uint4 r;
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(addr));
  return r;
}

struct H2Args {
  QuantArgs q;
  int nchunk;          // lanes (32-element chunks) per row = ld_codes / 32
  int lpg;             // lanes per group = g / 32
  int rpp;             // gather mode: rows per pass (threads own fixed positions); 0 = contiguous mode
  int rps;             // rows per stage
  int pitch;           // staged row pitch in elements (>= max(K, nchunk * 32) + kRowPadH2)
  int x_blocks;        // column blocks of the input (1 = plain row-major)
  int64_t blk_cols, blk_stride;
  float rq;            // rn(1 / qmax)
  // fused producer norm (mq_norm_quant_rows): 0 none, 1 RMSNorm, 2 LayerNorm; the staged rows are normalised in place
  // before they are quantised
  int norm;
  int norm_parts;      // consumer warps sharing one row in the norm phase (power of two)
  const void* norm_w;  // [K] T
  const void* norm_b;  // [K] T or NULL (LayerNorm)
  float eps;
  void* out_y;         // optional dump of the normalised rows [R, ld_y] T (tests; consumers outside the path)
  int64_t ld_y;
};

// reciprocal-mantissa table: entry m = bits to ADD to the exponent part of y = rn_T(1 / 1.m) (carry into the exponent
// for m == 0).  Built with integer division: q = round(2^(2*mant+1) / (2^mant + m)); a tie cannot occur for m > 0.
template <typename T>
__device__ __forceinline__ uint16_t rcp_mant_entry(int m) {
  constexpr int B = Pk<T>::kMant;
  if (m == 0) return (uint16_t)(1u << B);
  const uint32_t num = 1u << (2 * B + 1), d = (1u << B) + (uint32_t)m;
  uint32_t q = num / d;
  const uint32_t r = num - q * d;
  if (2 * r > d) ++q;
  return (uint16_t)(q - (1u << B));
}

template <typename T>
struct GroupK {
  uint32_t ns2, inv2, p2;   // -s' | -s',  y | y,  2^j | 2^j  (packed pairs)
  float scale;              // the reference's scale, widened (what mq_mixed_gemm reads)
  bool fast;
};

// scale = rn_T(rn_32(clamp(absmax, rn_T(1e-5)) / qmax))  -- clamp_(min=1e-5).div_(qmax) on a T tensor, fake_quant.py:92-94
template <typename T>
__device__ __forceinline__ GroupK<T> group_consts(uint32_t amax16, float qmax, float rq, const uint16_t* __restrict__ tbl) {
  GroupK<T> k;
  const float lo = Pk<T>::to_f(Pk<T>::from_f(1e-5f));
  const float c = fmaxf(Pk<T>::to_f(amax16 & 0xFFFFu), lo);
  const float q0 = c * rq;
  const float sf = fmaf(fmaf(-qmax, q0, c), rq, q0);          // correctly rounded c / qmax (qmax = 2^k - 1)
  const uint32_t sb = Pk<T>::from_f(sf);
  k.scale = Pk<T>::to_f(sb);
  constexpr int B = Pk<T>::kMant;
  const uint32_t E = sb >> B, m = sb & ((1u << B) - 1u);
  uint32_t sp, inv, p;
  if (Pk<T>::kScale) {                                          // fp16: s' = 1.m * 2^3, y = rn(1 / s'), p = 2^(3 - e)
    k.fast = E >= 3u && E <= 30u;
    sp = m | (18u << B);
    inv = (11u << B) + tbl[m];
    p = (33u - E) << B;
  } else {                                                      // bf16: wide exponent, no rescaling
    k.fast = E >= 16u && E <= 240u;
    sp = sb;
    inv = ((253u - E) << B) + tbl[m];
    p = 0;
  }
  k.ns2 = (sp | (sp << 16)) ^ 0x80008000u;
  k.inv2 = inv | (inv << 16);
  k.p2 = p | (p << 16);
  return k;
}

__device__ __forceinline__ void sts_u128(uint32_t addr, const uint4& v) {
  asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" :: "r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void cons_bar_sync() {      // the 16 consumer warps only (the producer warp never joins)
  asm volatile("bar.sync 1, %0;" :: "n"(kConsThreads) : "memory");
}
constexpr int kNormMaxRows = 64;                        // rows per stage the norm phase supports (partial-sum slots)

// Normalise the `nr` staged rows of one stage in place.  Work unit = (row, part): `parts` consumer warps share a row.
//   RMSNorm  (transformers LlamaRMSNorm.forward): y = rn_T(w * rn_T(x32 * rsqrt(mean(x32^2) + eps)))
//   LayerNorm (torch.nn.functional.layer_norm on a 16-bit tensor): y = rn_T(((x32 - mean) * rstd) * w32 + b32)
// Statistics are fp32 sums over all K columns of the row (salient ones included: the norm knows no partition).
template <typename T, int NORM>
__device__ __forceinline__ void norm_stage(const H2Args& a, uint32_t sbase, uint32_t pitch_b, int nr, int64_t r0, int K,
                                           uint32_t w_sm, uint32_t b_sm, float2* psum, int warp, int lane) {
  using V2 = typename Pk<T>::V2;
  const int P = a.norm_parts, part = warp & (P - 1), wrow = warp / P, rows_per_round = kConsWarps / P;
  const int nvec = K >> 3, vpp = (nvec + P - 1) / P;
  const int v0 = part * vpp, v1 = min(nvec, v0 + vpp);
  for (int rb = 0; rb < a.rps; rb += rows_per_round) {
    const int r = rb + wrow;
    const bool have = r < nr;
    const uint32_t row_a = sbase + (uint32_t)min(r, a.rps - 1) * pitch_b;
    float s1 = 0.f, s2 = 0.f;
    float2 q1 = make_float2(0.f, 0.f), q2 = make_float2(0.f, 0.f);
    if (have) {
      for (int v = v0 + lane; v < v1; v += 32) {
        const uint4 d = lds_u128(row_a + (uint32_t)v * 16u);
        const uint32_t wd[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float2 f = Pk<T>::pair_f(wd[j]);
          q2 = __ffma2_rn(f, f, q2);                     // two independent fp32 accumulators per lane
          if (NORM == 2) q1 = __fadd2_rn(q1, f);
        }
      }
    }
    s2 = q2.x + q2.y;
    s1 = q1.x + q1.y;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      s2 += __shfl_xor_sync(0xffffffffu, s2, o);
      if (NORM == 2) s1 += __shfl_xor_sync(0xffffffffu, s1, o);
    }
    if (lane == 0 && r < a.rps) psum[r * P + part] = make_float2(s1, s2);
    cons_bar_sync();
    if (have) {
      float t1 = 0.f, t2 = 0.f;
      for (int pp = 0; pp < P; ++pp) {
        const float2 v = psum[r * P + pp];
        t1 += v.x;
        t2 += v.y;
      }
      const float inv_k = 1.0f / (float)K;
      float mean = 0.f, var = t2 * inv_k;
      if (NORM == 2) {
        mean = t1 * inv_k;
        var = fmaxf(var - mean * mean, 0.f);
      }
      const float rstd = rsqrtf(var + a.eps);
      const float2 rstd2 = make_float2(rstd, rstd), nmean2 = make_float2(-mean, -mean);
      T* yrow = a.out_y ? static_cast<T*>(a.out_y) + (r0 + r) * a.ld_y : nullptr;
      for (int v = v0 + lane; v < v1; v += 32) {
        const uint4 d = lds_u128(row_a + (uint32_t)v * 16u);
        const uint4 wv = lds_u128(w_sm + (uint32_t)v * 16u);
        const uint32_t xd[4] = {d.x, d.y, d.z, d.w}, wd[4] = {wv.x, wv.y, wv.z, wv.w};
        uint32_t yd[4];
        if (NORM == 1) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float2 h = __fmul2_rn(Pk<T>::pair_f(xd[j]), rstd2);
            yd[j] = pmul<V2>(Pk<T>::pack_rn(h.x, h.y), wd[j]);    // rn_T(w * h): one rounding, as the fp32 product
          }
        } else {
          const uint4 bv = b_sm ? lds_u128(b_sm + (uint32_t)v * 16u) : make_uint4(0u, 0u, 0u, 0u);
          const uint32_t bd[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float2 n = __fmul2_rn(__fadd2_rn(Pk<T>::pair_f(xd[j]), nmean2), rstd2);
            const float2 y = __ffma2_rn(n, Pk<T>::pair_f(wd[j]), Pk<T>::pair_f(bd[j]));
            yd[j] = Pk<T>::pack_rn(y.x, y.y);
          }
        }
        const uint4 yv = make_uint4(yd[0], yd[1], yd[2], yd[3]);
        sts_u128(row_a + (uint32_t)v * 16u, yv);
        if (yrow) *reinterpret_cast<uint4*>(yrow + (int64_t)v * 8) = yv;
      }
    }
  }
}

template <typename T, bool GATHER, int NORM = 0>
__global__ void __launch_bounds__(kThreadsH2, 1)
quant_rows_h2(const H2Args a) {
  using V2 = typename Pk<T>::V2;
  extern __shared__ __align__(128) unsigned char smem[];
  const QuantArgs& q = a.q;
  griddep_launch();
  griddep_wait();              // programmatic dependent launch: the input rows come from the previous kernel of the stream
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int K = (int)q.K, Kn = (int)q.Kn;
  const uint32_t pitch_b = (uint32_t)a.pitch * 2u;
  const uint32_t zero_off = (uint32_t)(a.pitch - kRowPadH2) * 2u;       // byte offset of the zero pad inside a staged row

  // shared memory: [full[3] empty[3] | pad to 128][tbl u16 x 2^mant][ssal u16 x sal_cols][rows ...]
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem);
  uint64_t* empty_bar = full_bar + kStagesH2;
  uint64_t* norm_bar = empty_bar + kStagesH2;                            // NORM: rows of the stage are normalised
  uint16_t* tbl = reinterpret_cast<uint16_t*>(smem + 128);
  constexpr int kTbl = 1 << Pk<T>::kMant;
  uint16_t* ssal = tbl + kTbl;
  size_t rows_off = (128 + (size_t)kTbl * 2 + (size_t)q.sal_cols * 2 + 127) / 128 * 128;
  // NORM: [psum float2 x kNormMaxRows][norm weight T x K][norm bias T x K] in front of the rows
  float2* psum = reinterpret_cast<float2*>(smem + rows_off);
  const uint32_t w_sm = smem_u32(smem + rows_off + kNormMaxRows * 8);
  const uint32_t b_sm = (NORM == 2 && a.norm_b) ? w_sm + (uint32_t)K * 2u : 0u;
  if (NORM) rows_off += (kNormMaxRows * 8 + (size_t)K * 2 * (NORM == 2 ? 2 : 1) + 127) / 128 * 128;
  unsigned char* srows = smem + rows_off;
  const uint32_t srows_a = smem_u32(srows);
  const uint32_t stage_b = (uint32_t)a.rps * pitch_b;

  const int64_t nblocks = (q.R + a.rps - 1) / a.rps;
  const int64_t my_first = blockIdx.x;
  const int64_t my_count = my_first < nblocks ? (nblocks - my_first + gridDim.x - 1) / gridDim.x : 0;

  if (tid == 0) {
    for (int s = 0; s < kStagesH2; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], kConsWarps + 1);
      mbar_init(&norm_bar[s], kConsWarps);
    }
    fence_mbar_init();
  }
  for (int m = tid; m < kTbl; m += blockDim.x) tbl[m] = rcp_mant_entry<T>(m);
  if (NORM) {                                                     // norm weight (+ bias): K * 2 bytes each, K % 8 == 0
    for (int v = tid; v < (K >> 3); v += blockDim.x) {
      sts_u128(w_sm + (uint32_t)v * 16u, __ldg(reinterpret_cast<const uint4*>(a.norm_w) + v));
      if (b_sm) sts_u128(b_sm + (uint32_t)v * 16u, __ldg(reinterpret_cast<const uint4*>(a.norm_b) + v));
    }
  }
  for (int j = tid; j < q.sal_cols; j += blockDim.x)
    ssal[j] = j < q.Ks ? (uint16_t)(q.sal[j] * 2) : (uint16_t)zero_off;
  {  // zero everything behind the K copied elements of every staged row (never written by the bulk copies)
    const int tail = a.pitch - K;
    for (int j = tid; j < kStagesH2 * a.rps * tail; j += blockDim.x)
      reinterpret_cast<uint16_t*>(srows)[(size_t)(j / tail) * a.pitch + K + (j % tail)] = 0;
  }
  __syncthreads();

  if (warp == kConsWarps) {
    // ------------------------------------------------------------------ producer warp: loads + salient copy
    auto issue = [&](int64_t i) {
      const int s = (int)(i % kStagesH2);
      const int64_t r0 = (my_first + i * gridDim.x) * a.rps;
      const int nr = (int)min((int64_t)a.rps, q.R - r0);
      if (lane == 0) mbar_expect_tx(&full_bar[s], (uint32_t)nr * (uint32_t)K * 2u);
      __syncwarp();
      const T* x = static_cast<const T*>(q.x);
      unsigned char* dst = srows + (size_t)s * stage_b;
      for (int j = lane; j < nr * a.x_blocks; j += 32) {
        const int r = j / a.x_blocks, b = j - r * a.x_blocks;
        const int64_t c0 = (int64_t)b * a.blk_cols;
        const int64_t cols = min(a.blk_cols, (int64_t)K - c0);
        bulk_g2s(dst + (size_t)r * pitch_b + c0 * 2, x + (int64_t)b * a.blk_stride + (r0 + r) * q.ld + (a.x_blocks > 1 ? 0 : c0),
                 (uint32_t)cols * 2u, &full_bar[s]);
      }
    };
    for (int64_t i = 0; i < min((int64_t)kStagesH2 - 1, my_count); ++i) issue(i);
    for (int64_t i = 0; i < my_count; ++i) {
      const int s = (int)(i % kStagesH2);
      const int64_t nx = i + kStagesH2 - 1;                 // refill the stage block i-1 used
      if (nx < my_count) {
        if (nx >= kStagesH2 && !mbar_wait(&empty_bar[nx % kStagesH2], (uint32_t)((nx / kStagesH2 - 1) & 1), 0x120u)) return;
        issue(nx);
      }
      if (!mbar_wait(&full_bar[s], (uint32_t)((i / kStagesH2) & 1), 0x110u + s)) return;
      if (NORM && !mbar_wait(&norm_bar[s], (uint32_t)((i / kStagesH2) & 1), 0x130u + s)) return;
      if (q.sal_cols > 0) {
        const int64_t r0 = (my_first + i * gridDim.x) * a.rps;
        const int nr = (int)min((int64_t)a.rps, q.R - r0);
        const int nv = q.sal_cols / 8;
        const uint32_t base = srows_a + (uint32_t)s * stage_b;
        for (int it = lane; it < nr * nv; it += 32) {
          const int r = it / nv, c = (it - r * nv) * 8;
          const uint32_t rb = base + (uint32_t)r * pitch_b;
          const uint4 o4 = *reinterpret_cast<const uint4*>(&ssal[c]);
          const uint32_t o[4] = {o4.x, o4.y, o4.z, o4.w};
          uint32_t v[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) v[e] = __byte_perm(lds_u16(rb + (o[e] & 0xFFFFu)), lds_u16(rb + (o[e] >> 16)), 0x5410);
          T* dst = static_cast<T*>(q.out_sal) + (r0 + r) * q.ld_sal + c;
          if (c + 8 <= q.sal_width) {
            *reinterpret_cast<uint4*>(dst) = make_uint4(v[0], v[1], v[2], v[3]);
          } else {
            for (int e = 0; e < 8 && c + e < q.sal_width; ++e)
              reinterpret_cast<uint16_t*>(dst)[e] = (uint16_t)((v[e >> 1] >> ((e & 1) * 16)) & 0xFFFFu);
          }
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty_bar[s]);
    }
    return;
  }

  // -------------------------------------------------------------------- consumer warps
  const int lpg = a.lpg, lpg_sh = 31 - __clz(lpg), nchunk = a.nchunk;   // lpg is a power of two
  const int nact = GATHER ? a.rpp * nchunk : kConsThreads;      // threads that own work
  const bool owner = tid < nact;
  const int n_iter = (a.rps * nchunk + nact - 1) / nact;
  // piece ownership inside a group (conflict-free LDS.128 in contiguous mode): lane i of a group owns the 8-element
  // pieces i + lpg * ((k + rot) & 3), k = 0..3, rot = index of the group among the groups of an 8-lane phase
  const int gi = lane & (lpg - 1);                              // lpg divides 32
  const int rot = lpg < 8 ? (lane >> lpg_sh) & 3 : 0;
  int piece_b[4];                                               // byte offset of piece k inside the group's codes
#pragma unroll
  for (int k = 0; k < 4; ++k) piece_b[k] = (gi + lpg * ((k + rot) & 3)) * 8;
  // (row, lane slot u) of this thread's first item and the per-iteration step, without divisions in the loop
  int row0 = owner ? tid / nchunk : 0, u0 = owner ? tid - row0 * nchunk : 0;
  const int step_r = nact / nchunk, step_u = nact - step_r * nchunk;    // gather mode: (rpp, 0)
  uint32_t off[GATHER ? kElemsPerLane : 1];
  if (GATHER) {
    const int grp = u0 >> lpg_sh;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
#pragma unroll
      for (int t = 0; t < 8; ++t) {
        const int pos = grp * q.g + piece_b[k] + t;
        uint32_t o = zero_off;
        if (owner && pos < Kn) o = (uint32_t)(q.perm ? q.perm[pos] : pos) * 2u;
        off[k * 8 + t] = o + (uint32_t)row0 * pitch_b;
      }
    }
  }

  for (int64_t i = 0; i < my_count; ++i) {
    const int s = (int)(i % kStagesH2);
    if (!mbar_wait(&full_bar[s], (uint32_t)((i / kStagesH2) & 1), 0x100u + s)) return;
    const int64_t r0 = (my_first + i * gridDim.x) * a.rps;
    const int nr = (int)min((int64_t)a.rps, q.R - r0);
    const uint32_t sbase = srows_a + (uint32_t)s * stage_b;
    if (NORM) {
      norm_stage<T, NORM>(a, sbase, pitch_b, nr, r0, K, w_sm, b_sm, psum, warp, lane);
      fence_proxy_async_smem();     // the next bulk copy into this stage is ordered after these generic-proxy stores
      __syncwarp();
      if (lane == 0) mbar_arrive(&norm_bar[s]);
      if (!mbar_wait(&norm_bar[s], (uint32_t)((i / kStagesH2) & 1), 0x138u + s)) return;
    }

    int row = row0, u = u0;
    for (int ii = 0; ii < n_iter; ++ii) {
      const bool live = owner && row < nr;
      const int grp = u >> lpg_sh;
      uint32_t x[16];
      if (GATHER) {
        // CTA-uniform base of this pass; the lane-0 shuffle makes it provably warp-uniform so that every gather is
        // one LDS.U16 [R + UR] without an address add
        const uint32_t pbase = __shfl_sync(0xffffffffu, sbase + (uint32_t)(ii * a.rpp) * pitch_b, 0);
#pragma unroll
        for (int e = 0; e < 16; ++e)
          x[e] = __byte_perm(lds_u16(pbase + off[2 * e]), lds_u16(pbase + off[2 * e + 1]), 0x5410);
      } else {
        const uint32_t rb = sbase + (uint32_t)min(row, a.rps - 1) * pitch_b + (uint32_t)(grp * q.g) * 2u;
        const int e0 = grp * q.g;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const uint4 v = lds_u128(rb + (uint32_t)piece_b[k] * 2u);
          x[k * 4 + 0] = v.x; x[k * 4 + 1] = v.y; x[k * 4 + 2] = v.z; x[k * 4 + 3] = v.w;
          const int pe = e0 + piece_b[k];                // first permuted position of the piece
          if (pe + 8 > Kn) {                             // positions >= Kn are zero padding (fake_quant.py:84-86)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const uint32_t keep = (pe + 2 * j < Kn ? 0x0000FFFFu : 0u) | (pe + 2 * j + 1 < Kn ? 0xFFFF0000u : 0u);
              x[k * 4 + j] &= keep;
            }
          }
        }
      }
      // group abs-max: 16 packed max, fold the two halves, butterfly over the lanes of the group
      uint32_t am = 0u;
#pragma unroll
      for (int e = 0; e < 16; ++e) am = pabsmax<V2>(am, x[e]);
      am = pmax<V2>(am, __byte_perm(am, am, 0x1032));
      for (int o = 1; o < lpg; o <<= 1) am = pmax<V2>(am, __shfl_xor_sync(0xffffffffu, am, o));
      const GroupK<T> gk = group_consts<T>(am, q.qmax, a.rq, tbl);

      uint32_t cb[16];
      if (gk.fast) {
#pragma unroll
        for (int e = 0; e < 16; ++e) {
          const uint32_t xs = Pk<T>::kScale ? pmul<V2>(x[e], gk.p2) : x[e];
          const uint32_t q0 = pmul<V2>(xs, gk.inv2);
          const uint32_t rem = pfma<V2>(gk.ns2, q0, xs);
          const uint32_t q1 = pfma<V2>(rem, gk.inv2, q0);
          cb[e] = padd<V2>(q1, Pk<T>::kMagic2);
        }
      } else {                                     // exact fp32 path (denormal / extreme scales)
        const float inv = __frcp_rn(gk.scale);
#pragma unroll
        for (int e = 0; e < 16; ++e) {
          const uint32_t b0 = code_bits_exact<T>(Pk<T>::to_f(x[e] & 0xFFFFu), gk.scale, inv);
          const uint32_t b1 = code_bits_exact<T>(Pk<T>::to_f(x[e] >> 16), gk.scale, inv);
          cb[e] = ((b0 + Pk<T>::kBias) & 0xFFu) | (((b1 + Pk<T>::kBias) & 0xFFu) << 16);
        }
      }
      if (live) {
        int8_t* crow = q.codes + (r0 + row) * q.ld_codes + (int64_t)grp * q.g;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          uint2 w;
          w.x = __byte_perm(cb[k * 4 + 0], cb[k * 4 + 1], 0x6420);
          w.y = __byte_perm(cb[k * 4 + 2], cb[k * 4 + 3], 0x6420);
          if (Pk<T>::kBias) {          // bytes hold 64 + code in [0, 127]: remove the bias, sign-extend bit 6 into bit 7
            w.x ^= 0x40404040u; w.x |= (w.x & 0x40404040u) << 1;
            w.y ^= 0x40404040u; w.y |= (w.y & 0x40404040u) << 1;
          }
          *reinterpret_cast<uint2*>(crow + piece_b[k]) = w;
        }
        if (gi == 0 && grp < q.G) q.scales[(int64_t)grp * q.ld_scales + (r0 + row)] = gk.scale;
      }
      row += step_r;
      u += step_u;
      if (u >= nchunk) {
        u -= nchunk;
        ++row;
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty_bar[s]);
  }
}

struct H2Plan {
  bool ok;
  bool gather;
  H2Args args;
  size_t smem;
};

template <typename T>
H2Plan plan_h2(const QuantArgs& q, int n_bits, int x_blocks, int64_t blk_cols, int64_t blk_stride, int norm = 0) {
  H2Plan p{};
  p.ok = false;
  if (norm && (x_blocks > 1 || q.K > 16384)) return p;
  if (sizeof(T) != 2) return p;
  const int g = q.g;
  if (g < 32 || g > 1024 || (g & (g - 1)) != 0) return p;
  if (q.smooth || q.g_int != q.G || q.G < 1) return p;
  if (n_bits > Pk<T>::kMaxBits) return p;
  if (q.K < 8 || q.K % 8 != 0 || (q.ld * 2) % 16 != 0 || reinterpret_cast<uintptr_t>(q.x) % 16 != 0) return p;
  if (q.ld_codes % 32 != 0 || q.ld_codes < (int64_t)q.G * g || reinterpret_cast<uintptr_t>(q.codes) % 16 != 0) return p;
  if (q.Ks > 0 && ((q.ld_sal * 2) % 16 != 0 || reinterpret_cast<uintptr_t>(q.out_sal) % 16 != 0)) return p;
  if (x_blocks > 1 && (blk_cols % 8 != 0 || blk_stride % 8 != 0)) return p;
  if (x_blocks == 1 && q.ld < q.K) return p;
  H2Args& a = p.args;
  a.q = q;
  a.nchunk = (int)(q.ld_codes / 32);
  a.lpg = g / 32;
  if (a.nchunk % a.lpg != 0) return p;                          // dummy zero groups must be whole groups
  a.x_blocks = x_blocks;
  a.blk_cols = x_blocks > 1 ? blk_cols : q.K;
  a.blk_stride = x_blocks > 1 ? blk_stride : 0;
  a.rq = 1.0f / q.qmax;
  p.gather = q.perm != nullptr;
  const int64_t need = std::max<int64_t>(q.K, (int64_t)a.nchunk * 32);
  a.pitch = (int)((need + kRowPadH2 + 7) / 8 * 8);
  if ((size_t)a.pitch * 2 > 65535 * 2) return p;
  const size_t row_bytes = (size_t)a.pitch * 2;
  const int vec = 8;
  a.q.sal_cols = q.Ks > 0 ? (int)((q.sal_width + vec - 1) / vec * vec) : 0;
  const size_t fixed = 128 + (size_t)(1 << Pk<T>::kMant) * 2 + (size_t)a.q.sal_cols * 2 + 256 +
                       (norm ? kNormMaxRows * 8 + (size_t)q.K * 2 * (norm == 2 ? 2 : 1) + 128 : 0);
  const size_t budget = 226 * 1024;
  if (fixed + (size_t)kStagesH2 * row_bytes > budget) return p;
  int fit = (int)((budget - fixed) / ((size_t)kStagesH2 * row_bytes));
  if (p.gather) {
    if (a.nchunk > kConsThreads) return p;                      // one lane per chunk: rows wider than 16 K positions
    a.rpp = kConsThreads / a.nchunk;
    if ((double)(a.rpp * a.nchunk) / kConsThreads < 0.6) return p;   // too many idle lanes: quant_rows_fast is better
    if (a.rpp > fit) a.rpp = fit;
    if (a.rpp < 1) return p;
    int passes = std::min(fit / a.rpp, std::max(1, 8 / a.rpp));
    if (passes < 1) passes = 1;
    a.rps = a.rpp * passes;
    if ((int64_t)a.pitch * 2 * a.rpp > 0x7fffffff) return p;
  } else {
    a.rpp = 0;
    a.rps = std::min(fit, 8);
  }
  if (a.rps > q.R) a.rps = (int)std::max<int64_t>(1, q.R);
  if (p.gather && a.rps % a.rpp != 0) a.rps = (a.rps + a.rpp - 1) / a.rpp * a.rpp;   // whole passes (extra rows are masked)
  if ((size_t)a.rps * q.K * 2 >= (1u << 20)) return p;          // mbarrier tx-count limit
  p.smem = fixed / 128 * 128 + 128 + (size_t)kStagesH2 * a.rps * row_bytes;
  if (p.smem > 227 * 1024) return p;
  a.norm = norm;
  if (norm) {
    if (a.rps > kNormMaxRows) return p;
    int parts = 1;
    while (parts * 2 * a.rps <= kConsWarps) parts *= 2;           // consumer warps sharing a row in the norm phase
    a.norm_parts = parts;
  }
  p.ok = true;
  return p;
}

template <typename T>
int launch_h2(const H2Plan& p, cudaStream_t st) {
  const H2Args& a = p.args;
  const int64_t nblocks = (a.q.R + a.rps - 1) / a.rps;
  int64_t grid = mq_worker_sms();
  if (grid > nblocks) grid = nblocks;
  auto go = [&](auto kern) -> int {
    MQ_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.smem));
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(kThreadsH2);
    cfg.dynamicSmemBytes = p.smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = mq_get_option("pdl", 1) == 1 ? 1 : 0;
    MQ_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kern, a));
    MQ_CHECK_CUDA(cudaGetLastError());
    return MQ_OK;
  };
  if (a.norm == 1) return p.gather ? go(quant_rows_h2<T, true, 1>) : go(quant_rows_h2<T, false, 1>);
  if (a.norm == 2) return p.gather ? go(quant_rows_h2<T, true, 2>) : go(quant_rows_h2<T, false, 2>);
  return p.gather ? go(quant_rows_h2<T, true>) : go(quant_rows_h2<T, false>);
}

}  // namespace h2
