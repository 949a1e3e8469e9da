// K2d: the decode regime of the W4A4 MixedQuant linear (M <= 64 tokens; BASELINE config 5) -- weight-bandwidth bound.
//
//   y[m, n] = sum_g a_scale[g, m] * w_scale[g, n] * sum_{k in g} a_code[m, k] * w_code[n, k] + a_sal[m, :] . w_sal[n, :] + b[n]
//
// Replaces the same  F.linear(q_x, self.weight, self.bias)  (/root/reference/smoothquant/fake_quant.py:306) as
// mixed_gemm.cu, for the shapes where the prefill kernel's 128-row MMA tiles are almost empty and the time is the time
// to stream the weights once.  Here the weight operand is INT4-RESIDENT: two's-complement nibbles, two codes per byte
// (the reference's own accounting: 4 bits per non-salient weight, smoothquant/model_size.py:5-15), in a layout made
// offline by mq_pack_decode so that the kernel never shuffles a weight byte:
//
//   swap-AB: the W rows are the M axis of the MMA (16 rows per warp tile), the <= 64 tokens are its N axis (1..8 n8
//   tiles), so one pass over a weight byte serves every token.
//   nibbles [row tile of 16][64-column unit][lane][16 B]: a lane's 16 bytes are exactly its four A-fragment registers
//     of mma.m16n8k32.s8 for TWO k32 steps -- low nibbles step 0, high nibbles step 1.  Unpacking is
//     (p << 4) & 0xF0F0F0F0 and p & 0xF0F0F0F0: each byte becomes 16 * code as an int8, and the 1/16 is folded into
//     the weight scale (exact).  Three ALU instructions per 8 codes, no table, no sign fix-up.
//   salient slice [row tile][16-column step][lane][16 B] = the A fragments of mma.m16n8k16 (f16 / bf16).
//
// The arithmetic engine is mma.sync (IMMA.16832.S8 / HMMA.16816): at M <= 64 the tensor work is 1-2 % of tcgen05's
// peak and the kernel is bound by HBM, not by the tensor pipe; what matters is bytes in flight, so a producer warp keeps
// a 6-stage ring of 256-column blocks filled with bulk TMA copies (weights: one contiguous 2 KB copy per row tile and
// block; activation codes / salient columns: one 128B-swizzled tensor tile per block, re-read from L2 by every CTA),
// and 8 compute warps each own one 16-row tile of the round.  Every CTA gets a contiguous, byte-balanced range of row
// tiles.  The per-group rescale is the prefill kernel's, operation for operation (exact int32 -> fp32 through the
// 1.5 * 2^23 bias pattern, t = v * as exact, acc = fma(t, ws, acc) group by group), so both kernels agree to the last
// bit on the integer part.
#include <cuda.h>

#include "common.cuh"
#include "ptx.cuh"
#include "tma_host.cuh"

namespace {

using namespace mqptx;

constexpr int kDWarps = 8;                       // compute warps = row tiles per round
constexpr int kDThreads = (kDWarps + 1) * 32;    // + producer warp
constexpr int kDStages = 6;
constexpr int kBlockK = 256;                     // quantised columns per stage (4 units of 64)
constexpr int kTileWBytes = 2048;                // one row tile's weights per stage: 16 rows x 256 nibbles, or 16 x 64 halves
constexpr int kScaleSlots = 4;                   // groups that can end inside one block (g >= 64)

struct DecodeArgs {
  const uint8_t* w_nib;       // [tiles][units][32][16]
  const uint8_t* w_sal;       // [tiles][steps16 padded to 4][32][16]
  const float* a_scales;      // [G][ld_as]
  int64_t ld_as;
  const float* w_scales;      // [G][ld_ws]
  int64_t ld_ws;
  const void* bias;
  void* y;
  int64_t ldy;
  int M, N, G, g;
  int units;                  // 64-column units that hold quantised columns = ceil(G * g / 64)
  int nkb;                    // int blocks  = ceil(units / 4)
  int nsb;                    // salient blocks = ceil(Ks / 64)
  int tiles;                  // ceil(N / 16)
  int64_t nib_tile_stride;    // bytes between row tiles in w_nib  = nkb * 2048
  int64_t sal_tile_stride;    // bytes between row tiles in w_sal  = nsb * 2048
  int as_bytes;               // bytes of one activation-scale row copied per group (multiple of 16)
};

template <int NT>
struct DSmem {
  static constexpr int kMp = NT * 8;
  static constexpr int kActBytes = 2 * kMp * 128;            // two 128B-swizzled tiles [Mp x 128 B]
  static constexpr int kScaleBytes = 1024;                   // kScaleSlots x Mp floats (<= 1 KB)
  static constexpr int kWBytes = kDWarps * kTileWBytes;
  static constexpr int kStageBytes = kActBytes + kScaleBytes + kWBytes;   // multiple of 1024
  static constexpr int kBarOffset = kDStages * kStageBytes;
  static constexpr int kTotal = kBarOffset + 2 * kDStages * 8 + 1024;     // + slack to align the base to 1024 B
  static_assert(kStageBytes % 1024 == 0, "stage alignment");
  static_assert(kScaleSlots * kMp * 4 <= kScaleBytes, "scale slots");
};

__device__ __forceinline__ void ldmatrix_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldmatrix_x2(uint32_t addr, uint32_t& r0, uint32_t& r1) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x2.shared.b16 {%0, %1}, [%2];" : "=r"(r0), "=r"(r1) : "r"(addr));
}
__device__ __forceinline__ void imma_16832(int (&c)[4], const uint32_t* a, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
               : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
template <typename T>
__device__ __forceinline__ void hmma_16816(float (&c)[4], const uint4& a, uint32_t b0, uint32_t b1);
template <>
__device__ __forceinline__ void hmma_16816<__half>(float (&c)[4], const uint4& a, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a.x), "r"(a.y), "r"(a.z), "r"(a.w), "r"(b0), "r"(b1));
}
template <>
__device__ __forceinline__ void hmma_16816<__nv_bfloat16>(float (&c)[4], const uint4& a, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a.x), "r"(a.y), "r"(a.z), "r"(a.w), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint4 lds_v4(uint32_t addr) {
  uint4 r;
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(addr));
  return r;
}
__device__ __forceinline__ float2 lds_f2(uint32_t addr) {
  float2 r;
  asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(r.x), "=f"(r.y) : "r"(addr));
  return r;
}

// ------------------------------------------------------------------------------------------------ offline packers
// nibble word (tile, unit, lane, pi): byte j = code[r][k + j] & 15 | code[r][k + 32 + j] << 4 with
// r = 16 * tile + lane / 4 + 8 * (pi & 1),  k = 64 * unit + 16 * (pi >> 1) + 4 * (lane & 3)
__global__ void pack_decode_nib_kernel(const int8_t* __restrict__ codes, int64_t N, int64_t Kc, int64_t ldw, int units_padded,
                                       uint32_t* __restrict__ out, int64_t words) {
  for (int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; w < words; w += (int64_t)gridDim.x * blockDim.x) {
    const int pi = (int)(w & 3), lane = (int)((w >> 2) & 31);
    const int64_t tu = w >> 7;
    const int unit = (int)(tu % units_padded);
    const int64_t tile = tu / units_padded;
    const int64_t r = tile * 16 + (lane >> 2) + 8 * (pi & 1);
    const int64_t k = (int64_t)unit * 64 + 16 * (pi >> 1) + 4 * (lane & 3);
    uint32_t v = 0;
    if (r < N) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const uint32_t lo = k + j < Kc ? (uint32_t)codes[r * ldw + k + j] & 15u : 0u;
        const uint32_t hi = k + 32 + j < Kc ? (uint32_t)codes[r * ldw + k + 32 + j] & 15u : 0u;
        v |= (lo | (hi << 4)) << (8 * j);
      }
    }
    out[w] = v;
  }
}
// salient word (tile, step16, lane, pi): halves w_sal[r][k], w_sal[r][k + 1] with
// r = 16 * tile + lane / 4 + 8 * (pi & 1),  k = 16 * step + 8 * (pi >> 1) + 2 * (lane & 3)
__global__ void pack_decode_sal_kernel(const uint16_t* __restrict__ w_sal, int64_t N, int64_t Ks, int64_t ld, int steps_padded,
                                       uint32_t* __restrict__ out, int64_t words) {
  for (int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; w < words; w += (int64_t)gridDim.x * blockDim.x) {
    const int pi = (int)(w & 3), lane = (int)((w >> 2) & 31);
    const int64_t ts = w >> 7;
    const int step = (int)(ts % steps_padded);
    const int64_t tile = ts / steps_padded;
    const int64_t r = tile * 16 + (lane >> 2) + 8 * (pi & 1);
    const int64_t k = (int64_t)step * 16 + 8 * (pi >> 1) + 2 * (lane & 3);
    uint32_t v = 0;
    if (r < N) {
      if (k < Ks) v |= w_sal[r * ld + k];
      if (k + 1 < Ks) v |= (uint32_t)w_sal[r * ld + k + 1] << 16;
    }
    out[w] = v;
  }
}

// ------------------------------------------------------------------------------------------------ the kernel
constexpr float kMagicF = 12582912.0f;          // 1.5 * 2^23: bits(kMagicF) + v are the bits of the float kMagicF + v
constexpr uint32_t kMagicI = 0x4B400000u;

template <typename T, int NT>
__global__ void __launch_bounds__(kDThreads, 1)
decode_gemm_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_as,
                   const DecodeArgs p) {
  using S = DSmem<NT>;
  constexpr int Mp = S::kMp;
  extern __shared__ unsigned char smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  unsigned char* smem = smem_raw + (((raw_addr + 1023u) & ~1023u) - raw_addr);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + S::kBarOffset);
  uint64_t* empty_bar = full_bar + kDStages;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // contiguous, balanced range of 16-row tiles for this CTA; processed in rounds of kDWarps tiles (one per warp)
  const int t0 = (int)((int64_t)p.tiles * blockIdx.x / gridDim.x);
  const int t1 = (int)((int64_t)p.tiles * (blockIdx.x + 1) / gridDim.x);
  const int rounds = (t1 - t0 + kDWarps - 1) / kDWarps;
  const int blocks = p.nkb + p.nsb;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kDStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], kDWarps);
    }
    fence_mbar_init();
    tma_prefetch_desc(&map_a);
    tma_prefetch_desc(&map_as);
  }
  __syncthreads();

  if (warp == kDWarps) {
    // ------------------------------------------------------------------ producer warp (lane 0 issues)
    if (lane != 0) return;
    uint32_t it = 0;
    for (int r = 0; r < rounds; ++r) {
      const int tb = t0 + r * kDWarps;
      const int nt = min(kDWarps, t1 - tb);
      for (int b = 0; b < blocks; ++b, ++it) {
        const uint32_t s = it % kDStages;
        if (it >= (uint32_t)kDStages && !mbar_wait(&empty_bar[s], ((it / kDStages) - 1u) & 1u, 0x820u + s)) return;
        unsigned char* st = smem + (size_t)s * S::kStageBytes;
        unsigned char* sw = st + S::kActBytes + S::kScaleBytes;
        if (b < p.nkb) {
          const bool second = (b * 4 + 2) < p.units;              // the block's second 128-column tile holds real columns
          int ngrp = 0;                                            // groups whose last column lies in this block
          const int k_lo = b * kBlockK, k_hi = k_lo + kBlockK;
          const int g_first = k_lo / p.g;                          // first group intersecting the block
          for (int gi = g_first; gi < p.G && (gi + 1) * p.g <= k_hi; ++gi)
            if ((gi + 1) * p.g > k_lo) ++ngrp;
          mbar_expect_tx(&full_bar[s], (uint32_t)((second ? 2 : 1) * Mp * 128 + ngrp * p.as_bytes + nt * kTileWBytes));
          tma_load_2d(st, &map_a, k_lo, 0, &full_bar[s]);
          if (second) tma_load_2d(st + Mp * 128, &map_a, k_lo + 128, 0, &full_bar[s]);
          for (int gi = g_first; gi < p.G && (gi + 1) * p.g <= k_hi; ++gi) {
            const int end = (gi + 1) * p.g;
            if (end <= k_lo) continue;
            const int slot = p.g <= kBlockK ? (end - k_lo) / p.g - 1 : 0;
            bulk_g2s(st + S::kActBytes + slot * Mp * 4, p.a_scales + (int64_t)gi * p.ld_as, (uint32_t)p.as_bytes, &full_bar[s]);
          }
          for (int i = 0; i < nt; ++i)
            bulk_g2s(sw + i * kTileWBytes, p.w_nib + (int64_t)(tb + i) * p.nib_tile_stride + (int64_t)b * kTileWBytes,
                     kTileWBytes, &full_bar[s]);
        } else {
          const int sb = b - p.nkb;
          mbar_expect_tx(&full_bar[s], (uint32_t)(Mp * 128 + nt * kTileWBytes));
          tma_load_2d(st, &map_as, sb * 64, 0, &full_bar[s]);
          for (int i = 0; i < nt; ++i)
            bulk_g2s(sw + i * kTileWBytes, p.w_sal + (int64_t)(tb + i) * p.sal_tile_stride + (int64_t)sb * kTileWBytes,
                     kTileWBytes, &full_bar[s]);
        }
      }
    }
    return;
  }

  // -------------------------------------------------------------------- compute warps
  const int gid = lane >> 2, tq = lane & 3;
  // ldmatrix row address of this lane: matrix q = lane / 8 (x4: n-tile pair j + q / 2, 16-byte chunk q & 1), row lane % 8
  const int lq = lane >> 3, lr = lane & 7;
  const uint32_t smem_a = smem_u32(smem);
  uint32_t it = 0;
  for (int r = 0; r < rounds; ++r) {
    const int tb = t0 + r * kDWarps;
    const int nt = min(kDWarps, t1 - tb);
    const bool active = warp < nt;
    const int n0 = (tb + warp) * 16;
    const int nA = n0 + gid, nB = nA + 8;
    float acc[NT][4];
    int ci[NT][4];
#pragma unroll
    for (int j = 0; j < NT; ++j)
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        acc[j][c] = 0.f;
        ci[j][c] = 0;
      }
    float wsA = 0.f, wsB = 0.f;

    for (int b = 0; b < p.nkb; ++b, ++it) {
      const uint32_t s = it % kDStages;
      if (!mbar_wait(&full_bar[s], (it / kDStages) & 1u, 0x800u + s)) return;
      if (active) {
        const uint32_t st = smem_a + s * (uint32_t)S::kStageBytes;
        const uint32_t sw = st + (uint32_t)(S::kActBytes + S::kScaleBytes) + (uint32_t)warp * kTileWBytes + (uint32_t)lane * 16u;
#pragma unroll
        for (int uu = 0; uu < 4; ++uu) {
          const int u = b * 4 + uu;
          if (u < p.units) {                                       // warp-uniform
            const int k0 = u * 64;
            if (k0 % p.g == 0) {                                   // a group starts: fetch its weight scales (1/16 folded in)
              const int gi = k0 / p.g;
              wsA = nA < p.N ? p.w_scales[(int64_t)gi * p.ld_ws + nA] * 0.0625f : 0.f;
              wsB = nB < p.N ? p.w_scales[(int64_t)gi * p.ld_ws + nB] * 0.0625f : 0.f;
            }
            const uint4 pk = lds_v4(sw + (uint32_t)uu * 512u);
            const uint32_t a_lo[4] = {(pk.x << 4) & 0xF0F0F0F0u, (pk.y << 4) & 0xF0F0F0F0u, (pk.z << 4) & 0xF0F0F0F0u,
                                      (pk.w << 4) & 0xF0F0F0F0u};
            const uint32_t a_hi[4] = {pk.x & 0xF0F0F0F0u, pk.y & 0xF0F0F0F0u, pk.z & 0xF0F0F0F0u, pk.w & 0xF0F0F0F0u};
            const uint32_t tile = st + (uint32_t)(uu >> 1) * (uint32_t)(Mp * 128);
#pragma unroll
            for (int hs = 0; hs < 2; ++hs) {                       // the two k32 steps of the unit
              const int chunk = ((uu & 1) * 2 + hs) * 2;           // first 16-byte chunk of the step inside the 128-B row
              if (NT == 1) {
                uint32_t b0, b1;
                ldmatrix_x2(tile + (uint32_t)lr * 128u + (uint32_t)(((chunk + (lq & 1)) ^ lr) << 4), b0, b1);
                imma_16832(ci[0], hs ? a_hi : a_lo, b0, b1);
              } else {
#pragma unroll
                for (int j = 0; j < NT; j += 2) {
                  uint32_t b0, b1, b2, b3;
                  const int m = 8 * (j + (lq >> 1)) + lr;
                  ldmatrix_x4(tile + (uint32_t)m * 128u + (uint32_t)(((chunk + (lq & 1)) ^ lr) << 4), b0, b1, b2, b3);
                  imma_16832(ci[j], hs ? a_hi : a_lo, b0, b1);
                  imma_16832(ci[j + 1], hs ? a_hi : a_lo, b2, b3);
                }
              }
            }
            if ((k0 + 64) % p.g == 0) {                            // the group ends with this unit: rescale into fp32
              const int end = k0 + 64 - b * kBlockK;
              const int slot = p.g <= kBlockK ? end / p.g - 1 : 0;
              const uint32_t sa = st + (uint32_t)S::kActBytes + (uint32_t)(slot * Mp * 4) + (uint32_t)(tq * 8);
#pragma unroll
              for (int j = 0; j < NT; ++j) {
                const float2 as2 = lds_f2(sa + (uint32_t)(j * 32));
                const float ax = __uint_as_float(__float_as_uint(as2.x) & 0xFFFFFFFCu);
                const float ay = __uint_as_float(__float_as_uint(as2.y) & 0xFFFFFFFCu);
                const float cx = -kMagicF * ax, cy = -kMagicF * ay;
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                  const float f = __uint_as_float((uint32_t)ci[j][c] + kMagicI);
                  const float t = (c & 1) ? fmaf(f, ay, cy) : fmaf(f, ax, cx);       // == 16 * sum * as exactly
                  acc[j][c] = fmaf(t, (c & 2) ? wsB : wsA, acc[j][c]);
                  ci[j][c] = 0;
                }
              }
            }
          }
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty_bar[s]);
    }

    // salient slice: fp32 accumulators of its own (the prefill kernel adds the slice's sum to the integer part last)
    float cs[NT][4];
#pragma unroll
    for (int j = 0; j < NT; ++j)
#pragma unroll
      for (int c = 0; c < 4; ++c) cs[j][c] = 0.f;
    for (int sb = 0; sb < p.nsb; ++sb, ++it) {
      const uint32_t s = it % kDStages;
      if (!mbar_wait(&full_bar[s], (it / kDStages) & 1u, 0x810u + s)) return;
      if (active) {
        const uint32_t st = smem_a + s * (uint32_t)S::kStageBytes;
        const uint32_t sw = st + (uint32_t)(S::kActBytes + S::kScaleBytes) + (uint32_t)warp * kTileWBytes + (uint32_t)lane * 16u;
#pragma unroll
        for (int k16 = 0; k16 < 4; ++k16) {
          const uint4 a = lds_v4(sw + (uint32_t)k16 * 512u);
          const int chunk = k16 * 2;
          if (NT == 1) {
            uint32_t b0, b1;
            ldmatrix_x2(st + (uint32_t)lr * 128u + (uint32_t)(((chunk + (lq & 1)) ^ lr) << 4), b0, b1);
            hmma_16816<T>(cs[0], a, b0, b1);
          } else {
#pragma unroll
            for (int j = 0; j < NT; j += 2) {
              uint32_t b0, b1, b2, b3;
              const int m = 8 * (j + (lq >> 1)) + lr;
              ldmatrix_x4(st + (uint32_t)m * 128u + (uint32_t)(((chunk + (lq & 1)) ^ lr) << 4), b0, b1, b2, b3);
              hmma_16816<T>(cs[j], a, b0, b1);
              hmma_16816<T>(cs[j + 1], a, b2, b3);
            }
          }
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty_bar[s]);
    }

    if (active) {
      const T* bias = static_cast<const T*>(p.bias);
      T* y = static_cast<T*>(p.y);
      const float bA = bias && nA < p.N ? Num<T>::to_f(bias[nA]) : 0.f;
      const float bB = bias && nB < p.N ? Num<T>::to_f(bias[nB]) : 0.f;
#pragma unroll
      for (int j = 0; j < NT; ++j)
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const int m = 8 * j + 2 * tq + (c & 1);
          const int n = (c & 2) ? nB : nA;
          if (m < p.M && n < p.N)
            y[(int64_t)m * p.ldy + n] = Num<T>::from_f(acc[j][c] + cs[j][c] + ((c & 2) ? bB : bA));
        }
    }
  }
}

// ------------------------------------------------------------------------------------------------ host side
template <typename T, int NT>
int launch_decode(const CUtensorMap& ma, const CUtensorMap& mas, const DecodeArgs& p, cudaStream_t st) {
  using S = DSmem<NT>;
  auto kern = decode_gemm_kernel<T, NT>;
  if (int rc = mq_ensure_smem_attr(reinterpret_cast<const void*>(kern), S::kTotal)) return rc;
  int grid = mq_sm_count();
  if (grid > p.tiles) grid = p.tiles;
  kern<<<grid, kDThreads, S::kTotal, st>>>(ma, mas, p);
  MQ_CHECK_CUDA(cudaGetLastError());
  return MQ_OK;
}

template <typename T>
int dispatch_decode(const CUtensorMap& ma, const CUtensorMap& mas, const DecodeArgs& p, cudaStream_t st) {
  if (p.M <= 8) return launch_decode<T, 1>(ma, mas, p, st);
  if (p.M <= 16) return launch_decode<T, 2>(ma, mas, p, st);
  if (p.M <= 32) return launch_decode<T, 4>(ma, mas, p, st);
  return launch_decode<T, 8>(ma, mas, p, st);
}

int mp_of(int64_t M) { return M <= 8 ? 8 : M <= 16 ? 16 : M <= 32 ? 32 : 64; }

}  // namespace

extern "C" {

int mq_decode_weight_bytes(int64_t N, int64_t G, int group_size, int64_t Ks, size_t* nib_bytes, size_t* sal_bytes) {
  MQ_CHECK_ARG(N >= 1 && G >= 1 && group_size >= 64 && group_size % 64 == 0 && Ks >= 0,
               "mq_decode_weight_bytes: needs N >= 1, G >= 1, group_size a multiple of 64, Ks >= 0");
  const int64_t tiles = (N + 15) / 16;
  const int64_t units = G * group_size / 64, nkb = (units + 3) / 4, nsb = (Ks + 63) / 64;
  if (nib_bytes) *nib_bytes = (size_t)(tiles * nkb * kTileWBytes);
  if (sal_bytes) *sal_bytes = (size_t)(tiles * nsb * kTileWBytes);
  return MQ_OK;
}

int mq_pack_decode(const int8_t* w_codes, int64_t ldw, const void* w_sal, int64_t ld_wsal, int64_t N, int64_t G,
                   int group_size, int64_t Ks, int dtype, uint8_t* out_nib, uint8_t* out_sal, void* stream) {
  MQ_CHECK_ARG(w_codes && out_nib, "mq_pack_decode: null pointer");
  MQ_CHECK_ARG(N >= 1 && G >= 1 && group_size >= 64 && group_size % 64 == 0,
               "mq_pack_decode: needs N >= 1, G >= 1 and a group size that is a multiple of 64");
  MQ_CHECK_ARG(ldw >= G * group_size, "mq_pack_decode: ldw=%lld < G*g", (long long)ldw);
  MQ_CHECK_ARG(Ks == 0 || (w_sal && out_sal && ld_wsal >= Ks), "mq_pack_decode: bad salient arguments");
  MQ_CHECK_ARG(dtype == MQ_F16 || dtype == MQ_BF16, "mq_pack_decode: dtype must be MQ_F16 or MQ_BF16");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int64_t tiles = (N + 15) / 16;
  const int64_t units = G * group_size / 64, nkb = (units + 3) / 4, nsb = (Ks + 63) / 64;
  const int64_t cap = (int64_t)mq_sm_count() * 16;
  {
    const int64_t words = tiles * nkb * 4 * 128;
    int64_t blocks = (words + 255) / 256;
    if (blocks > cap) blocks = cap;
    pack_decode_nib_kernel<<<(unsigned)blocks, 256, 0, st>>>(w_codes, N, G * group_size, ldw, (int)(nkb * 4),
                                                            reinterpret_cast<uint32_t*>(out_nib), words);
    MQ_CHECK_CUDA(cudaGetLastError());
  }
  if (Ks > 0) {
    const int64_t words = tiles * nsb * 4 * 128;
    int64_t blocks = (words + 255) / 256;
    if (blocks > cap) blocks = cap;
    pack_decode_sal_kernel<<<(unsigned)blocks, 256, 0, st>>>(static_cast<const uint16_t*>(w_sal), N, Ks, ld_wsal,
                                                            (int)(nsb * 4), reinterpret_cast<uint32_t*>(out_sal), words);
    MQ_CHECK_CUDA(cudaGetLastError());
  }
  return MQ_OK;
}

int mq_decode_gemm(const int8_t* a_codes, int64_t lda, const float* a_scales, int64_t ld_as, const void* a_sal,
                   int64_t ld_asal, const uint8_t* w_nib, const float* w_scales, int64_t ld_ws, const uint8_t* w_sal_dec,
                   const void* bias, int64_t M, int64_t N, int64_t G, int group_size, int64_t Ks, int dtype, void* y,
                   int64_t ldy, void* stream) {
  MQ_CHECK_ARG(y && M >= 0 && M <= 64 && N >= 1 && ldy >= N, "mq_decode_gemm: needs 0 <= M <= 64, N >= 1, ldy >= N (M=%lld)",
               (long long)M);
  MQ_CHECK_ARG(dtype == MQ_F16 || dtype == MQ_BF16, "mq_decode_gemm: dtype must be MQ_F16 or MQ_BF16");
  MQ_CHECK_ARG(G >= 1 && group_size >= 64 && group_size % 64 == 0 && group_size <= 4096,
               "mq_decode_gemm: needs G >= 1 and a group size that is a multiple of 64 (got G=%lld g=%d)", (long long)G,
               group_size);
  MQ_CHECK_ARG(a_codes && a_scales && w_nib && w_scales, "mq_decode_gemm: null code / scale pointer");
  MQ_CHECK_ARG(lda >= G * group_size && lda % 16 == 0 && reinterpret_cast<uintptr_t>(a_codes) % 16 == 0,
               "mq_decode_gemm: lda=%lld must be a multiple of 16 and >= G*g", (long long)lda);
  MQ_CHECK_ARG(ld_as >= M && ld_as % 4 == 0 && ld_ws >= N && reinterpret_cast<uintptr_t>(a_scales) % 16 == 0,
               "mq_decode_gemm: scales need ld_as >= M, ld_as %% 4 == 0, ld_ws >= N, 16-byte aligned a_scales");
  MQ_CHECK_ARG(reinterpret_cast<uintptr_t>(w_nib) % 16 == 0, "mq_decode_gemm: w_nib must be 16-byte aligned");
  if (Ks > 0) {
    MQ_CHECK_ARG(a_sal && w_sal_dec && ld_asal >= Ks && ld_asal % 8 == 0 && reinterpret_cast<uintptr_t>(a_sal) % 16 == 0 &&
                     reinterpret_cast<uintptr_t>(w_sal_dec) % 16 == 0,
                 "mq_decode_gemm: salient operands need ld_asal >= Ks, ld_asal %% 8 == 0 and 16-byte alignment");
  }
  if (M == 0) return MQ_OK;
  const int mp = mp_of(M);
  CUtensorMap ma, mas;
  int rc;
  // activation codes / salient columns as 128B-swizzled tiles of [mp rows x 128 bytes]; rows past M and columns past
  // the valid width are zero-filled by TMA
  if ((rc = mq_make_map(&ma, a_codes, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, M, G * group_size, lda, mp))) return rc;
  if (Ks > 0) {
    const CUtensorMapDataType dt16 = dtype == MQ_F16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
    if ((rc = mq_make_map(&mas, a_sal, dt16, 2, M, Ks, ld_asal, mp))) return rc;
  } else {
    mas = ma;
  }
  DecodeArgs p{};
  p.w_nib = w_nib; p.w_sal = w_sal_dec; p.a_scales = a_scales; p.ld_as = ld_as; p.w_scales = w_scales; p.ld_ws = ld_ws;
  p.bias = bias; p.y = y; p.ldy = ldy; p.M = (int)M; p.N = (int)N; p.G = (int)G; p.g = group_size;
  p.units = (int)(G * group_size / 64);
  p.nkb = (p.units + 3) / 4;
  p.nsb = (int)((Ks + 63) / 64);
  p.tiles = (int)((N + 15) / 16);
  p.nib_tile_stride = (int64_t)p.nkb * kTileWBytes;
  p.sal_tile_stride = (int64_t)p.nsb * kTileWBytes;
  p.as_bytes = (int)((M + 3) / 4 * 4 * 4);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == MQ_F16) return dispatch_decode<__half>(ma, mas, p, st);
  return dispatch_decode<__nv_bfloat16>(ma, mas, p, st);
}

}  // extern "C"
