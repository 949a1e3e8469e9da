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

// compute warps = row tiles per round (template parameter NW): 16 for M <= 16 (more independent IMMA chains per
// scheduler: the legacy tensor path is latency bound at 2 warps per scheduler), 8 above (register budget); + 1 producer
constexpr int kBlockK = 256;                     // quantised columns per weight block (4 units of 64)
constexpr int kNibBytes = 2048;                  // nibbles of one row tile and block: 16 rows x 256 columns / 2
constexpr int kSalBytes = 2048;                  // salient block of one row tile: 16 rows x 64 columns x 2 B
constexpr int kMaxStages = 8;

// bytes of one (row tile, block) record in the nibble stream: the nibbles, then the weight scales of the groups that
// intersect the block -- max(1, 256 / g) slots of 16 rows, in the 16-bit compute dtype (exact: they are values of it)
__host__ __device__ constexpr int ws_slots(int g) { return g >= kBlockK ? 1 : kBlockK / g; }
__host__ __device__ constexpr int nib_record_bytes(int g) { return kNibBytes + ws_slots(g) * 32; }

struct DecodeArgs {
  const uint8_t* w_nib;       // [tiles][nkb] records of nib_record_bytes(g)
  const uint8_t* w_sal;       // [tiles][nsb][4 steps][32][16]
  const float* a_scales;      // [G][ld_as]
  int64_t ld_as;
  const void* bias;
  void* y;
  int64_t ldy;
  int M, N, G, g;
  int units;                  // 64-column units that hold quantised columns = G * g / 64
  int nkb;                    // weight blocks of 256 columns = ceil(units / 4)
  int nsb;                    // salient blocks of 64 columns = ceil(Ks / 64)
  int tiles;                  // ceil(N / 16)
  int rec;                    // nib_record_bytes(g)
  int64_t nib_tile_stride;    // bytes between row tiles in w_nib  = nkb * rec
  int64_t sal_tile_stride;    // bytes between row tiles in w_sal  = nsb * 2048
  int as_bytes;               // bytes of one activation-scale row copied per group (multiple of 16)
  // stage geometry (runtime): a stage holds `bps` consecutive blocks of every row tile of the round
  int bps, stages;
  int stage_bytes;            // multiple of 1024
  int w_off;                  // byte offset of the weight records inside a stage (act tiles come first)
  int w_pitch;                // bytes reserved per row tile inside a stage (>= bps * max(rec, 2048), multiple of 16)
  int as_off;                 // byte offset of the activation-scale slots inside a stage
  int tiles_per_cta;          // row tiles per CTA (the grid is ceil(tiles / tiles_per_cta))
  int gspan_sh;               // g >= 256: log2(g / 256) = log2(blocks per group); else 0
  int g_sh;                   // log2(g)
  int split;                  // CTAs per cluster sharing the K range of the same row tiles (1, 2, 4, 8)
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
__device__ __forceinline__ uint32_t lds_u16d(uint32_t addr) {
  uint32_t r;
  asm volatile("ld.shared.u16 %0, [%1];" : "=r"(r) : "r"(addr));
  return r;
}

// ------------------------------------------------------------------------------------------------ offline packers
// One thread per 4-byte word of the nibble stream.  Record (tile, block): [2048 B nibbles][slots x 16 scales].
// nibble word (unit uu, lane, pi): byte j = code[r][k + j] & 15 | code[r][k + 32 + j] << 4 with
// r = 16 * tile + lane / 4 + 8 * (pi & 1),  k = 256 * block + 64 * uu + 16 * (pi >> 1) + 4 * (lane & 3)
template <typename T>
__global__ void pack_decode_nib_kernel(const int8_t* __restrict__ codes, const float* __restrict__ w_scales, int64_t ld_ws,
                                       int64_t N, int64_t Kc, int64_t ldw, int G, int g, int nkb, uint32_t* __restrict__ out,
                                       int64_t words) {
  const int rec_words = nib_record_bytes(g) / 4;
  for (int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; w < words; w += (int64_t)gridDim.x * blockDim.x) {
    const int64_t rb = w / rec_words;                  // record index = tile * nkb + block
    const int wi = (int)(w - rb * rec_words);
    const int block = (int)(rb % nkb);
    const int64_t tile = rb / nkb;
    uint32_t v = 0;
    if (wi < kNibBytes / 4) {
      const int pi = wi & 3, lane = (wi >> 2) & 31, uu = wi >> 7;
      const int64_t r = tile * 16 + (lane >> 2) + 8 * (pi & 1);
      const int64_t k = (int64_t)block * kBlockK + uu * 64 + 16 * (pi >> 1) + 4 * (lane & 3);
      if (r < N) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const uint32_t lo = k + j < Kc ? (uint32_t)codes[r * ldw + k + j] & 15u : 0u;
          const uint32_t hi = k + 32 + j < Kc ? (uint32_t)codes[r * ldw + k + 32 + j] & 15u : 0u;
          v |= (lo | (hi << 4)) << (8 * j);
        }
      }
    } else {                                           // two consecutive rows' scales of one slot
      const int si = wi - kNibBytes / 4;               // 0 .. slots * 8 - 1
      const int slot = si >> 3, r2 = (si & 7) * 2;
      const int gi = g >= kBlockK ? (block * kBlockK) / g : block * (kBlockK / g) + slot;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int64_t r = tile * 16 + r2 + h;
        float s = 0.f;
        if (r < N && gi < G) s = w_scales[(int64_t)gi * ld_ws + r];
        const T t = Num<T>::from_f(s);
        v |= (uint32_t)(*reinterpret_cast<const uint16_t*>(&t)) << (16 * h);
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

template <typename T> __device__ __forceinline__ float bits16_to_f(uint32_t b);
template <> __device__ __forceinline__ float bits16_to_f<__half>(uint32_t b) { return __half2float(__ushort_as_half((unsigned short)b)); }
template <> __device__ __forceinline__ float bits16_to_f<__nv_bfloat16>(uint32_t b) { return __uint_as_float(b << 16); }

// GU = 64-column units per group inside one 256-column block: 1 (g = 64), 2 (g = 128), 4 (g >= 256: a group spans
// g / 256 whole blocks).  Compile-time so that the unit loop has no data-dependent branch except "is this a real group".
template <typename T, int NT, int GU, int NW>
__global__ void __launch_bounds__((NW + 1) * 32, 1)
decode_gemm_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_as,
                   const DecodeArgs p) {
  constexpr int Mp = NT * 8;
  constexpr uint32_t kActTile = Mp * 128;       // one 128B-swizzled activation tile [Mp rows x 128 B]
  extern __shared__ unsigned char smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  unsigned char* smem = smem_raw + (((raw_addr + 1023u) & ~1023u) - raw_addr);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + (size_t)p.stages * p.stage_bytes);
  uint64_t* empty_bar = full_bar + kMaxStages;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // split-K: the `split` CTAs of a cluster own the same row tiles and consecutive ranges of the K blocks (cut on group
  // boundaries); their fp32 partial tiles are summed in rank order through the leader's shared memory at the end
  const int split = p.split;
  const int crank = split > 1 ? (int)cluster_ctarank() : 0;
  const int t0 = (int)(blockIdx.x / (unsigned)split) * p.tiles_per_cta;
  const int t1 = min(p.tiles, t0 + p.tiles_per_cta);
  const int rounds = (t1 - t0 + NW - 1) / NW;             // 1 when split > 1
  const int gspan = 1 << p.gspan_sh;
  const int kb_lo = crank == 0 ? 0 : (int)((int64_t)p.nkb * crank / split) / gspan * gspan;
  const int kb_hi = crank == split - 1 ? p.nkb : (int)((int64_t)p.nkb * (crank + 1) / split) / gspan * gspan;
  const int sb_lo = (int)((int64_t)p.nsb * crank / split), sb_hi = (int)((int64_t)p.nsb * (crank + 1) / split);
  const int nis = (kb_hi - kb_lo + p.bps - 1) / p.bps;    // integer stages per round
  const int nss = (sb_hi - sb_lo + p.bps - 1) / p.bps;    // salient stages per round
  const uint32_t nst = (uint32_t)p.stages;

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], NW);
    }
    fence_mbar_init();
    tma_prefetch_desc(&map_a);
    tma_prefetch_desc(&map_as);
  }
  __syncthreads();
  // barriers and descriptors are set up; everything below touches memory the previous kernel of the stream may have
  // written (activation codes / scales) or may still be reading (y): wait for it here, not at the launch boundary
  griddep_launch();
  griddep_wait();

  if (warp == NW) {
    // ------------------------------------------------------------------ producer warp
    // Every copy of a stage is issued by a different lane: a bulk / tensor copy costs its issuing thread ~100 cycles,
    // and a stage has 10-30 of them (one thread issuing them all made the stage period 1.5 us, whatever its size).
    uint32_t it = 0;
    for (int r = 0; r < rounds; ++r) {
      const int tb = t0 + r * NW;
      const int nt = min(NW, t1 - tb);
      for (int si = 0; si < nis + nss; ++si, ++it) {
        const uint32_t s = it % nst;
        if (it >= nst && !mbar_wait(&empty_bar[s], ((it / nst) - 1u) & 1u, 0x820u + s)) return;
        unsigned char* st = smem + (size_t)s * p.stage_bytes;
        unsigned char* sw = st + p.w_off;
        if (si < nis) {
          const int b0 = kb_lo + si * p.bps, nb = min(p.bps, kb_hi - b0);
          const int k_lo = b0 * kBlockK, k_hi = k_lo + nb * kBlockK;
          // 128-column activation tiles with real columns; groups whose last column lies in this stage
          const int ntile = min(2 * nb, (p.units * 64 - k_lo + 127) / 128);
          const int g_first = (k_lo >> p.g_sh);                       // first group ending after k_lo
          const int g_last = min(p.G, k_hi >> p.g_sh);                // one past the last group ending at or before k_hi
          const int ngrp = max(0, g_last - g_first);
          if (lane == 0)
            mbar_expect_tx(&full_bar[s], (uint32_t)(ntile * (int)kActTile + ngrp * p.as_bytes + nt * nb * p.rec));
          __syncwarp();
          for (int op = lane; op < ntile + ngrp + nt; op += 32) {
            if (op < ntile) {
              tma_load_2d(st + op * kActTile, &map_a, k_lo + op * 128, 0, &full_bar[s]);
            } else if (op < ntile + ngrp) {
              const int gi = g_first + (op - ntile);
              const int end = (gi + 1) << p.g_sh;
              const int slot = ((end - k_lo + p.g - 1) >> p.g_sh) - 1;
              bulk_g2s(st + p.as_off + slot * Mp * 4, p.a_scales + (int64_t)gi * p.ld_as, (uint32_t)p.as_bytes, &full_bar[s]);
            } else {
              const int i = op - ntile - ngrp;
              bulk_g2s(sw + i * p.w_pitch, p.w_nib + (int64_t)(tb + i) * p.nib_tile_stride + (int64_t)b0 * p.rec,
                       (uint32_t)(nb * p.rec), &full_bar[s]);
            }
          }
        } else {
          const int b0 = sb_lo + (si - nis) * p.bps, nb = min(p.bps, sb_hi - b0);
          if (lane == 0) mbar_expect_tx(&full_bar[s], (uint32_t)(nb * (int)kActTile + nt * nb * kSalBytes));
          __syncwarp();
          for (int op = lane; op < nb + nt; op += 32) {
            if (op < nb) {
              tma_load_2d(st + op * kActTile, &map_as, (b0 + op) * 64, 0, &full_bar[s]);
            } else {
              const int i = op - nb;
              bulk_g2s(sw + i * p.w_pitch, p.w_sal + (int64_t)(tb + i) * p.sal_tile_stride + (int64_t)b0 * kSalBytes,
                       (uint32_t)(nb * kSalBytes), &full_bar[s]);
            }
          }
        }
      }
    }
    if (split > 1) {                 // the two cluster barriers of the partial-sum exchange (see the compute warps)
      cluster_sync_all();
      cluster_sync_all();
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
    const int tb = t0 + r * NW;
    const int nt = min(NW, t1 - tb);
    const bool active = warp < nt;
    const int n0 = (tb + warp) * 16;
    const int nA = n0 + gid, nB = nA + 8;
    float acc[NT][4];
    int ci[NT][4], cj[NT][4];              // two integer accumulator sets (even / odd k32 steps): independent IMMA chains
#pragma unroll
    for (int j = 0; j < NT; ++j)
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        acc[j][c] = 0.f;
        ci[j][c] = 0;
        cj[j][c] = 0;
      }

    for (int si = 0; si < nis; ++si, ++it) {
      const uint32_t s = it % nst;
      if (!mbar_wait(&full_bar[s], (it / nst) & 1u, 0x800u + s)) return;
      if (active) {
        const uint32_t st = smem_a + s * (uint32_t)p.stage_bytes;
        const int b0 = kb_lo + si * p.bps, nb = min(p.bps, kb_hi - b0);
        for (int bi = 0; bi < nb; ++bi) {
          const int b = b0 + bi;
          const uint32_t rec = st + (uint32_t)p.w_off + (uint32_t)(warp * p.w_pitch + bi * p.rec);
          const uint32_t sw = rec + (uint32_t)lane * 16u;
          // blocks are padded with zero nibbles up to 4 units, so every unit is computed (a padding unit adds 0)
#pragma unroll
          for (int uu = 0; uu < 4; ++uu) {
            const uint4 pk = lds_v4(sw + (uint32_t)uu * 512u);
            const uint32_t a_lo[4] = {(pk.x << 4) & 0xF0F0F0F0u, (pk.y << 4) & 0xF0F0F0F0u, (pk.z << 4) & 0xF0F0F0F0u,
                                      (pk.w << 4) & 0xF0F0F0F0u};
            const uint32_t a_hi[4] = {pk.x & 0xF0F0F0F0u, pk.y & 0xF0F0F0F0u, pk.z & 0xF0F0F0F0u, pk.w & 0xF0F0F0F0u};
            const uint32_t tile = st + (uint32_t)(bi * 2 + (uu >> 1)) * kActTile;
#pragma unroll
            for (int hs = 0; hs < 2; ++hs) {                       // the two k32 steps of the unit
              const int chunk = ((uu & 1) * 2 + hs) * 2;           // first 16-byte chunk of the step inside the 128-B row
              if (NT == 1) {
                uint32_t b0r, b1r;
                ldmatrix_x2(tile + (uint32_t)lr * 128u + (uint32_t)(((chunk + (lq & 1)) ^ lr) << 4), b0r, b1r);
                imma_16832(hs ? cj[0] : ci[0], hs ? a_hi : a_lo, b0r, b1r);
              } else {
#pragma unroll
                for (int j = 0; j < NT; j += 2) {
                  uint32_t b0r, b1r, b2r, b3r;
                  const int m = 8 * (j + (lq >> 1)) + lr;
                  ldmatrix_x4(tile + (uint32_t)m * 128u + (uint32_t)(((chunk + (lq & 1)) ^ lr) << 4), b0r, b1r, b2r, b3r);
                  imma_16832(hs ? cj[j] : ci[j], hs ? a_hi : a_lo, b0r, b1r);
                  imma_16832(hs ? cj[j + 1] : ci[j + 1], hs ? a_hi : a_lo, b2r, b3r);
                }
              }
            }
            if ((uu + 1) % GU == 0) {                              // compile-time: a group may end with this unit
              // group index / slots without divisions: GU < 4 -> 4 / GU groups per block; GU == 4 -> gspan blocks per group
              const int gi = GU == 4 ? b >> p.gspan_sh : b * (4 / GU) + uu / GU;
              const bool ends = GU < 4 || ((b + 1) & ((1 << p.gspan_sh) - 1)) == 0;
              if (ends && gi < p.G) {                              // warp-uniform; padding groups have no scales
                const int wslot = GU == 4 ? 0 : uu / GU;
                const uint32_t wsa = rec + (uint32_t)kNibBytes + (uint32_t)(wslot * 32 + gid * 2);
                const float wsA = bits16_to_f<T>(lds_u16d(wsa)) * 0.0625f;          // the 16 of the nibble trick
                const float wsB = bits16_to_f<T>(lds_u16d(wsa + 16u)) * 0.0625f;
                // activation-scale slot = index of the group among those ending inside this stage
                const int aslot = GU == 4 ? gi - (b0 >> p.gspan_sh) : bi * (4 / GU) + uu / GU;
                const uint32_t sa = st + (uint32_t)p.as_off + (uint32_t)(aslot * Mp * 4) + (uint32_t)(tq * 8);
#pragma unroll
                for (int j = 0; j < NT; ++j) {
                  const float2 as2 = lds_f2(sa + (uint32_t)(j * 32));
                  const float ax = __uint_as_float(__float_as_uint(as2.x) & 0xFFFFFFFCu);
                  const float ay = __uint_as_float(__float_as_uint(as2.y) & 0xFFFFFFFCu);
                  const float cx = -kMagicF * ax, cy = -kMagicF * ay;
#pragma unroll
                  for (int c = 0; c < 4; ++c) {
                    const float f = __uint_as_float((uint32_t)(ci[j][c] + cj[j][c]) + kMagicI);
                    const float t = (c & 1) ? fmaf(f, ay, cy) : fmaf(f, ax, cx);     // == 16 * sum * as exactly
                    acc[j][c] = fmaf(t, (c & 2) ? wsB : wsA, acc[j][c]);
                    ci[j][c] = 0;
                    cj[j][c] = 0;
                  }
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
    for (int si = 0; si < nss; ++si, ++it) {
      const uint32_t s = it % nst;
      if (!mbar_wait(&full_bar[s], (it / nst) & 1u, 0x810u + s)) return;
      if (active) {
        const uint32_t st = smem_a + s * (uint32_t)p.stage_bytes;
        const int b0 = sb_lo + si * p.bps, nb = min(p.bps, sb_hi - b0);
        for (int bi = 0; bi < nb; ++bi) {
          const uint32_t sw = st + (uint32_t)p.w_off + (uint32_t)(warp * p.w_pitch + bi * kSalBytes) + (uint32_t)lane * 16u;
          const uint32_t tile = st + (uint32_t)bi * kActTile;
#pragma unroll
          for (int k16 = 0; k16 < 4; ++k16) {
            const uint4 a = lds_v4(sw + (uint32_t)k16 * 512u);
            const int chunk = k16 * 2;
            if (NT == 1) {
              uint32_t b0r, b1r;
              ldmatrix_x2(tile + (uint32_t)lr * 128u + (uint32_t)(((chunk + (lq & 1)) ^ lr) << 4), b0r, b1r);
              hmma_16816<T>(cs[0], a, b0r, b1r);
            } else {
#pragma unroll
              for (int j = 0; j < NT; j += 2) {
                uint32_t b0r, b1r, b2r, b3r;
                const int m = 8 * (j + (lq >> 1)) + lr;
                ldmatrix_x4(tile + (uint32_t)m * 128u + (uint32_t)(((chunk + (lq & 1)) ^ lr) << 4), b0r, b1r, b2r, b3r);
                hmma_16816<T>(cs[j], a, b0r, b1r);
                hmma_16816<T>(cs[j + 1], a, b2r, b3r);
              }
            }
          }
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty_bar[s]);
    }

    if (split > 1) {
      // every CTA of the cluster is done with its ring -> ranks 1.. store their partial tiles into the leader's ring
      // memory ([rank - 1][value][thread], coalesced) -> the leader adds them in rank order: deterministic
      cluster_sync_all();
      const uint32_t tidc = (uint32_t)(warp * 32 + lane);
      if (crank != 0 && active) {
        const uint32_t base = mapa_u32(smem_a, 0u) + (uint32_t)((crank - 1) * NT * 4) * (uint32_t)(NW * 32 * 4) + tidc * 4u;
#pragma unroll
        for (int j = 0; j < NT; ++j)
#pragma unroll
          for (int c = 0; c < 4; ++c)
            asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(base + (uint32_t)((j * 4 + c) * NW * 32 * 4)),
                         "f"(acc[j][c] + cs[j][c]) : "memory");
      }
      cluster_sync_all();
      if (crank != 0) continue;
      if (active) {
        for (int rk = 1; rk < split; ++rk) {
          const uint32_t base = smem_a + (uint32_t)((rk - 1) * NT * 4) * (uint32_t)(NW * 32 * 4) + tidc * 4u;
#pragma unroll
          for (int j = 0; j < NT; ++j)
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              float v;
              asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(base + (uint32_t)((j * 4 + c) * NW * 32 * 4)) : "memory");
              acc[j][c] += v;
            }
        }
      }
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
int mp_of(int64_t M) { return M <= 8 ? 8 : M <= 16 ? 16 : M <= 32 ? 32 : 64; }

// stage geometry for `mp` padded tokens and `bps` blocks per stage; returns false if fewer than 2 stages fit
bool plan_stages(DecodeArgs& p, int mp, int bps, int nw) {
  const int act = bps * 2 * mp * 128;
  const int per_tile = bps * (p.rec > kSalBytes ? p.rec : kSalBytes);
  p.bps = bps;
  p.w_off = act;
  p.w_pitch = (per_tile + 15) / 16 * 16;
  p.as_off = p.w_off + nw * p.w_pitch;
  const int stage_k = bps * kBlockK;
  const int slots = p.g >= stage_k ? 1 : stage_k / p.g;
  p.stage_bytes = (p.as_off + slots * mp * 4 + 1023) / 1024 * 1024;
  const int budget = 227 * 1024 - 1024 - 2 * kMaxStages * 8;
  p.stages = budget / p.stage_bytes;
  if (p.stages > kMaxStages) p.stages = kMaxStages;
  return p.stages >= 2;
}

template <typename T, int NT, int GU>
int launch_decode_g(const CUtensorMap& ma, const CUtensorMap& mas, DecodeArgs p, cudaStream_t st) {
  constexpr int NW = NT <= 2 ? 16 : 8;
  int bps = mq_get_option("decode_bps", 0);
  if (bps <= 0) bps = NT == 8 ? 1 : 2;
  while (bps > 1 && !plan_stages(p, NT * 8, bps, NW)) bps /= 2;
  if (!plan_stages(p, NT * 8, bps, NW)) {
    mq_set_error("mq_decode_gemm: no stage geometry fits shared memory (g=%d)", p.g);
    return MQ_ERR_UNSUPPORTED;
  }
  const int smem = p.stages * p.stage_bytes + 2 * kMaxStages * 8 + 1024;
  auto kern = decode_gemm_kernel<T, NT, GU, NW>;
  if (int rc = mq_ensure_smem_attr(reinterpret_cast<const void*>(kern), 227 * 1024)) return rc;
  // Decomposition.  Many row tiles (tiles / NW >= ~0.7 * SMs): every CTA owns a contiguous range of tiles and the
  // whole K (rounds of NW tiles).  Few tiles (small N, long K): clusters of `split` CTAs share NW tiles and cut K, so
  // that all SMs stream and no CTA walks more than a handful of stages (a stage costs ~1 us of latency however small).
  const int sms = mq_sm_count();
  int tpc = mq_get_option("decode_tpc", 0), split = mq_get_option("decode_split", 0);
  if (tpc <= 0) {
    const int groups = (p.tiles + NW - 1) / NW;            // clusters if every cluster takes NW tiles
    if (groups * 10 >= sms * 7) {
      tpc = (p.tiles + sms - 1) / sms;
      if (split <= 0) split = 1;
    } else {
      tpc = p.tiles < NW ? p.tiles : NW;
      if (split <= 0) {
        split = 1;
        while (split < 8 && groups * split * 2 <= sms && p.nkb / (split * 2) >= 2 * (1 << p.gspan_sh)) split *= 2;
      }
    }
  }
  if (split <= 0) split = 1;
  if (split > 1 && tpc > NW) tpc = NW;                     // one round per CTA when K is split
  // the partial tiles of ranks 1.. are exchanged through the leader's ring memory: (split - 1) x NT*4 x NW*32 floats
  while (split > 1 && (size_t)(split - 1) * NT * 4 * NW * 32 * 4 > (size_t)p.stages * p.stage_bytes) split /= 2;
  p.tiles_per_cta = tpc;
  p.split = split;
  const int grid = (p.tiles + tpc - 1) / tpc * split;
  {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3((NW + 1) * 32);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    int na = 0;
    if (split > 1) {
      attr[na].id = cudaLaunchAttributeClusterDimension;
      attr[na].val.clusterDim.x = (unsigned)split;
      attr[na].val.clusterDim.y = 1;
      attr[na].val.clusterDim.z = 1;
      ++na;
    }
    if (mq_get_option("pdl", 1) == 1) {   // overlap this kernel's launch + prologue with the tail of its predecessor
      attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
      attr[na].val.programmaticStreamSerializationAllowed = 1;
      ++na;
    }
    cfg.attrs = attr;
    cfg.numAttrs = na;
    MQ_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kern, ma, mas, p));
  }
  MQ_CHECK_CUDA(cudaGetLastError());
  return MQ_OK;
}

template <typename T, int NT>
int launch_decode(const CUtensorMap& ma, const CUtensorMap& mas, DecodeArgs p, cudaStream_t st) {
  p.gspan_sh = 0;
  p.g_sh = 0;
  for (int v = p.g; v > 1; v >>= 1) ++p.g_sh;
  for (int v = p.g / kBlockK; v > 1; v >>= 1) ++p.gspan_sh;
  if (p.g == 64) return launch_decode_g<T, NT, 1>(ma, mas, p, st);
  if (p.g == 128) return launch_decode_g<T, NT, 2>(ma, mas, p, st);
  return launch_decode_g<T, NT, 4>(ma, mas, p, st);
}

template <typename T>
int dispatch_decode(const CUtensorMap& ma, const CUtensorMap& mas, const DecodeArgs& p, cudaStream_t st) {
  if (p.M <= 8) return launch_decode<T, 1>(ma, mas, p, st);
  if (p.M <= 16) return launch_decode<T, 2>(ma, mas, p, st);
  if (p.M <= 32) return launch_decode<T, 4>(ma, mas, p, st);
  return launch_decode<T, 8>(ma, mas, p, st);
}

bool pow2(int v) { return v > 0 && (v & (v - 1)) == 0; }

}  // namespace

extern "C" {

int mq_decode_weight_bytes(int64_t N, int64_t G, int group_size, int64_t Ks, size_t* nib_bytes, size_t* sal_bytes) {
  MQ_CHECK_ARG(N >= 1 && G >= 1 && group_size >= 64 && pow2(group_size) && Ks >= 0,
               "mq_decode_weight_bytes: needs N >= 1, G >= 1, a power-of-two group_size >= 64, Ks >= 0");
  const int64_t tiles = (N + 15) / 16;
  const int64_t units = G * group_size / 64, nkb = (units + 3) / 4, nsb = (Ks + 63) / 64;
  if (nib_bytes) *nib_bytes = (size_t)(tiles * nkb * nib_record_bytes(group_size));
  if (sal_bytes) *sal_bytes = (size_t)(tiles * nsb * kSalBytes);
  return MQ_OK;
}

int mq_pack_decode(const int8_t* w_codes, int64_t ldw, const float* w_scales, int64_t ld_ws, const void* w_sal,
                   int64_t ld_wsal, int64_t N, int64_t G, int group_size, int64_t Ks, int dtype, uint8_t* out_nib,
                   uint8_t* out_sal, void* stream) {
  MQ_CHECK_ARG(w_codes && w_scales && out_nib, "mq_pack_decode: null pointer");
  MQ_CHECK_ARG(N >= 1 && G >= 1 && group_size >= 64 && pow2(group_size),
               "mq_pack_decode: needs N >= 1, G >= 1 and a power-of-two group size >= 64");
  MQ_CHECK_ARG(ldw >= G * group_size && ld_ws >= N, "mq_pack_decode: ldw=%lld < G*g or ld_ws < N", (long long)ldw);
  MQ_CHECK_ARG(Ks == 0 || (w_sal && out_sal && ld_wsal >= Ks), "mq_pack_decode: bad salient arguments");
  MQ_CHECK_ARG(dtype == MQ_F16 || dtype == MQ_BF16, "mq_pack_decode: dtype must be MQ_F16 or MQ_BF16");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int64_t tiles = (N + 15) / 16;
  const int64_t units = G * group_size / 64, nkb = (units + 3) / 4, nsb = (Ks + 63) / 64;
  const int64_t cap = (int64_t)mq_sm_count() * 16;
  {
    const int64_t words = tiles * nkb * (nib_record_bytes(group_size) / 4);
    int64_t blocks = (words + 255) / 256;
    if (blocks > cap) blocks = cap;
    if (dtype == MQ_F16)
      pack_decode_nib_kernel<__half><<<(unsigned)blocks, 256, 0, st>>>(w_codes, w_scales, ld_ws, N, G * group_size, ldw, (int)G,
                                                                      group_size, (int)nkb, reinterpret_cast<uint32_t*>(out_nib), words);
    else
      pack_decode_nib_kernel<__nv_bfloat16><<<(unsigned)blocks, 256, 0, st>>>(w_codes, w_scales, ld_ws, N, G * group_size, ldw,
                                                                              (int)G, group_size, (int)nkb,
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
                   int64_t ld_asal, const uint8_t* w_nib, const uint8_t* w_sal_dec, const void* bias, int64_t M,
                   int64_t N, int64_t G, int group_size, int64_t Ks, int dtype, void* y, int64_t ldy, void* stream) {
  MQ_CHECK_ARG(y && M >= 0 && M <= 64 && N >= 1 && ldy >= N, "mq_decode_gemm: needs 0 <= M <= 64, N >= 1, ldy >= N (M=%lld)",
               (long long)M);
  MQ_CHECK_ARG(dtype == MQ_F16 || dtype == MQ_BF16, "mq_decode_gemm: dtype must be MQ_F16 or MQ_BF16");
  MQ_CHECK_ARG(G >= 1 && group_size >= 64 && pow2(group_size) && group_size <= 4096,
               "mq_decode_gemm: needs G >= 1 and a power-of-two group size in [64, 4096] (got G=%lld g=%d)", (long long)G,
               group_size);
  MQ_CHECK_ARG(a_codes && a_scales && w_nib, "mq_decode_gemm: null code / scale pointer");
  MQ_CHECK_ARG(lda >= G * group_size && lda % 16 == 0 && reinterpret_cast<uintptr_t>(a_codes) % 16 == 0,
               "mq_decode_gemm: lda=%lld must be a multiple of 16 and >= G*g", (long long)lda);
  MQ_CHECK_ARG(ld_as >= M && ld_as % 4 == 0 && reinterpret_cast<uintptr_t>(a_scales) % 16 == 0,
               "mq_decode_gemm: scales need ld_as >= M, ld_as %% 4 == 0 and a 16-byte aligned pointer");
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
  p.w_nib = w_nib; p.w_sal = w_sal_dec; p.a_scales = a_scales; p.ld_as = ld_as;
  p.bias = bias; p.y = y; p.ldy = ldy; p.M = (int)M; p.N = (int)N; p.G = (int)G; p.g = group_size;
  p.units = (int)(G * group_size / 64);
  p.nkb = (p.units + 3) / 4;
  p.nsb = (int)((Ks + 63) / 64);
  p.tiles = (int)((N + 15) / 16);
  p.rec = nib_record_bytes(group_size);
  p.nib_tile_stride = (int64_t)p.nkb * p.rec;
  p.sal_tile_stride = (int64_t)p.nsb * kSalBytes;
  p.as_bytes = (int)((M + 3) / 4 * 4 * 4);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == MQ_F16) return dispatch_decode<__half>(ma, mas, p, st);
  return dispatch_decode<__nv_bfloat16>(ma, mas, p, st);
}

}  // extern "C"
