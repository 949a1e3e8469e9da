// K2: fused mixed-precision GEMM for the W4A4 MixedQuant linear layer (sm_100a: TMA + tcgen05 + TMEM).
//
//   y[m, n] = sum_g a_scale[g, m] * w_scale[g, n] * sum_{k in group g} a_code[m, k] * w_code[n, k]     (kind::i8)
//           + sum_j a_sal[m, j] * w_sal[n, j]                                                           (kind::f16)
//           + bias[n]
//
// Replaces  torch.functional.F.linear(q_x, self.weight, self.bias)   /root/reference/smoothquant/fake_quant.py:306
// where q_x / self.weight hold dequantised code*scale values in the non-salient columns and FP16 originals in the
// salient ones (fake_quant.py:291-301, :347-365).
//
// Persistent kernel: gridDim.x = min(#tiles, #SMs) CTAs of 384 threads, each looping over 128 x BN output tiles
// (tile = blockIdx.x + i * gridDim.x; n-block fastest so that concurrently running CTAs share A rows in L2).
//   warp 0      TMA producer: 128-byte-wide K blocks of A (128 rows) and W (BN rows) into a kStages ring
//               (SWIZZLE_128B, one full/empty mbarrier pair per stage)
//   warp 1      MMA issuer (one thread): per accumulation unit (a K group of codes, or the salient slice) it issues
//               tcgen05.mma kind::i8 / kind::f16 (M=128, N=BN, K=32 bytes) into one of two TMEM accumulator stages,
//               commits the stage to the epilogue and moves on to the other stage
//   warp 2      TMEM allocation / deallocation (2*BN columns);  warp 3 idle
//   warps 4-11  epilogue (two warpgroups, registers raised with setmaxnreg): thread t of a warp owns TMEM lane
//               (= tile row) 32*(warp%4)+t and BN/2 columns.  Per unit it reads the int32 partial sums with
//               tcgen05.ld, converts them with the 2^23 magic-number trick (|sum| <= g*49 << 2^22, exact), scales by
//               a_scale[g,m]*w_scale[g,n] with packed f32x2 math and accumulates in fp32 registers; it releases the
//               TMEM stage as soon as its loads have completed, so the MMAs of the next unit (and of the next tile,
//               while this tile's output is being rounded and stored) overlap the rescale.
#include <cuda.h>

#include "common.cuh"
#include "tma_host.cuh"
#include "ptx.cuh"

namespace {

using namespace mqptx;

constexpr int BM = 128;
constexpr int kKBlockBytes = 128;  // one swizzle atom wide
// Warp roles.  The epilogue warps come FIRST: the SM sub-partition schedulers favour the highest warp id among the
// eligible warps, and the two issue warps (TMA producer, MMA issuer) have little to issue but are on the critical
// path -- with low ids they were starved by the epilogue warps sharing their sub-partition.
#ifndef MQ_GEMM_EPI_WARPS
#define MQ_GEMM_EPI_WARPS 8
#endif
constexpr int kEpiWarps = MQ_GEMM_EPI_WARPS;  // 8: thread = row x BN/2 columns (232 regs); 16: row x BN/4 (112 regs)
constexpr int kColSplit = kEpiWarps / 4;      // column slices of a tile (each TMEM lane quarter has kColSplit warps)
constexpr int kProducerWarp = kEpiWarps;
constexpr int kMmaWarp = kEpiWarps + 1;
constexpr int kAllocWarp = kEpiWarps + 2;
// + one idle warp: setmaxnreg is a warpgroup-wide instruction, so the issue warps must form a complete group of 4
// (an 11-warp CTA hung on the B200)
constexpr int kThreads = (kEpiWarps + 4) * 32;
#ifndef MQ_GEMM_EPI_REGS
// 8 epilogue warps: 384 threads launch with 168 registers = 64 512; 8 x 32 x 240 + 4 x 32 x 24 = 64 512 as well, so the
// epilogue gets 240 registers when the four issue warps shrink to the minimum of 24 (they spill nothing at 24).  240
// is what lets the TMEM drain run as double-buffered x32 loads (64 registers of in-flight sums) without spills:
// 0.588 -> 0.576 ms at 32768 x 4096 x 4096 (r02; single-buffered x32 / x64 at 232 registers spill and lose).
// 16 epilogue warps: ptxas launches 608 threads with 96 registers -> 104 after the issue warps shrink to 40.
#define MQ_GEMM_EPI_REGS (MQ_GEMM_EPI_WARPS == 8 ? 240 : 104)
#endif
#ifndef MQ_GEMM_ISSUE_REGS
#define MQ_GEMM_ISSUE_REGS (MQ_GEMM_EPI_WARPS == 8 ? 24 : 40)
#endif
#ifndef MQ_GEMM_NBUF
#define MQ_GEMM_NBUF 2   // TMEM drain chunks in flight per epilogue thread: 2 = load chunk c+1 while chunk c is rescaled, 1 = none
#endif
constexpr int kEpiRegs = MQ_GEMM_EPI_REGS;
constexpr int kNBuf = MQ_GEMM_NBUF;
static_assert(kEpiWarps == 8 || kEpiWarps == 16, "8 or 16 epilogue warps");

struct GemmArgs {
  const float* a_scales;
  int64_t ld_as;
  const float* w_scales;
  int64_t ld_ws;
  const void* bias;
  void* y;
  int64_t ldy;
  int M, N, G, g;
  int kb_int;  // 128-column K blocks of codes holding the G*g quantised columns
  int kb_sal;  // 64-column K blocks of the salient slice
  int tiles_m, tiles_n;
  int profile;
  int debug;  // bit 1: the MMA warp issues no MMAs (timing experiments only)
};

// kPreload: the epilogue refills every drained TMEM accumulator stage with the bit pattern of 1.5 * 2^23 and the
// integer MMAs always accumulate, so that the drained int32 words ARE the floats kMagic + sum: the per-element
// integer add of the magic-number conversion disappears from the epilogue (1/3 of its instructions).
#ifndef MQ_GEMM_PRELOAD
#define MQ_GEMM_PRELOAD 1
#endif
constexpr bool kPreload = MQ_GEMM_PRELOAD != 0;
// kMagicMma: the magic-number pattern is not stored by the epilogue threads (16 STTM + ~67 constant re-materialising
// moves per unit and warp) but produced by the tensor core: before the integer MMAs of a unit the issuer runs ONE
// kind::f16 MMA (K = 16, accumulate off) whose two operands are the same constant shared-memory atom -- every 16-byte
// chunk holds six 1024.0 and two 0.0 halves, so every accumulator becomes 12 * 2^20 = 1.5 * 2^23 exactly (fp32), i.e.
// the bit pattern 0x4B400000 the integer MMAs then accumulate onto.  Costs a quarter of a unit's tensor time at
// g = 128 and removes a fifth of the epilogue's instructions.
#ifndef MQ_GEMM_MAGIC_MMA
#define MQ_GEMM_MAGIC_MMA 1
#endif
// Chosen per launch (template parameter kMagicMma): units of up to 256 columns use the constant MMA, longer units keep
// the epilogue refill (at g = 1024 the kernel is operand-stream bound and the extra MMA per unit costs 4-6 %).
// MQ_GEMM_MAGIC_MMA / option gemm_magic_mma: 1 = by group size, 2 = always, 3 = never.
constexpr int kMagicMmaMode = MQ_GEMM_MAGIC_MMA;
constexpr int kMagicMmaMaxGroup = 256;
#ifndef MQ_GEMM_UNIFORM
#define MQ_GEMM_UNIFORM 2   // 0: plain thread-derived addresses (experiments), 1: uniform warp index, 2: + opaque uniform addresses
#endif
#ifndef MQ_GEMM_CH
#define MQ_GEMM_CH (MQ_GEMM_EPI_WARPS == 8 ? 32 : 16)
#endif
#ifndef MQ_GEMM_SCALE_ON_FULL
#define MQ_GEMM_SCALE_ON_FULL 1
#endif
constexpr bool kScaleOnFull = MQ_GEMM_SCALE_ON_FULL != 0;
__device__ __forceinline__ void tmem_ld_chunk(uint32_t taddr, uint32_t (&v)[16]) { tmem_ld_32x16(taddr, v); }
__device__ __forceinline__ void tmem_ld_chunk(uint32_t taddr, uint32_t (&v)[32]) { tmem_ld_32x32(taddr, v); }
__device__ __forceinline__ void tmem_ld_chunk(uint32_t taddr, uint32_t (&v)[64]) {   // two x32 loads, one wait
  tmem_ld_32x32(taddr, *reinterpret_cast<uint32_t (*)[32]>(&v[0]));
  tmem_ld_32x32(taddr + 32u, *reinterpret_cast<uint32_t (*)[32]>(&v[32]));
}
#ifndef MQ_GEMM_PAIR_DEFAULT
#define MQ_GEMM_PAIR_DEFAULT 3  // 1: CTA-pair kernel whenever possible;  2: single-CTA kernel;  3: by tile count
#endif
constexpr int kScaleSlots = 8;  // ring of per-unit scale vectors (ws[BN] | as[BM], fp32) staged in shared memory by TMA

// optional per-CTA wait-cycle profile (option gemm_profile): 8 x u64 per CTA
//   [0] kernel cycles  [1] kernel ns  [2] producer waiting for a free smem stage  [3] MMA warp waiting for operands
//   [4] MMA warp waiting for a free TMEM stage  [5] epilogue warp 4 waiting for an accumulator  [6] .. for scales
//   [7] epilogue warp 4 busy cycles (drain + rescale + store)
constexpr int kProfSlots = 8;
constexpr int kProfCtas = 256;
__device__ unsigned long long g_gemm_prof[kProfCtas * kProfSlots];

// kPair: CTA-pair kernel (cta_group::2, a 256 x BN tile per pair): each CTA stages its own 128 rows of A and HALF of
// the W tile per K block, so that the L2 -> shared-memory operand stream per SM drops by a third.
template <int BN, int kStages, bool kPair = false>
struct SmemLayout {
  static constexpr int kABytes = BM * kKBlockBytes;
  static constexpr int kBRows = kPair ? BN / 2 : BN;  // W rows staged by this CTA
  static constexpr int kBBytes = kBRows * kKBlockBytes;
  static constexpr int kStageBytes = kABytes + kBBytes;
  // per scale slot: ws[BN] | as[BM]   (fp32)
  static constexpr int kSlotFloats = BN + BM;
  static constexpr int kConstOffset = kStages * kStageBytes;  // 1024-byte constant atom (kMagicMma), 1024-aligned
  static constexpr int kScaleOffset = kConstOffset + 1024;
  static constexpr int kBarOffset = kScaleOffset + kScaleSlots * kSlotFloats * 4;
  // full[kStages], empty[kStages], tmem_full[2], tmem_empty[2], sfull[kScaleSlots], sempty[kScaleSlots], tmem base
  static constexpr int kBarBytes = (2 * kStages + 4 + 2 * kScaleSlots) * 8 + 16;
  // output staging: per epilogue warp 32 rows x 32 columns of 16-bit values, row pitch 80 B (64 + 16: conflict-free
  // 16-byte accesses both by row and by 8-row x 4-chunk read-back)
  static constexpr int kOutPitch = 80;
  static constexpr int kOutWarpBytes = 32 * kOutPitch;
  static constexpr int kOutOffset = (kBarOffset + kBarBytes + 15) / 16 * 16;
  static constexpr int kTotal = kOutOffset + kEpiWarps * kOutWarpBytes + 1024;  // + slack to align the base to 1024 B
  static_assert(kTotal <= 232448, "shared memory budget");
};

template <int kRegs>
__device__ __forceinline__ void reg_dealloc() {
  asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRegs));
}
template <int kRegs>
__device__ __forceinline__ void reg_alloc() {
  asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(kRegs));
}

// int32 -> fp32 by bit pattern: for |v| < 2^22, bits(kMagic) + v are the bits of the float kMagic + v.  The rescale
// then needs no subtraction of its own: t = fma(kMagic + v, as, -kMagic * as) is ONE rounding of the exact value
// v * as, and v * as (<= 13 + 11 bits for fp16-valued scales) is representable, so t == v * as exactly; `as` and
// -kMagic * as are per-thread scalars of the unit, so the instruction reads one new register.  -kMagic * as =
// -3 * 2^22 * as is exact whenever `as` has at most 22 significant bits (fp16 / bf16-valued scales have 11 / 8); the
// two lowest mantissa bits of wider activation scales are cleared so that the identity holds for any input.
constexpr float kMagic = 12582912.0f;     // 1.5 * 2^23
constexpr uint32_t kMagicBits = 0x4B400000u;

// bounded wait on the mbarrier at shared address `bar`; returns -1 if it expired, else the cycles it took when
// profiling (0 otherwise)
template <bool kProf>
__device__ __forceinline__ long long wait_timed(bool prof, uint32_t bar, uint32_t parity, uint32_t code) {
  if (!kProf || !prof) return mbar_wait_a(bar, parity, code) ? 0 : -1;
  const long long t = clock64();  // try_wait may block for a while by itself: time the whole wait
  const bool ok = mbar_wait_a(bar, parity, code);
  return ok ? clock64() - t : -1;
}

template <typename OutT, int BN, int kStages, bool kProf, bool kPair, bool kMagicMma>
__global__ void __launch_bounds__(kThreads, 1)
mixed_gemm_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_w,
                  const __grid_constant__ CUtensorMap map_as, const __grid_constant__ CUtensorMap map_ws,
                  const __grid_constant__ CUtensorMap map_sa, const __grid_constant__ CUtensorMap map_sw,
                  const GemmArgs p) {
  using L = SmemLayout<BN, kStages, kPair>;
  constexpr bool kRefill = kPreload && !kMagicMma;  // the epilogue stores the pattern itself
  constexpr bool kBiased = kPreload || kMagicMma;   // drained words are the floats kMagic + sum
  constexpr int CW = BN / kColSplit;  // columns per epilogue thread
  // work distribution: a CTA (or a CTA pair) loops over 128 (256) x BN tiles; `wid` / `wstride` index the workers
  const uint32_t cta_rank = kPair ? cluster_ctarank() : 0u;
  const bool leader = cta_rank == 0u;
  const int wid = kPair ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int wstride = kPair ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  constexpr int TM = kPair ? 2 * BM : BM;  // tile rows per worker
  constexpr int kTmemCols = 2 * BN;
  static_assert(CW % 16 == 0, "bad epilogue split");
  static_assert(kTmemCols == 256 || kTmemCols == 512 || kTmemCols == 128, "TMEM columns must be a power of two");

  extern __shared__ unsigned char smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  unsigned char* smem = smem_raw + (((raw_addr + 1023u) & ~1023u) - raw_addr);

  float* sscale = reinterpret_cast<float*>(smem + L::kScaleOffset);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::kBarOffset);
  uint64_t* empty_bar = full_bar + kStages;
  uint64_t* tfull_bar = empty_bar + kStages;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint64_t* sfull_bar = tempty_bar + 2;
  uint64_t* sempty_bar = sfull_bar + kScaleSlots;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sempty_bar + kScaleSlots);

#if MQ_GEMM_UNIFORM
  // warp index through a lane-0 shuffle: ptxas then knows the value (and every TMEM / barrier address derived from
  // it) is warp-uniform and keeps it in uniform registers instead of one R2UR per tcgen05 instruction
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
#else
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#endif
  const bool prof = kProf && p.profile != 0 && blockIdx.x < kProfCtas;
  const long long k_t0 = kProf ? clock64() : 0;
  const unsigned long long k_ns0 = kProf ? globaltimer_ns() : 0ull;
  long long w0 = 0, w1 = 0, dt;  // wait-cycle accumulators of this thread's role
  const int kb_total = p.kb_int + p.kb_sal;
  const int num_tiles = p.tiles_m * p.tiles_n;

  if (warp == kProducerWarp && lane == 0) {
    tma_prefetch_desc(&map_a);
    tma_prefetch_desc(&map_w);
    tma_prefetch_desc(&map_as);
    tma_prefetch_desc(&map_ws);
    tma_prefetch_desc(&map_sa);
    tma_prefetch_desc(&map_sw);
  }
  if (warp == kMmaWarp && lane == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int t = 0; t < 2; ++t) {
      mbar_init(&tfull_bar[t], 1);
      mbar_init(&tempty_bar[t], kPair ? 2 * kEpiWarps : kEpiWarps);  // pair: the epilogues of both CTAs
    }
    for (int s = 0; s < kScaleSlots; ++s) {
      mbar_init(&sfull_bar[s], 1);
      mbar_init(&sempty_bar[s], kEpiWarps);
    }
    fence_mbar_init();
  }
  if (kMagicMma && warp == kAllocWarp + 1) {  // constant atom: per 16-byte chunk six halves of 1024.0 and two of 0.0
    const uint4 chunk = make_uint4(0x64006400u, 0x64006400u, 0x64006400u, 0u);
    uint4* atom = reinterpret_cast<uint4*>(smem + L::kConstOffset);
    atom[lane] = chunk;
    atom[lane + 32] = chunk;
    fence_proxy_async_smem();  // generic-proxy stores -> visible to the tensor core's async-proxy reads
  }
  if (warp == kAllocWarp) {
    if (kPair) tmem_alloc_pair<kTmemCols>(tmem_slot);
    else tmem_alloc<kTmemCols>(tmem_slot);
  }
  tc_fence_before();
  if (kPair) cluster_sync_all();  // barriers of both CTAs initialised before any remote arrive / multicast commit
  else __syncthreads();
  tc_fence_after();
#if MQ_GEMM_UNIFORM
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *reinterpret_cast<volatile uint32_t*>(tmem_slot), 0);
#else
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(tmem_slot);
#endif

  if (warp >= kEpiWarps) {
    reg_dealloc<MQ_GEMM_ISSUE_REGS>();
    // Producer and MMA issuer run as whole warps with warp-uniform control flow; the asynchronous instructions are
    // issued by one elected lane.  (A `lane == 0` branch around the whole loop makes ptxas treat every operand of
    // UTMALDG / UTCIMMA as potentially divergent and wrap each issue in a vote / R2UR waterfall.)
    if (warp == kProducerWarp) {
      // ------------------------------------------------------------------ TMA producer
      uint32_t kbc = 0, sc = 0;  // K blocks / scale slots issued so far (ring positions)
      const uint32_t full_leader = kPair ? mapa_u32(smem_u32(full_bar), 0u) : 0u;
      for (int tile = wid; tile < num_tiles; tile += wstride) {
        const int m0 = (tile / p.tiles_n) * TM + (int)cta_rank * BM, n0 = (tile % p.tiles_n) * BN;
        const int nw = n0 + (int)cta_rank * L::kBRows;  // first W row this CTA stages
        int u_next = 0;  // next integer unit whose scale vectors have not been requested yet
        for (int kb = 0; kb < kb_total; ++kb, ++kbc) {
          const uint32_t s = kbc % kStages, ph = (kbc / kStages) & 1u;
          if ((dt = wait_timed<kProf>(prof, smem_u32(&empty_bar[s]), ph ^ 1u, 0x200u + s)) < 0) goto teardown;
          if (kProf) w0 += dt;
          // Scale vectors of every unit that starts inside this K block (ws[u, n0 .. n0+BN) and as[u, m0 .. m0+BM))
          // complete on the SAME full barrier as the block's operands (MQ_GEMM_SCALE_ON_FULL): the MMAs of the unit
          // cannot start before they have landed, so the accumulator-complete barrier the epilogue waits for anyway
          // also covers the scales and the epilogue needs no second wait per unit.
          const uint32_t sc0 = sc;
          while (u_next < p.G && u_next * p.g < (kb + 1) * 128) {
            const uint32_t slot = sc % kScaleSlots;
            if (!mbar_wait(&sempty_bar[slot], ((sc / kScaleSlots) & 1u) ^ 1u, 0x600u + slot)) goto teardown;
            ++u_next;
            ++sc;
          }
          const uint32_t n_new = sc - sc0;
          const int u0 = u_next - (int)n_new;
          unsigned char* sa = smem + (size_t)s * L::kStageBytes;
          unsigned char* sb = sa + L::kABytes;
          if (elect_one()) {
            const uint32_t scale_bytes = kScaleOnFull ? n_new * (uint32_t)(L::kSlotFloats * 4) : 0u;
            if (kPair) {  // both CTAs' bytes complete on the leader's barrier, which the leader arms for both
              if (leader) mbar_expect_tx(&full_bar[s], 2u * ((uint32_t)L::kStageBytes + scale_bytes));
              const uint32_t fb = full_leader + s * 8u;
              for (uint32_t q = 0; q < n_new; ++q) {
                float* sl = sscale + ((sc0 + q) % kScaleSlots) * L::kSlotFloats;
                if (kScaleOnFull) {
                  tma_load_2d_pair(sl, &map_sw, n0, u0 + (int)q, fb);
                  tma_load_2d_pair(sl + BN, &map_sa, m0, u0 + (int)q, fb);
                } else {
                  uint64_t* sf = &sfull_bar[(sc0 + q) % kScaleSlots];
                  mbar_expect_tx(sf, (uint32_t)(L::kSlotFloats * 4));
                  tma_load_2d(sl, &map_sw, n0, u0 + (int)q, sf);
                  tma_load_2d(sl + BN, &map_sa, m0, u0 + (int)q, sf);
                }
              }
              if (kb < p.kb_int) {
                tma_load_2d_pair(sa, &map_a, kb * 128, m0, fb);
                tma_load_2d_pair(sb, &map_w, kb * 128, nw, fb);
              } else {
                const int ks = (kb - p.kb_int) * 64;
                tma_load_2d_pair(sa, &map_as, ks, m0, fb);
                tma_load_2d_pair(sb, &map_ws, ks, nw, fb);
              }
            } else {
              mbar_expect_tx(&full_bar[s], (uint32_t)L::kStageBytes + scale_bytes);
              for (uint32_t q = 0; q < n_new; ++q) {
                float* sl = sscale + ((sc0 + q) % kScaleSlots) * L::kSlotFloats;
                uint64_t* sf = kScaleOnFull ? &full_bar[s] : &sfull_bar[(sc0 + q) % kScaleSlots];
                if (!kScaleOnFull) mbar_expect_tx(sf, (uint32_t)(L::kSlotFloats * 4));
                tma_load_2d(sl, &map_sw, n0, u0 + (int)q, sf);
                tma_load_2d(sl + BN, &map_sa, m0, u0 + (int)q, sf);
              }
              if (kb < p.kb_int) {
                tma_load_2d(sa, &map_a, kb * 128, m0, &full_bar[s]);
                tma_load_2d(sb, &map_w, kb * 128, n0, &full_bar[s]);
              } else {
                const int ks = (kb - p.kb_int) * 64;
                tma_load_2d(sa, &map_as, ks, m0, &full_bar[s]);
                tma_load_2d(sb, &map_ws, ks, n0, &full_bar[s]);
              }
            }
          }
        }
      }
    } else if (warp == kMmaWarp && leader) {
      // ------------------------------------------------------------------ MMA issuer
      // Whole warp, warp-uniform control flow, one elected lane issues.  Per K block the elected lane runs ONE asm
      // block (4 MMAs) plus the commits; descriptors are 32-bit low words advanced by constants.
      constexpr uint32_t idesc_i8 = umma_idesc(2u, 1u, TM, BN);
      const uint32_t idesc_f16 = umma_idesc(1u, Num<OutT>::kDtype == MQ_BF16 ? 1u : 0u, TM, BN);
      // pair: the MMAs span both SMs and every completion is multicast to the same barrier of both CTAs
      auto commit = [](uint32_t bar) {
        if (kPair) tc_commit_pair(bar);
        else tc_commit_a(bar);
      };
#ifdef MQ_DEBUG_NOMMA
      const bool issue = false;
#else
      const bool issue = !(kProf && (p.debug & 2));
#endif
      // without preload the first use of a TMEM stage finds it free (wait on the inverted parity); with preload the
      // epilogue's initial fill is completion #0 of the stage's empty barrier
      constexpr uint32_t kEmptyFlip = kRefill ? 0u : 1u;
      const uint32_t idesc_fill = umma_idesc(1u, 0u, TM, BN);  // f16 x f16 -> f32
      const uint32_t const_lo = umma_desc_lo(smem_u32(smem) + (uint32_t)L::kConstOffset);
      auto fill = [&](uint32_t taddr) {
        if (kPair) mma_const_fill_pair(taddr, const_lo, idesc_fill);
        else mma_const_fill(taddr, const_lo, idesc_fill);
      };
      const uint32_t smem_a = smem_u32(smem);
      const uint32_t full_a = smem_u32(full_bar), empty_a = smem_u32(empty_bar);
      const uint32_t tfull_a = smem_u32(tfull_bar), tempty_a = smem_u32(tempty_bar);
      const int bpu = p.g >= 128 ? p.g / 128 : 1;   // K blocks per unit       (g >= 128)
      const int upb = p.g >= 128 ? 1 : 128 / p.g;   // units per K block       (g  < 128)
      const int spu = p.g >= 128 ? 4 : p.g / 32;    // MMAs per unit inside one K block
      uint32_t kbc = 0, uc = 0;  // K blocks consumed, units issued (ring positions)
      for (int tile = wid; tile < num_tiles; tile += wstride) {
        if (p.g >= 128) {
          for (int u = 0; u < p.G; ++u, ++uc) {
            const uint32_t t = uc & 1u;
            if ((dt = wait_timed<kProf>(prof, tempty_a + t * 8u, ((uc >> 1) & 1u) ^ kEmptyFlip, 0x300u + t)) < 0) goto teardown;
            if (kProf) w1 += dt;
            tc_fence_after();
            for (int b = 0; b < bpu; ++b, ++kbc) {
              const uint32_t s = kbc % kStages;
              if ((dt = wait_timed<kProf>(prof, full_a + s * 8u, (kbc / kStages) & 1u, 0x400u + s)) < 0) goto teardown;
              if (kProf) w0 += dt;
              tc_fence_after();
              const uint32_t sa = smem_a + s * (uint32_t)L::kStageBytes;
              if (elect_one()) {
                if (issue) {
                  if (kMagicMma && b == 0) fill(tmem_base + t * BN);
                  if (kPair)
                    mma_i8_block4_pair(tmem_base + t * BN, umma_desc_lo(sa), umma_desc_lo(sa + L::kABytes), idesc_i8,
                                       (kBiased || b > 0) ? 1u : 0u);
                  else
                    mma_i8_block4(tmem_base + t * BN, umma_desc_lo(sa), umma_desc_lo(sa + L::kABytes), idesc_i8,
                                  (kBiased || b > 0) ? 1u : 0u);
                }
                commit(empty_a + s * 8u);                    // smem stage free once these MMAs have read it
                if (b == bpu - 1) commit(tfull_a + t * 8u);  // accumulator stage complete -> epilogue
              }
              __syncwarp();
            }
          }
        } else {
          int u = 0;
          for (int kb = 0; kb < p.kb_int; ++kb, ++kbc) {
            const uint32_t s = kbc % kStages;
            if ((dt = wait_timed<kProf>(prof, full_a + s * 8u, (kbc / kStages) & 1u, 0x400u + s)) < 0) goto teardown;
            if (kProf) w0 += dt;
            tc_fence_after();
            const uint32_t sa = smem_a + s * (uint32_t)L::kStageBytes;
            const uint32_t la = umma_desc_lo(sa), lb = umma_desc_lo(sa + L::kABytes);
            for (int h = 0; h < upb && u < p.G; ++h, ++u, ++uc) {  // units past G are zero padding of the K block
              const uint32_t t = uc & 1u;
              if ((dt = wait_timed<kProf>(prof, tempty_a + t * 8u, ((uc >> 1) & 1u) ^ kEmptyFlip, 0x300u + t)) < 0) goto teardown;
              if (kProf) w1 += dt;
              tc_fence_after();
              if (elect_one()) {
                for (int sub = 0; sub < spu; ++sub) {
                  const uint32_t off = (uint32_t)((h * spu + sub) * 2);
                  if (issue) {
                    if (kMagicMma && sub == 0) fill(tmem_base + t * BN);
                    if (kPair) mma_i8_lo_pair(tmem_base + t * BN, la + off, lb + off, idesc_i8, (kBiased || sub > 0) ? 1u : 0u);
                    else mma_i8_lo(tmem_base + t * BN, la + off, lb + off, idesc_i8, (kBiased || sub > 0) ? 1u : 0u);
                  }
                }
                commit(tfull_a + t * 8u);
              }
              __syncwarp();
            }
            if (elect_one()) commit(empty_a + s * 8u);
            __syncwarp();
          }
        }
        if (p.kb_sal > 0) {  // the salient slice: one more accumulation unit, kind::f16
          const uint32_t t = uc & 1u;
          if ((dt = wait_timed<kProf>(prof, tempty_a + t * 8u, ((uc >> 1) & 1u) ^ kEmptyFlip, 0x300u + t)) < 0) goto teardown;
          if (kProf) w1 += dt;
          tc_fence_after();
          for (int kb = 0; kb < p.kb_sal; ++kb, ++kbc) {
            const uint32_t s = kbc % kStages;
            if ((dt = wait_timed<kProf>(prof, full_a + s * 8u, (kbc / kStages) & 1u, 0x400u + s)) < 0) goto teardown;
            if (kProf) w0 += dt;
            tc_fence_after();
            const uint32_t sa = smem_a + s * (uint32_t)L::kStageBytes;
            if (elect_one()) {
              if (issue) {
                if (kPair)
                  mma_f16_block4_pair(tmem_base + t * BN, umma_desc_lo(sa), umma_desc_lo(sa + L::kABytes), idesc_f16,
                                      kb > 0 ? 1u : 0u);
                else
                  mma_f16_block4(tmem_base + t * BN, umma_desc_lo(sa), umma_desc_lo(sa + L::kABytes), idesc_f16,
                                 kb > 0 ? 1u : 0u);
              }
              commit(empty_a + s * 8u);
              if (kb == p.kb_sal - 1) commit(tfull_a + t * 8u);
            }
            __syncwarp();
          }
          ++uc;
        }
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue (warps 4..11)
    reg_alloc<kEpiRegs>();
    const int ew = warp;
    const int quarter = warp & 3;      // TMEM lane quarter this warp may access
    const int col0 = (ew >> 2) * CW;   // first tile column owned by this thread
    const int row = quarter * 32 + lane;
    // shared-window byte addresses (32-bit) of everything the unit loop touches: no generic-pointer arithmetic inside
#if MQ_GEMM_UNIFORM >= 2
    // lane-0 shuffles: warp-uniform AND opaque, so ptxas keeps the addresses in uniform registers instead of
    // re-deriving the aligned shared-memory base (CgaCtaId, shifts, masks) three times per unit
    auto uni = [](uint32_t v) { return __shfl_sync(0xffffffffu, v, 0); };
    const uint32_t sscale_a = uni(smem_u32(sscale) + (uint32_t)col0 * 4u);
    const uint32_t sas_a = uni(smem_u32(sscale) + (uint32_t)(BN + quarter * 32) * 4u) + (uint32_t)lane * 4u;
    const uint32_t tfull_a = uni(smem_u32(tfull_bar)), tempty_a = uni(smem_u32(tempty_bar));
    const uint32_t sfull_a = uni(smem_u32(sfull_bar)), sempty_a = uni(smem_u32(sempty_bar));
#else
    const uint32_t sscale_a = smem_u32(sscale) + (uint32_t)col0 * 4u;
    const uint32_t sas_a = smem_u32(sscale) + (uint32_t)(BN + row) * 4u;
    const uint32_t tfull_a = smem_u32(tfull_bar), tempty_a = smem_u32(tempty_bar);
    const uint32_t sfull_a = smem_u32(sfull_bar), sempty_a = smem_u32(sempty_bar);
#endif
    // a drained accumulator stage is handed back to the MMA issuer: the leader CTA's barrier in a pair
    const uint32_t tempty_l = kPair ? mapa_u32(tempty_a, 0u) : tempty_a;
    auto release_tmem = [tempty_l](uint32_t t) {
      if (kPair) mbar_arrive_cluster(tempty_l + t * 8u);
      else mbar_arrive_a(tempty_l + t * 8u);
    };
#if MQ_GEMM_UNIFORM >= 2
    const uint32_t tmem_lane = uni(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)col0);
#else
    const uint32_t tmem_lane = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)col0;
#endif
    uint32_t uc = 0, sc = 0;
    uint32_t magic[8];
    if (kRefill) {
#pragma unroll
      for (int i = 0; i < 8; ++i) asm volatile("mov.u32 %0, 0x4B400000;" : "=r"(magic[i]));  // opaque: stays in registers
#pragma unroll 1
      for (int c = 0; c < 2 * BN / 8; ++c)
        if ((c * 8) % BN >= col0 && (c * 8) % BN < col0 + CW) tmem_st_32x8(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(c * 8), magic);
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        release_tmem(0u);
        release_tmem(1u);
      }
    }
    for (int tile = wid; tile < num_tiles; tile += wstride) {
      const int m0 = (tile / p.tiles_n) * TM + (int)cta_rank * BM, n0 = (tile % p.tiles_n) * BN;
      const int m = m0 + row;
      const bool row_ok = m < p.M;
      float2 acc[CW / 2];
#pragma unroll
      for (int j = 0; j < CW / 2; ++j) acc[j] = make_float2(0.f, 0.f);

      // ---- integer units: acc[n] += (v[n] * as) * ws[n]
      for (int u = 0; u < p.G; ++u, ++uc, ++sc) {
        const uint32_t t = uc & 1u, slot = sc % kScaleSlots;
        const uint32_t slot_off = slot * (uint32_t)(L::kSlotFloats * 4);
        if (!kScaleOnFull) {
          if ((dt = wait_timed<kProf>(prof, sfull_a + slot * 8u, (sc / kScaleSlots) & 1u, 0x700u + slot)) < 0) goto teardown;
          if (kProf) w1 += dt;
        }
        if ((dt = wait_timed<kProf>(prof, tfull_a + t * 8u, (uc >> 1) & 1u, 0x500u + t)) < 0) goto teardown;
        if (kProf) w0 += dt;
        tc_fence_after();
        const float as_raw = lds_f32(sas_a + slot_off);  // landed before the unit's MMAs could start (kScaleOnFull)
        const uint32_t taddr = tmem_lane + t * BN;
        const uint32_t ws_a = sscale_a + slot_off;
        const float as = __uint_as_float(__float_as_uint(as_raw) & 0xFFFFFFFCu);
        const float cas = -kMagic * as;
        constexpr int CH = MQ_GEMM_CH;  // columns per TMEM load (double buffered: 2 x CH registers)
        uint32_t v[kNBuf][CH];
#ifdef MQ_DEBUG_NOLDTM
        const bool dbg_nomath = kProf && (p.debug & 4), dbg_noload = true;
#else
        const bool dbg_nomath = kProf && (p.debug & 4), dbg_noload = kProf && (p.debug & 8);
#endif
        if (kNBuf == 2 && !dbg_noload) tmem_ld_chunk(taddr, v[0]);
#pragma unroll
        for (int c = 0; c < CW / CH; ++c) {
          if (kNBuf == 1 && !dbg_noload) tmem_ld_chunk(taddr + (uint32_t)(c * CH), v[0]);
          tmem_ld_wait();
          if (kNBuf == 2 && c + 1 < CW / CH && !dbg_noload) tmem_ld_chunk(taddr + (uint32_t)((c + 1) * CH), v[(c + 1) & (kNBuf - 1)]);
          if (kRefill) {  // chunk c has been read: refill it
#pragma unroll
            for (int r = 0; r < CH / 8; ++r) tmem_st_32x8(taddr + (uint32_t)(c * CH + r * 8), magic);
          }
          if (c + 1 == CW / CH) {  // every load of this thread has landed: hand the TMEM stage back to the MMA warp
            if (kRefill) tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) release_tmem(t);
          }
          const uint32_t* vv = v[c & (kNBuf - 1)];
#ifdef MQ_DEBUG_NOMATH
          acc[c].x += __uint_as_float(vv[0]) + __uint_as_float(vv[15]);
          continue;
#endif
          if (dbg_nomath) {
            acc[c].x += __uint_as_float(vv[0]);
            continue;
          }
#pragma unroll
          for (int q = 0; q < CH / 4; ++q) {
#ifdef MQ_DEBUG_NOLDS   // timing experiment only: no shared-memory read of the weight scales (wrong results)
            const float4 ws = make_float4(as_raw, cas, as_raw, cas);
#else
            const float4 ws = lds_f32x4(ws_a + (uint32_t)((c * CH + q * 4) * 4));
#endif
            const int j = q * 4;
            constexpr uint32_t kAdd = kBiased ? 0u : kMagicBits;
            const float2 f0 = make_float2(__uint_as_float(vv[j + 0] + kAdd), __uint_as_float(vv[j + 1] + kAdd));
            const float2 f1 = make_float2(__uint_as_float(vv[j + 2] + kAdd), __uint_as_float(vv[j + 3] + kAdd));
            const float2 t0 = __ffma2_rn(f0, make_float2(as, as), make_float2(cas, cas));  // == v * as exactly
            const float2 t1 = __ffma2_rn(f1, make_float2(as, as), make_float2(cas, cas));
            // scalar FMAs here: a packed fma.f32x2 with three distinct 64-bit operands is register-file bound
            // (5 clk per SMSP, tools/ubench pipe3) while two scalar FMAs take 2.6
            float2& a0 = acc[(c * CH + j) / 2 + 0];
            float2& a1 = acc[(c * CH + j) / 2 + 1];
            a0.x = fmaf(t0.x, ws.x, a0.x);
            a0.y = fmaf(t0.y, ws.y, a0.y);
            a1.x = fmaf(t1.x, ws.z, a1.x);
            a1.y = fmaf(t1.y, ws.w, a1.y);
          }
        }
        __syncwarp();  // scale slot consumed by every lane
        if (lane == 0) mbar_arrive_a(sempty_a + slot * 8u);
      }

      // ---- salient unit (kind::f16, fp32 accumulator): acc[n] += v[n]
      if (p.kb_sal > 0) {
        const uint32_t t = uc & 1u;
        if ((dt = wait_timed<kProf>(prof, tfull_a + t * 8u, (uc >> 1) & 1u, 0x500u + t)) < 0) goto teardown;
        if (kProf) w0 += dt;
        tc_fence_after();
        const uint32_t taddr = tmem_lane + t * BN;
        uint32_t v[2][16];
        tmem_ld_32x16(taddr, v[0]);
#pragma unroll
        for (int c = 0; c < CW / 16; ++c) {
          tmem_ld_wait();
          if (c + 1 < CW / 16) tmem_ld_32x16(taddr + (uint32_t)((c + 1) * 16), v[(c + 1) & 1]);
          if (kRefill && p.G > 0) {
            tmem_st_32x8(taddr + (uint32_t)(c * 16), magic);
            tmem_st_32x8(taddr + (uint32_t)(c * 16 + 8), magic);
          }
          if (c + 1 == CW / 16) {
            if (kRefill) tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) release_tmem(t);
          }
          const uint32_t* vv = v[c & 1];
#pragma unroll
          for (int j = 0; j < 8; ++j)
            acc[c * 8 + j] = __fadd2_rn(acc[c * 8 + j], make_float2(__uint_as_float(vv[2 * j]),
                                                                    __uint_as_float(vv[2 * j + 1])));
        }
        ++uc;
      }

      // ---- bias, rounding, store (overlaps the next tile's first MMAs).  Each thread owns one row, so a direct store
      // would touch 32 different rows per instruction; instead 32 x 32 slices are transposed through a per-warp
      // shared-memory buffer and leave as row-contiguous 16-byte chunks (8 rows x 64 B per instruction).
      {
        const OutT* bias = static_cast<const OutT*>(p.bias);
        const bool vec_ok = (p.ldy % 8 == 0) && (reinterpret_cast<uintptr_t>(p.y) % 16 == 0);
#ifdef MQ_DEBUG_NOSTORE
        if (acc[0].x == 1.2345e-30f)
#endif
        if (vec_ok) {
          unsigned char* stg = smem + L::kOutOffset + warp * L::kOutWarpBytes;
          const int rr = lane >> 2, cc = lane & 3;  // read-back: row within a group of 8, 16-byte chunk within 64 B
#pragma unroll
          for (int sl = 0; sl < CW / 32; ++sl) {
            const int nb = n0 + col0 + sl * 32;  // first output column of this slice
            if (nb < p.N) {                       // warp-uniform
#pragma unroll
              for (int c4 = 0; c4 < 4; ++c4) {
                __align__(16) OutT o[8];
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                  const float2 av = acc[sl * 16 + c4 * 4 + e / 2];
                  float r = (e & 1) ? av.y : av.x;
                  if (bias && nb + c4 * 8 + e < p.N) r += Num<OutT>::to_f(bias[nb + c4 * 8 + e]);
                  o[e] = Num<OutT>::from_f(r);
                }
                *reinterpret_cast<uint4*>(stg + lane * L::kOutPitch + c4 * 16) = *reinterpret_cast<const uint4*>(o);
              }
              __syncwarp();
#pragma unroll
              for (int it = 0; it < 4; ++it) {
                const int r = it * 8 + rr;
                const int gm = m0 + quarter * 32 + r, gn = nb + cc * 8;
                const uint4 val = *reinterpret_cast<const uint4*>(stg + r * L::kOutPitch + cc * 16);
                if (gm < p.M && gn < p.N) {
                  OutT* dst = static_cast<OutT*>(p.y) + (int64_t)gm * p.ldy + gn;
                  if (gn + 8 <= p.N) {
                    *reinterpret_cast<uint4*>(dst) = val;
                  } else {
                    const OutT* ve = reinterpret_cast<const OutT*>(&val);
                    for (int e = 0; e < 8 && gn + e < p.N; ++e) dst[e] = ve[e];
                  }
                }
              }
              __syncwarp();
            }
          }
        } else if (row_ok) {  // unaligned output: element stores straight from the row owner
          OutT* yrow = static_cast<OutT*>(p.y) + (int64_t)m * p.ldy + n0 + col0;
#pragma unroll
          for (int j = 0; j < CW; ++j) {  // fully unrolled: the accumulators must stay in registers
            const int n = n0 + col0 + j;
            if (n < p.N) {
              float r = (j & 1) ? acc[j / 2].y : acc[j / 2].x;
              if (bias) r += Num<OutT>::to_f(bias[n]);
              yrow[j] = Num<OutT>::from_f(r);
            }
          }
        }
      }
    }
  }

teardown:
  if (kProf && prof && lane == 0) {
    unsigned long long* pr = g_gemm_prof + blockIdx.x * kProfSlots;
    const long long total = clock64() - k_t0;
    if (warp == kProducerWarp) {
      pr[0] = (unsigned long long)total;
      pr[1] = globaltimer_ns() - k_ns0;
      pr[2] = (unsigned long long)w0;
    } else if (warp == kMmaWarp) {
      pr[3] = (unsigned long long)w0;
      pr[4] = (unsigned long long)w1;
    } else if (warp == 0) {
      pr[5] = (unsigned long long)w0;
      pr[6] = (unsigned long long)w1;
      pr[7] = (unsigned long long)(total - w0 - w1);
    }
  }
  tc_fence_before();
  if (kPair) cluster_sync_all();  // no CTA leaves while its peer may still signal its barriers or read its operands
  else __syncthreads();
  if (warp == kAllocWarp) {
    tc_fence_after();
    if (kPair) tmem_dealloc_pair<kTmemCols>(tmem_base);
    else tmem_dealloc<kTmemCols>(tmem_base);
  }
}

// ------------------------------------------------------------------------------------------------ host side
// tensor-map cache (mq_make_map) and the per-kernel shared-memory attribute helper live in tma_host.cu

template <typename OutT, int BN, int kStages, bool kPair, bool kMagicMma>
int launch_gemm_m(const CUtensorMap& ma, const CUtensorMap& mw, const CUtensorMap& mas, const CUtensorMap& mws,
                  const CUtensorMap& msa, const CUtensorMap& msw, GemmArgs p, cudaStream_t st) {
  using L = SmemLayout<BN, kStages, kPair>;
  p.debug = mq_get_option("gemm_debug", 0);
  p.profile = mq_get_option("gemm_profile", 0);
  auto kern = p.profile ? mixed_gemm_kernel<OutT, BN, kStages, true, kPair, kMagicMma>
                        : mixed_gemm_kernel<OutT, BN, kStages, false, kPair, kMagicMma>;
  if (int rc = mq_ensure_smem_attr(reinterpret_cast<const void*>(kern), L::kTotal)) return rc;
  constexpr int TM = kPair ? 2 * BM : BM;
  p.tiles_m = (p.M + TM - 1) / TM;
  p.tiles_n = (p.N + BN - 1) / BN;
  const int64_t tiles = (int64_t)p.tiles_m * p.tiles_n;
  const int64_t workers = kPair ? (int64_t)mq_worker_sms() / 2 : (int64_t)mq_worker_sms();
  int64_t grid = mq_get_option("gemm_persistent", 0) == 2 ? tiles : workers;
  if (grid > tiles) grid = tiles;
  if (!kPair) {
    kern<<<(unsigned)grid, kThreads, L::kTotal, st>>>(ma, mw, mas, mws, msa, msw, p);
  } else {  // one cluster of two CTAs (the two SMs of a TPC) per worker
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)(2 * grid));
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = L::kTotal;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    MQ_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kern, ma, mw, mas, mws, msa, msw, p));
  }
  MQ_CHECK_CUDA(cudaGetLastError());
  return MQ_OK;
}

template <typename OutT, int BN, int kStages, bool kPair = false>
int launch_gemm(const CUtensorMap& ma, const CUtensorMap& mw, const CUtensorMap& mas, const CUtensorMap& mws,
                const CUtensorMap& msa, const CUtensorMap& msw, const GemmArgs& p, cudaStream_t st) {
  const int mode = mq_get_option("gemm_magic_mma", kMagicMmaMode);
  const bool magic = mode == 2 || (mode == 1 && p.g <= kMagicMmaMaxGroup);
  if (magic) return launch_gemm_m<OutT, BN, kStages, kPair, true>(ma, mw, mas, mws, msa, msw, p, st);
  return launch_gemm_m<OutT, BN, kStages, kPair, false>(ma, mw, mas, mws, msa, msw, p, st);
}

template <typename OutT>
int dispatch_gemm(const CUtensorMap& ma, const CUtensorMap& mw, const CUtensorMap& mas, const CUtensorMap& mws,
                  const CUtensorMap& msa, const CUtensorMap& msw, const GemmArgs& p, int bn, bool pair,
                  cudaStream_t st) {
  if (bn == 256 && pair) return launch_gemm<OutT, 256, 6, true>(ma, mw, mas, mws, msa, msw, p, st);
  if (bn == 256) return launch_gemm<OutT, 256, 4>(ma, mw, mas, mws, msa, msw, p, st);
  return launch_gemm<OutT, 128, 6>(ma, mw, mas, mws, msa, msw, p, st);
}

}  // namespace

extern "C" {

int mq_mixed_gemm(const int8_t* a_codes, int64_t lda, const float* a_scales, int64_t ld_as, const void* a_sal,
                  int64_t ld_asal, const int8_t* w_codes, int64_t ldw, const float* w_scales, int64_t ld_ws,
                  const void* w_sal, int64_t ld_wsal, const void* bias, int64_t M, int64_t N, int64_t G,
                  int group_size, int64_t Ks, int dtype, void* y, int64_t ldy, void* stream) {
  MQ_CHECK_ARG(y && M >= 0 && N >= 1 && ldy >= N, "mq_mixed_gemm: bad output (M=%lld N=%lld ldy=%lld)", (long long)M,
               (long long)N, (long long)ldy);
  MQ_CHECK_ARG(dtype == MQ_F16 || dtype == MQ_BF16, "mq_mixed_gemm: dtype must be MQ_F16 or MQ_BF16");
  MQ_CHECK_ARG(G >= 0 && Ks >= 0 && (G > 0 || Ks > 0), "mq_mixed_gemm: empty contraction");
  MQ_CHECK_ARG(M < (1ll << 31) && N < (1ll << 31), "mq_mixed_gemm: shape too large");
  if (M == 0) return MQ_OK;
  int64_t kc = 0;
  if (G > 0) {
    MQ_CHECK_ARG(group_size >= 32 && group_size % 32 == 0 && (128 % group_size == 0 || group_size % 128 == 0),
                 "mq_mixed_gemm: group_size=%d must be 32, 64 or a multiple of 128", group_size);
    kc = (G * group_size + 127) / 128 * 128;
    MQ_CHECK_ARG(a_codes && w_codes && a_scales && w_scales, "mq_mixed_gemm: null code / scale pointer");
    MQ_CHECK_ARG(lda >= kc && ldw >= kc && lda % 16 == 0 && ldw % 16 == 0,
                 "mq_mixed_gemm: lda=%lld ldw=%lld must be multiples of 16 and >= %lld", (long long)lda,
                 (long long)ldw, (long long)kc);
    MQ_CHECK_ARG(reinterpret_cast<uintptr_t>(a_codes) % 16 == 0 && reinterpret_cast<uintptr_t>(w_codes) % 16 == 0,
                 "mq_mixed_gemm: code pointers must be 16-byte aligned");
    MQ_CHECK_ARG(ld_as >= M && ld_ws >= N && ld_ws % 4 == 0 && ld_as % 4 == 0 &&
                     reinterpret_cast<uintptr_t>(w_scales) % 16 == 0 && reinterpret_cast<uintptr_t>(a_scales) % 16 == 0,
                 "mq_mixed_gemm: scales need ld_as >= M, ld_ws >= N, ld %% 4 == 0 and 16-byte aligned pointers");
  }
  if (Ks > 0) {
    MQ_CHECK_ARG(a_sal && w_sal && ld_asal >= Ks && ld_wsal >= Ks && ld_asal % 8 == 0 && ld_wsal % 8 == 0,
                 "mq_mixed_gemm: salient operands need ld >= Ks and ld %% 8 == 0");
    MQ_CHECK_ARG(reinterpret_cast<uintptr_t>(a_sal) % 16 == 0 && reinterpret_cast<uintptr_t>(w_sal) % 16 == 0,
                 "mq_mixed_gemm: salient pointers must be 16-byte aligned");
  }
  int bn = mq_get_option("gemm_bn", 0);
  if (bn != 128 && bn != 256) bn = N >= 256 ? 256 : 128;
  // CTA pairs (256 x 256 tiles over two SMs, each staging half of the W tile).  Option gemm_pair: 1 always (when
  // BN = 256 and M > 128), 2 never; default: whenever the pairs still get at least two tiles each -- below that the
  // single-CTA kernel spreads the few tiles over twice as many workers (2048 x 4096: 0.066 vs 0.067 ms).
  const int pair_opt = mq_get_option("gemm_pair", MQ_GEMM_PAIR_DEFAULT);
  const int64_t pair_tiles = ((M + 2 * BM - 1) / (2 * BM)) * ((N + 255) / 256);
  const bool pair = bn == 256 && M > BM && (pair_opt == 1 || (pair_opt == 3 && pair_tiles >= (int64_t)mq_sm_count()));
  const int wbox = pair ? bn / 2 : bn;  // W rows per TMA box

  CUtensorMap ma, mw, mas, mws, msa, msw;
  int rc;
  if (G > 0) {
    if ((rc = mq_make_map(&ma, a_codes, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, M, kc, lda, BM))) return rc;
    if ((rc = mq_make_map(&mw, w_codes, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, N, kc, ldw, wbox))) return rc;
    // scale vectors: [G, M] / [G, N] fp32, one row of BM / BN values per unit (rows and columns past M / N read 0)
    if ((rc = mq_make_map(&msa, a_scales, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, G, M, ld_as, 1, BM))) return rc;
    if ((rc = mq_make_map(&msw, w_scales, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, G, N, ld_ws, 1, bn))) return rc;
  }
  const CUtensorMapDataType dt16 = dtype == MQ_F16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  if (Ks > 0) {  // the tensor maps cover exactly Ks columns: TMA zero-fills the tail of the last 64-column block
    if ((rc = mq_make_map(&mas, a_sal, dt16, 2, M, Ks, ld_asal, BM))) return rc;
    if ((rc = mq_make_map(&mws, w_sal, dt16, 2, N, Ks, ld_wsal, wbox))) return rc;
  } else {
    mas = ma;
    mws = mw;
  }
  if (G == 0) {
    ma = mas;
    mw = mws;
    msa = mas;
    msw = mws;
  }
  GemmArgs p{};
  p.a_scales = a_scales; p.ld_as = ld_as; p.w_scales = w_scales; p.ld_ws = ld_ws; p.bias = bias;
  p.y = y; p.ldy = ldy; p.M = (int)M; p.N = (int)N; p.G = (int)G; p.g = G > 0 ? group_size : 32;
  p.kb_int = (int)(kc / 128); p.kb_sal = (int)((Ks + 63) / 64);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == MQ_F16) return dispatch_gemm<__half>(ma, mw, mas, mws, msa, msw, p, bn, pair, st);
  return dispatch_gemm<__nv_bfloat16>(ma, mw, mas, mws, msa, msw, p, bn, pair, st);
}

// debug: per-CTA wait-cycle profile of the last mixed GEMM launched with option gemm_profile != 0 (see kProfSlots).
// Copies min(n_ctas, 256) x 8 u64 to out_host.  Synchronises the device.
int mq_gemm_profile(unsigned long long* out_host, int n_ctas) {
  if (n_ctas > kProfCtas) n_ctas = kProfCtas;
  if (n_ctas <= 0 || !out_host) return MQ_OK;
  MQ_CHECK_CUDA(cudaMemcpyFromSymbol(out_host, g_gemm_prof, sizeof(unsigned long long) * kProfSlots * n_ctas));
  return MQ_OK;
}

// debug: fault record of the GEMM's bounded waits (see ptx.cuh); clears it.  Synchronises the device.
int mq_gemm_fault(unsigned int* out8_host) {
  unsigned int z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  MQ_CHECK_CUDA(cudaMemcpyFromSymbol(out8_host, g_mq_fault, sizeof(z)));
  MQ_CHECK_CUDA(cudaMemcpyToSymbol(g_mq_fault, z, sizeof(z)));
  return MQ_OK;
}

}  // extern "C"
