// Thin inline-PTX wrappers used by the sm_100a kernels: mbarrier, bulk / tensor TMA copies, tcgen05 (TMEM alloc,
// MMA, commit, ld), proxy fences.  Every wait is bounded (see mbar_wait).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace mqptx {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// one lane of the (fully active) warp; ptxas keeps the operands of the guarded instruction in uniform registers
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(pred));
  return pred != 0;
}

// ------------------------------------------------------------------------------------------------ mbarrier
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}

// Fault record of this translation unit: [0] code (0 = none), [1] block, [2] thread, [3] parity, [4] spins>>10.
// A wait that does not complete within kWaitLimitNs records itself here and then TRAPS: the launch ends with a sticky
// CUDA error that the next synchronising call of the host (torch) raises, so a broken pipeline can neither hang the
// GPU nor hand back uninitialised outputs silently.  The record survives the trap only for a debugger; hosts that
// want it read it with the TU's own mq_*_fault() helper before the context is torn down.
static __device__ unsigned int g_mq_fault[8];
constexpr unsigned long long kWaitLimitNs = 5000ull * 1000ull * 1000ull;

static __device__ __noinline__ void mq_wait_expired(uint32_t parity, uint32_t code, uint32_t spins) {
  if (atomicCAS(&g_mq_fault[0], 0u, code) == 0u) {
    g_mq_fault[1] = blockIdx.x + (blockIdx.y << 16);
    g_mq_fault[2] = threadIdx.x;
    g_mq_fault[3] = parity;
    g_mq_fault[4] = spins >> 10;
    __threadfence_system();
  }
  __trap();
}

__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

static __device__ __forceinline__ bool mbar_wait_slow(uint64_t* bar, uint32_t parity, uint32_t code) {
  // the global timer is slow to read: it is first sampled after 1024 failed polls, never on the common path
  unsigned long long t0 = 0;
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if ((++spins & 0x3FFu) == 0) {
      if (spins == 0x400u) t0 = globaltimer_ns();
      if (globaltimer_ns() - t0 > kWaitLimitNs) {
        mq_wait_expired(parity, code, spins);
        return false;
      }
    }
  }
  return true;
}
__device__ __forceinline__ bool mbar_wait(uint64_t* bar, uint32_t parity, uint32_t code) {
  if (mbar_try_wait(bar, parity)) return true;
  return mbar_wait_slow(bar, parity, code);
}

// Variants on 32-bit shared-window addresses (hot loops: no generic -> shared conversion per use)
__device__ __forceinline__ void mbar_arrive_a(uint32_t bar_addr) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar_addr) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait_a(uint32_t bar_addr, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar_addr), "r"(parity)
      : "memory");
  return ok != 0;
}
static __device__ __forceinline__ bool mbar_wait_slow_a(uint32_t bar_addr, uint32_t parity, uint32_t code) {
  // the global timer is slow to read: it is first sampled after 1024 failed polls, never on the common path
  unsigned long long t0 = 0;
  uint32_t spins = 0;
  while (!mbar_try_wait_a(bar_addr, parity)) {
    if ((++spins & 0x3FFu) == 0) {
      if (spins == 0x400u) t0 = globaltimer_ns();
      if (globaltimer_ns() - t0 > kWaitLimitNs) {
        mq_wait_expired(parity, code, spins);
        return false;
      }
    }
  }
  return true;
}
__device__ __forceinline__ bool mbar_wait_a(uint32_t bar_addr, uint32_t parity, uint32_t code) {
  if (mbar_try_wait_a(bar_addr, parity)) return true;
  return mbar_wait_slow_a(bar_addr, parity, code);
}
__device__ __forceinline__ float lds_f32(uint32_t addr) {
  float r;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(r) : "r"(addr) : "memory");
  return r;
}
__device__ __forceinline__ float4 lds_f32x4(uint32_t addr) {
  float4 r;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "r"(addr) : "memory");
  return r;
}

// ------------------------------------------------------------------------------------------------ fences
// Programmatic dependent launch: a kernel launched with cudaLaunchAttributeProgrammaticStreamSerialization may start
// while its predecessor in the stream is still running; griddep_wait() blocks until the predecessor has completed and
// its memory is visible (a no-op for an ordinary launch), griddep_launch() lets the successor's CTAs be scheduled as
// soon as resources free up.
__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void griddep_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// ------------------------------------------------------------------------------------------------ bulk copies (TMA)
// 1-D bulk copy global -> shared, completion counted in bytes on an mbarrier.  16-byte aligned, size % 16 == 0.
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
// 2-D tiled tensor copy global -> shared through a CUtensorMap.
__device__ __forceinline__ void tma_load_2d(void* dst_smem, const CUtensorMap* map, int32_t c0, int32_t c1,
                                            uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_u32(dst_smem)),
      "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}

// ------------------------------------------------------------------------------------------------ tcgen05 / TMEM
template <int kCols>
__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem) {  // whole warp
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)),
               "n"(kCols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <int kCols>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr) {  // whole warp
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(kCols) : "memory");
}
// arrive on an mbarrier once all previously issued tcgen05.mma of this thread have completed
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc]; kind::i8 -> int32 accumulate, kind::f16 -> fp32 accumulate
__device__ __forceinline__ void mma_i8_ss(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void mma_f16_ss(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                           uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// kind::f8f6f4 with both operands E4M3 bytes: the integer codes -7..7 are exact in E4M3, products and the fp32
// sums of a K group stay integers far below 2^24, so this path yields the same sums as kind::i8 -- as floats
__device__ __forceinline__ void mma_f8_ss(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// TMEM -> registers: this warp's 32 lanes x 32 consecutive 32-bit columns (thread t gets lane base+t)
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_32x16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_st_32x16(uint32_t taddr, const uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%16], "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15};" ::"r"(v[0]),
      "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]), "r"(v[10]),
      "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_st_32x8(uint32_t taddr, const uint32_t (&v)[8]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x8.b32 [%8], {%0, %1, %2, %3, %4, %5, %6, %7};" ::"r"(v[0]),
      "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// ------------------------------------------------------------------------------------------------ descriptors
// Shared-memory matrix descriptor for a K-major operand tile stored as rows of exactly 128 bytes with the
// 128-byte swizzle (what TMA writes with CU_TENSOR_MAP_SWIZZLE_128B): 8-row atoms of 1024 bytes, so the stride
// between 8-row groups (SBO) is 1024 bytes; LBO is unused for swizzled K-major layouts.  Bits: [0,14) address>>4,
// [16,30) LBO>>4, [32,46) SBO>>4, [46,48) version = 1 (Blackwell), [61,64) layout type = 2 (SWIZZLE_128B).
__device__ __forceinline__ uint64_t umma_desc_k128(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
// Instruction descriptor (upper 32 bits of idesc): [4,6) D format (1 = f32, 2 = s32), [7,10) A format,
// [10,13) B format (i8: 1 = signed; f16 kind: 0 = f16, 1 = bf16), [15] A major, [16] B major (0 = K),
// [17,23) N >> 3, [24,29) M >> 4.
__host__ __device__ constexpr uint32_t umma_idesc(uint32_t d_fmt, uint32_t ab_fmt, uint32_t M, uint32_t N) {
  return (d_fmt << 4) | (ab_fmt << 7) | (ab_fmt << 10) | ((N >> 3) << 17) | ((M >> 4) << 24);
}

// Four consecutive 32-byte K slices of one 128-byte-swizzled K block in ONE asm statement (descriptor low words
// advance by 32 B >> 4 = 2 per slice; the high word -- SBO 1024 B, version 1, SWIZZLE_128B -- is constant), so that
// the issuing thread spends a handful of instructions per K block instead of a descriptor rebuild per MMA.
constexpr uint32_t kDescHiK128 = 0x40004040u;  // == umma_desc_k128(x) >> 32
#define MQ_MMA_BLOCK4(NAME, KIND, GRP)                                                                                 \
  __device__ __forceinline__ void NAME(uint32_t tmem_d, uint32_t lo_a, uint32_t lo_b, uint32_t idesc,              \
                                       uint32_t acc_first) {                                                        \
    asm volatile(                                                                                                   \
        "{\n\t.reg .pred p, q;\n\t.reg .b64 da, db;\n\t.reg .b32 la, lb;\n\t"                                      \
        "setp.ne.b32 p, %4, 0;\n\t"                                                                                 \
        "setp.eq.b32 q, 0, 0;\n\t"                                                                                  \
        "mov.b64 da, {%1, %5};\n\tmov.b64 db, {%2, %5};\n\t"                                                        \
        "tcgen05.mma.cta_group::" GRP ".kind::" KIND " [%0], da, db, %3, p;\n\t"                                        \
        "add.u32 la, %1, 2;\n\tadd.u32 lb, %2, 2;\n\tmov.b64 da, {la, %5};\n\tmov.b64 db, {lb, %5};\n\t"            \
        "tcgen05.mma.cta_group::" GRP ".kind::" KIND " [%0], da, db, %3, q;\n\t"                                        \
        "add.u32 la, %1, 4;\n\tadd.u32 lb, %2, 4;\n\tmov.b64 da, {la, %5};\n\tmov.b64 db, {lb, %5};\n\t"            \
        "tcgen05.mma.cta_group::" GRP ".kind::" KIND " [%0], da, db, %3, q;\n\t"                                        \
        "add.u32 la, %1, 6;\n\tadd.u32 lb, %2, 6;\n\tmov.b64 da, {la, %5};\n\tmov.b64 db, {lb, %5};\n\t"            \
        "tcgen05.mma.cta_group::" GRP ".kind::" KIND " [%0], da, db, %3, q;\n\t}" ::"r"(tmem_d),                        \
        "r"(lo_a), "r"(lo_b), "r"(idesc), "r"(acc_first), "r"(kDescHiK128)                                          \
        : "memory");                                                                                                \
  }
MQ_MMA_BLOCK4(mma_i8_block4, "i8", "1")
MQ_MMA_BLOCK4(mma_f16_block4, "f16", "1")
// CTA-pair forms (M = 256 over two SMs; issued by the leader CTA only)
MQ_MMA_BLOCK4(mma_i8_block4_pair, "i8", "2")
MQ_MMA_BLOCK4(mma_f16_block4_pair, "f16", "2")
#undef MQ_MMA_BLOCK4
// low word of the K-major SWIZZLE_128B descriptor of the tile at shared address `smem_addr`
__device__ __forceinline__ uint32_t umma_desc_lo(uint32_t smem_addr) {
  return ((smem_addr & 0x3FFFFu) >> 4) | (1u << 16);
}
__device__ __forceinline__ void mma_i8_lo(uint32_t tmem_d, uint32_t lo_a, uint32_t lo_b, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "mov.b64 da, {%1, %5};\n\tmov.b64 db, {%2, %5};\n\t"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], da, db, %3, p;\n\t}" ::"r"(tmem_d),
      "r"(lo_a), "r"(lo_b), "r"(idesc), "r"(accumulate), "r"(kDescHiK128)
      : "memory");
}
// One kind::f16 MMA (K = 16) whose A and B descriptors both point at ONE 1024-byte constant atom with a stride of 0
// between 8-row groups: every row of both operands reads the same 128 bytes, so D[m][n] = sum_k c_k * c_k for all
// (m, n).  Used to set a whole accumulator stage to a constant without touching it from the epilogue threads.
constexpr uint32_t kDescHiConst = 0x40004000u;  // SBO 0, version 1, SWIZZLE_128B
#define MQ_MMA_CONST(NAME, GRP)                                                                                        \
  __device__ __forceinline__ void NAME(uint32_t tmem_d, uint32_t lo, uint32_t idesc) {                                 \
    asm volatile(                                                                                                      \
        "{\n\t.reg .pred p;\n\t.reg .b64 da;\n\t"                                                                    \
        "setp.ne.b32 p, 0, 0;\n\t"                                                                                     \
        "mov.b64 da, {%1, %3};\n\t"                                                                                    \
        "tcgen05.mma.cta_group::" GRP ".kind::f16 [%0], da, da, %2, p;\n\t}" ::"r"(tmem_d),                              \
        "r"(lo), "r"(idesc), "r"(kDescHiConst)                                                                         \
        : "memory");                                                                                                   \
  }
MQ_MMA_CONST(mma_const_fill, "1")
MQ_MMA_CONST(mma_const_fill_pair, "2")
#undef MQ_MMA_CONST
__device__ __forceinline__ void tc_commit_a(uint32_t bar_addr) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar_addr) : "memory");
}

// ------------------------------------------------------------------------------------------------ CTA pairs
// cta_group::2: two CTAs of one cluster (the two SMs of a TPC) run one M = 256 MMA; each CTA stages its own 128 rows
// of A and HALF of the B tile, owns the 128 accumulator lanes of its rows, and only the leader (rank 0) issues.
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {  // every thread of both CTAs
  __syncwarp();
  asm volatile("barrier.cluster.arrive.release;\n\tbarrier.cluster.wait.acquire;" ::: "memory");
}
// shared::cluster address of the same shared-memory offset in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t mapa_u32(uint32_t cta_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(cta_addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// tile into THIS CTA's shared memory, completion bytes onto an mbarrier given by its shared::cluster address (the
// leader's full barrier)
__device__ __forceinline__ void tma_load_2d_pair(void* dst_smem, const CUtensorMap* map, int32_t c0, int32_t c1,
                                                 uint32_t bar_cluster_addr) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], "
      "[%2];" ::"r"(smem_u32(dst_smem)),
      "l"(map), "r"(bar_cluster_addr), "r"(c0), "r"(c1)
      : "memory");
}
template <int kCols>
__device__ __forceinline__ void tmem_alloc_pair(uint32_t* dst_smem) {  // the same warp of BOTH CTAs
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)),
               "n"(kCols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
template <int kCols>
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(kCols) : "memory");
}
// arrive on the mbarrier at this shared-memory offset in BOTH CTAs once the pair's MMAs issued so far have completed
__device__ __forceinline__ void tc_commit_pair(uint32_t bar_addr) {
  asm volatile(
      "{\n\t.reg .b16 m;\n\tmov.b16 m, 3;\n\t"
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], m;\n\t}" ::"r"(
          bar_addr)
      : "memory");
}
__device__ __forceinline__ void mma_i8_lo_pair(uint32_t tmem_d, uint32_t lo_a, uint32_t lo_b, uint32_t idesc,
                                               uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "mov.b64 da, {%1, %5};\n\tmov.b64 db, {%2, %5};\n\t"
      "tcgen05.mma.cta_group::2.kind::i8 [%0], da, db, %3, p;\n\t}" ::"r"(tmem_d),
      "r"(lo_a), "r"(lo_b), "r"(idesc), "r"(accumulate), "r"(kDescHiK128)
      : "memory");
}

}  // namespace mqptx
