/*
 * mq.h -- C ABI of libmq.so: the B200 (sm_100a) implementation of the W4A4 MixedQuant linear hot path.
 *
 * The reference (adithyab100/smoothquant-mixedprecision) is pure PyTorch and has no FFI; the "binding" a
 * maintainer adds is a ctypes stub.  Each entry point below names the reference code it
 * replaces, relative to /root/reference.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless the name ends in _host; sizes are element counts unless named
 *     *_bytes; `ld*` are leading dimensions in ELEMENTS of the pointed-to type
 *   - every call is asynchronous on `stream` (a cudaStream_t passed as void*), never synchronises, never
 *     allocates device memory, never throws, never exits the process
 *   - return value: 0 on success, a negative MQ_ERR_* otherwise; mq_last_error() gives a thread-local message
 *   - dtype codes: MQ_F16 / MQ_BF16 / MQ_F32 (the dtype the reference tensor lives in; all arithmetic is rounded
 *     to that dtype exactly like the reference's in-place torch ops)
 *   - there is no CPU fallback: without a CUDA device every compute call returns MQ_ERR_CUDA
 */
#ifndef MQ_H_
#define MQ_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MQ_ABI_VERSION 2

enum { MQ_F16 = 0, MQ_BF16 = 1, MQ_F32 = 2 };

enum {
  MQ_OK = 0,
  MQ_ERR_ARG = -1,        /* invalid argument (shape, alignment, dtype, null pointer)                     */
  MQ_ERR_CUDA = -2,       /* CUDA runtime / driver error, or no device                                      */
  MQ_ERR_UNSUPPORTED = -3,/* combination not implemented by this build (message says which)                 */
  MQ_ERR_WORKSPACE = -4   /* workspace too small                                                            */
};

/* column statistic kinds for mq_colstat */
enum {
  MQ_STAT_ABSMAX = 0,     /* max_m |x_mk|                      fake_quant.py:113 / :164 ; calibration.py:18-25 */
  MQ_STAT_ABSMEAN = 1,    /* mean_m |x_mk|                     run_experiments.py:60 (importance)              */
  MQ_STAT_MEAN3STD = 2,   /* mean + 3*std (unbiased) of |x|    README.md:36 -- parity unpinned                 */
  MQ_STAT_ARGMAX = 3      /* first row index attaining max|x|  README.md:37 -- parity unpinned                 */
};

int         mq_version(void);
const char* mq_last_error(void);
/* sm_count / cc_major / cc_minor of the current device; any pointer may be NULL */
int         mq_device_info(int* sm_count, int* cc_major, int* cc_minor);

/* ------------------------------------------------------------------------------------------------------------
 * K0  column statistics over the rows of x [M, K] (row-major, leading dimension ld).
 *     Replaces  t.abs().max(dim=0).values            fake_quant.py:113 (activations), :164 (weights)
 *               x.view(-1,K).abs().mean(dim=0)       run_experiments.py:60
 *     out_key: float32 [K].  workspace: mq_colstat_workspace_bytes(K, kind) bytes.
 * ---------------------------------------------------------------------------------------------------------- */
size_t mq_colstat_workspace_bytes(int64_t K, int kind);
int mq_colstat(const void* x, int64_t M, int64_t K, int64_t ld, int dtype, int kind,
               float* out_key, void* workspace, size_t workspace_bytes, void* stream);

/* mq_colstat for callers that may not allocate (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc): no workspace argument -- the per-split partial statistics go to a file-scope static __device__
 * array of mq_colstat_static_max_bytes() = 4 MiB inside the library (held by every GPU that loads it, until exit).
 * No counterpart in the reference.
 *     Same result contract as mq_colstat (ABSMAX / ARGMAX identical; ABSMEAN / MEAN3STD may differ in the last
 *     float bit when the split count had to be reduced to fit the scratch).
 *     Limit: K * 24 bytes <= 4 MiB, i.e. K <= 174 762 columns; beyond that MQ_ERR_WORKSPACE with a message.
 *     Calls no cudaMalloc / cudaFree / allocating routine.  All calls on a device SHARE the scratch: calls whose
 *     kernels may overlap on the GPU (different streams) race -- serialise them.  The device must be current. */
size_t mq_colstat_static_max_bytes(void);
int mq_colstat_static(const void* x, int64_t M, int64_t K, int64_t ld, int dtype, int kind,
                      float* out_key, void* stream);

/* K0b stable ascending argsort of n float32 keys (ties -> ascending index; the canonical order of SURVEY F3).
 *     Replaces  torch.argsort(col_max_val)           fake_quant.py:116, :167
 *     With descending != 0: descending keys, ties still by ascending index  (fake_quant.py:268).
 *     `subset` (nullable, int32 [n]) : sort only these column ids, keyed by key[subset[i]]; the output then
 *     holds column ids (used to sort the non-salient columns, fake_quant.py:293-300).
 *     out_perm: int32 [n].  workspace: mq_argsort_workspace_bytes(n). */
size_t mq_argsort_workspace_bytes(int64_t n);
int mq_argsort_stable(const float* key, const int32_t* subset, int64_t n, int descending, int32_t* out_perm,
                      void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * K1 / P0  fused row quantiser (activations: rows = tokens; weights: rows = output channels).
 *     For every row r:  v_j = x[r, perm[j]] (optionally / smooth[perm[j]]),  j in [0, Kn)
 *                       groups of `group_size` consecutive j (last one zero-padded),
 *                       scale = clamp(absmax, 1e-5) / qmax,  code = round_half_even(v / scale)     (dtype-rounded)
 *     Replaces  quantize_activation_per_group_absmax   fake_quant.py:78-101  applied to x[:, perm]
 *               quantize_weight_per_group_absmax       fake_quant.py:30-53   applied to W[:, perm]
 *               the boolean-mask salient split         fake_quant.py:291-298, :347
 *     Outputs   codes  int8  [R, ld_codes]   codes in [-qmax, qmax]; columns [Kn, ld_codes) are written as 0
 *               scales float [G, ld_scales]  TRANSPOSED (group-major) and widened to fp32: scales[g*ld + r]
 *               sal    dtype [R, ld_sal]     x[r, sal[j]] for j < Ks, zeros in [Ks, ld_sal)  (nullable if Ks==0)
 *     perm: int32 [Kn] original column of permuted position j; sal: int32 [Ks]; smooth: dtype [K] or NULL.
 *     Constraints: K <= 65535; ld_codes >= G*group_size (the GEMM wants it rounded up to 128); ld_scales >= R.
 * ---------------------------------------------------------------------------------------------------------- */
int mq_quant_rows(const void* x, int64_t R, int64_t K, int64_t ld, int dtype, const void* smooth,
                  const int32_t* perm, int64_t Kn, const int32_t* sal, int64_t Ks,
                  int group_size, int n_bits,
                  int8_t* out_codes, int64_t ld_codes, float* out_scales, int64_t ld_scales,
                  void* out_sal, int64_t ld_sal, void* stream);

/* Hybrid split of the same quantiser: the first `int_groups` groups leave as integer codes + scales exactly as above
 * (for the tcgen05 kind::i8 units of mq_mixed_gemm), the remaining groups leave DEQUANTISED -- code * scale rounded to
 * dtype, i.e. the very values quantize_*_per_group_absmax returns (fake_quant.py:97, :49) -- in out_sal columns
 * [t0, t0 + (G - int_groups) * group_size), t0 = Ks rounded up to 8, where mq_mixed_gemm contracts them together with
 * the salient columns as kind::f16 MMAs.  The per-group rescale of the integer units runs on the CUDA cores and
 * bounds the GEMM for group sizes <= 128; moving part of K to the 16-bit tensor path balances the two pipes
 * (DESIGN.md 3).  Needs ld_codes >= int_groups*group_size and ld_sal >= t0 + (G - int_groups)*group_size.
 * int_groups == G is mq_quant_rows. */
int mq_quant_rows_split(const void* x, int64_t R, int64_t K, int64_t ld, int dtype, const void* smooth,
                        const int32_t* perm, int64_t Kn, const int32_t* sal, int64_t Ks,
                        int group_size, int n_bits, int64_t int_groups,
                        int8_t* out_codes, int64_t ld_codes, float* out_scales, int64_t ld_scales,
                        void* out_sal, int64_t ld_sal, void* stream);

/* mq_quant_rows on a column-BLOCKED input: the K columns of a row live in x_blocks blocks of blk_cols columns
 * (the last one may be narrower), block b at x + b * blk_stride elements, rows of a block `ld` elements apart --
 * [x_blocks][R][blk_cols] is what an all-gather of the column shards of a column-parallel producer leaves behind
 * (torch.distributed.all_gather_into_tensor; the reference has no tensor parallelism, its
 * F.linear -> quantize_activation_per_group_absmax chain, fake_quant.py:306 -> :78-101, sees a dense tensor).  The
 * kernel's bulk copies reassemble every row in shared memory, so the gathered activations are never re-laid-out
 * in HBM.  Results are bit-identical to mq_quant_rows on the dense tensor.  16-bit dtypes, power-of-two group
 * sizes 32..1024, blk_cols % 8 == 0, blk_stride % 8 == 0, no smoothing divisor; anything else returns
 * MQ_ERR_UNSUPPORTED (gather into a dense tensor and call mq_quant_rows).  x_blocks == 1 is mq_quant_rows. */
int mq_quant_rows_blocked(const void* x, int64_t R, int64_t K, int x_blocks, int64_t blk_cols, int64_t blk_stride,
                          int64_t ld, int dtype, const int32_t* perm, int64_t Kn, const int32_t* sal, int64_t Ks,
                          int group_size, int n_bits,
                          int8_t* out_codes, int64_t ld_codes, float* out_scales, int64_t ld_scales,
                          void* out_sal, int64_t ld_sal, void* stream);

/* Producer-side fusion: the norm in front of q/k/v and gate/up fused into the row quantiser -- x is
 * read once, the normalised row lives only in shared memory.
 *     Replaces  LlamaRMSNorm / nn.LayerNorm (whose weights smooth_lm has already divided by the smoothing scale,
 *               smooth.py:126-141, :12-29, :33-54) followed by mq_quant_rows on its output:
 *       MQ_NORM_RMS    y = rn(w * rn(x32 * rsqrt(mean(x32^2) + eps)))        (16-bit roundings as in transformers)
 *       MQ_NORM_LAYER  y = rn(((x32 - mean) * rsqrt(var + eps)) * w32 + b32)  (torch.nn.functional.layer_norm)
 *     then codes / scales / salient copy of y exactly as mq_quant_rows (bit-identical given the same y).
 *     Row statistics are one-pass fp32 sums over all K columns (LayerNorm: var = E[x^2] - mean^2, clamped at 0): a row
 *     whose |mean| exceeds ~1e3 x its standard deviation loses digits against torch's Welford update -- not a regime
 *     transformer activations reach; y agrees with the torch norm to last-place units of the 16-bit type otherwise
 *     (tests/test_ops_gpu.py::test_norm_quant_rows_is_norm_then_row_quantiser).
 *     out_y (nullable): the normalised rows [R, ld_y], for consumers outside the path and for the parity tests.
 *     16-bit dtypes, K % 8 == 0, K <= 16384, power-of-two group sizes 32..1024; otherwise MQ_ERR_UNSUPPORTED. */
enum mq_norm_kind { MQ_NORM_RMS = 1, MQ_NORM_LAYER = 2 };
int mq_norm_quant_rows(const void* x, int64_t R, int64_t K, int64_t ld, int dtype, int norm_kind,
                       const void* norm_weight, const void* norm_bias, float eps,
                       const int32_t* perm, int64_t Kn, const int32_t* sal, int64_t Ks, int group_size, int n_bits,
                       int8_t* out_codes, int64_t ld_codes, float* out_scales, int64_t ld_scales,
                       void* out_sal, int64_t ld_sal, void* out_y, int64_t ld_y, void* stream);

/* Same arithmetic, reference-shaped output: out [R, K] (dtype) holds code*scale at the ORIGINAL column of every
 * permuted position and x itself at every column not listed in perm (salient pass-through).
 *     Replaces  quantize_activation_per_group_absmax_sort   fake_quant.py:105-154  (perm = argsort of the key)
 *               quantize_weight_per_group_absmax_sort       fake_quant.py:157-207
 *               per-token / per-channel quantisers          fake_quant.py:57-64, :10-16 (group_size = Kn, identity perm)
 *               the scatter-back                            fake_quant.py:301, :363-365
 *     `out` may alias `x` (in-place, like the reference's per_token path). */
int mq_fakequant_rows(const void* x, int64_t R, int64_t K, int64_t ld, int dtype,
                      const int32_t* perm, int64_t Kn, int group_size, int n_bits,
                      void* out, int64_t ld_out, void* stream);

/* per-tensor fake quantiser (one global scale): fake_quant.py:68-75, :20-26.  workspace >= 16 bytes. */
int mq_fakequant_tensor(const void* x, int64_t n, int dtype, int n_bits, void* out,
                        void* workspace, size_t workspace_bytes, void* stream);

/* int4 storage format of the weight packer: two's-complement nibbles, even column in the low nibble.
 * codes [R, ld_codes] int8  <->  packed [R, ld_packed] uint8 (cols/2 bytes used per row). */
int mq_pack_int4(const int8_t* codes, int64_t R, int64_t cols, int64_t ld_codes,
                 uint8_t* packed, int64_t ld_packed, void* stream);
int mq_unpack_int4(const uint8_t* packed, int64_t R, int64_t cols, int64_t ld_packed,
                   int8_t* codes, int64_t ld_codes, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * K2  fused mixed-precision GEMM  (tcgen05 kind::i8 per K-group + kind::f16 salient slice, one CTA, TMEM accum.)
 *     y[m, n] = sum_g a_scales[g, m] * w_scales[g, n] * sum_{k in g} a_codes[m, k] * w_codes[n, k]
 *             + sum_j a_sal[m, j] * w_sal[n, j]  + bias[n]
 *     Replaces  torch.functional.F.linear(q_x, self.weight, self.bias)     fake_quant.py:306
 *     a_codes [M, lda] int8, w_codes [N, ldw] int8    (K-major, G*group_size valid columns, zero padded to 128)
 *     a_scales [G, ld_as] fp32, w_scales [G, ld_ws] fp32 (group-major, as written by mq_quant_rows)
 *     a_sal [M, ld_asal], w_sal [N, ld_wsal] in `dtype` (MQ_F16 or MQ_BF16), Ks valid columns (any count; the
 *           kernel walks them in blocks of 64 and TMA zero-fills the tail), ld % 8 == 0, 16-byte aligned
 *     bias [N] in `dtype` or NULL;  y [M, ldy] in `dtype`.
 *     G == 0 runs the pure 16-bit GEMM (mode="reference": dequantised operands in a_sal / w_sal);
 *     Ks == 0 runs the pure integer path.  group_size must be 32, 64 or a multiple of 128.
 *     No workspace: the kernel needs none.  TMA descriptors are cached inside the library per (pointer, shape, ld),
 *     and the call touches neither the device nor the stream besides the launch, so it is CUDA-graph capturable.
 * ---------------------------------------------------------------------------------------------------------- */
int mq_mixed_gemm(const int8_t* a_codes, int64_t lda, const float* a_scales, int64_t ld_as,
                  const void* a_sal, int64_t ld_asal,
                  const int8_t* w_codes, int64_t ldw, const float* w_scales, int64_t ld_ws,
                  const void* w_sal, int64_t ld_wsal,
                  const void* bias, int64_t M, int64_t N, int64_t G, int group_size, int64_t Ks,
                  int dtype, void* y, int64_t ldy, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * K2d  decode regime (M <= 64 tokens) of the same F.linear (fake_quant.py:306) with INT4-RESIDENT weights: the
 *      weight operand is read as two's-complement nibbles -- 4 bits per non-salient weight, the reference's own
 *      accounting (smoothquant/model_size.py:5-15) -- from a layout made once by mq_pack_decode:
 *        nibbles [ceil(N/16) row tiles][ceil(G*g/256) blocks] records of 2048 + 32 * max(1, 256/g) bytes: the
 *                block's nibbles as [4 units of 64 columns][32 lanes][16 bytes], then the weight scales of the
 *                groups in the block as [slot][16 rows] values of `dtype` (exact: the scales are values of it)
 *        salient [ceil(N/16) row tiles][ceil(Ks/64) blocks][4 steps of 16 columns][32 lanes][16 bytes]
 *      (a lane's 16 bytes are its mma.sync A-fragment registers).  mq_decode_weight_bytes gives
 *      the two buffer sizes.  Activation operands, bias and y are exactly those of mq_mixed_gemm; all groups must be
 *      integer groups (no hybrid tail), quant_bits <= 4, group_size a power of two >= 64.
 * ---------------------------------------------------------------------------------------------------------- */
int mq_decode_weight_bytes(int64_t N, int64_t G, int group_size, int64_t Ks, size_t* nib_bytes, size_t* sal_bytes);
int mq_pack_decode(const int8_t* w_codes, int64_t ldw, const float* w_scales, int64_t ld_ws,
                   const void* w_sal, int64_t ld_wsal, int64_t N, int64_t G,
                   int group_size, int64_t Ks, int dtype, uint8_t* out_nib, uint8_t* out_sal, void* stream);
int mq_decode_gemm(const int8_t* a_codes, int64_t lda, const float* a_scales, int64_t ld_as,
                   const void* a_sal, int64_t ld_asal,
                   const uint8_t* w_nib, const uint8_t* w_sal_dec,
                   const void* bias, int64_t M, int64_t N, int64_t G, int group_size, int64_t Ks,
                   int dtype, void* y, int64_t ldy, void* stream);

/* tuning knobs (process-global, for benchmarks/tests): name in {"gemm_bn", "gemm_stages", "quant_rows_per_stage",
 * "quant_warps", "quant_force_generic", ...}; value 0 restores the default.  MQ_ERR_ARG for unknown names. */
int mq_set_option(const char* name, int value);

/* debug: every in-kernel barrier wait is bounded (5 s); one that expires records {code, block, thread, parity,
 * spins>>10, 0, 0, 0} and traps (sticky CUDA error on the stream) instead of hanging the GPU or returning garbage.  These copy the record of the quantiser / GEMM kernels to
 * out8_host (8 x uint32, HOST pointer) and clear it.  code == 0 means no fault.  They synchronise the device. */
int mq_quant_fault(unsigned int* out8_host);
int mq_gemm_fault(unsigned int* out8_host);

/* debug: with option "gemm_profile" != 0 the mixed GEMM records, per CTA, 8 x uint64: kernel cycles, kernel ns,
 * cycles the TMA producer waited for a free stage, the MMA warp waited for operands / for a free accumulator, the
 * first epilogue warp waited for an accumulator / for scales, and that warp's busy cycles.  Copies
 * min(n_ctas, 256) records to out_host (HOST pointer).  Synchronises the device. */
int mq_gemm_profile(unsigned long long* out_host, int n_ctas);

/* ------------------------------------------------------------------------------------------------------------
 * Column-parallel plumbing; the reference is single-GPU (its only multi-GPU touchpoint is
 * device_map="auto", ppl_eval.py:70), so these have no counterpart there.
 *
 * mq_peer_gather: all-gather of the [M, S] column blocks of a column-parallel layer over peer-mapped memory
 * (cuMem mappings with peer access, e.g. torch.distributed._symmetric_memory), done by the copy engines and
 * ordered by stream memory operations -- no kernel, so it runs underneath the persistent GEMM of the next chunk.
 *     bufs[world]   device address, in THIS process, of every rank's gather buffer [world][M][S]; the caller's GEMM
 *                   has written block `rank` of bufs[rank] in stream order before the call
 *     flags[world]  device address of every rank's flag block, uint32 [4 + ring buffers][world], zero at start
 *     seq           1, 2, 3, ... : the same sequence of calls on every rank
 *     streams[n]    cudaStream_t handles (as integers): the peers are dealt round-robin over them so that several
 *                   copy engines pull in parallel; every one must already be ordered after the caller's GEMM
 *     On return the streams hold: publish ready (streams[0]), and per peer wait-ready / DMA its block into
 *     bufs[rank] / publish pulled.  Work ordered after ALL the streams sees the full [world][M][S].
 * mq_peer_wait_pulled: makes `stream` wait until every peer has copied this rank's block of gather `seq` out --
 *     call it before overwriting a gather buffer last used by gather `seq`. */
int mq_peer_gather(const uint64_t* bufs, const uint64_t* flags, int rank, int world, size_t block_bytes,
                   uint32_t seq, const uint64_t* streams, int n_streams);
int mq_peer_wait_pulled(const uint64_t* flags, int rank, int world, uint32_t seq, void* stream);
/* Push variant of the same gather (flag block: uint32 [4 + ring buffers][world] = ready, pulled, arrived, unused, then
 * one row consumed[buf_id] per ring buffer -- the gathers of ONE buffer are consumed in order).  mq_peer_push:
 * once every peer has consumed gather `prev` (the previous user of this ring buffer; 0 = none), copy this rank's block
 * into block `rank` of every peer's buffer -- or with ONE copy to mc_buf, the multicast address of the buffers, when
 * mc_buf != 0 -- and publish arrived[rank] = seq.  mq_peer_wait_arrived: the stream that reads the gathered buffer waits
 * for every peer's block.  mq_peer_publish_consumed: called on the compute stream before a ring buffer is reused. */
int mq_peer_push(const uint64_t* bufs, const uint64_t* flags, uint64_t mc_buf, int rank, int world,
                 size_t block_bytes, uint32_t seq, uint32_t prev, int buf_id, const uint64_t* streams,
                 int n_streams);
int mq_peer_wait_arrived(const uint64_t* flags, int rank, int world, uint32_t seq, void* stream);
int mq_peer_publish_consumed(const uint64_t* flags, int rank, int world, uint32_t seq, int buf_id, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* MQ_H_ */
