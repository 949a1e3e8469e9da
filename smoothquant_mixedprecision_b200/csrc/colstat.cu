// K0 column statistics over the rows of x [M, K] and K0b stable argsort of the resulting keys.
//
// Replaces (paths relative to /root/reference):
//   t.abs().max(dim=0).values                 smoothquant/fake_quant.py:113 (activations), :164 (weights)
//   x.view(-1, K).abs().mean(dim=0)           run_experiments/run_experiments.py:60   (importance)
//   torch.argsort(col_max_val)                smoothquant/fake_quant.py:116, :167, :268
//
// HBM-bound: one coalesced pass over x (16-byte loads, a warp covers 512 contiguous bytes of a row), per-CTA
// partial statistics in a workspace, then a fixed-order final reduction so that results are deterministic.
#include <mutex>

#include "common.cuh"

// Scratch of mq_colstat_static (the no-allocation variant at the end of this file): per-split partial statistics.
// Lives in every GPU context that loads the library until the process exits.  In HOST code this name is only a
// stand-in for the symbol -- the device address comes from cudaGetSymbolAddress (static_scratch below).
#define MQ_COLSTAT_STATIC_BYTES (4u << 20)
static __device__ __align__(256) unsigned char g_colstat_static_scratch[MQ_COLSTAT_STATIC_BYTES];

namespace {

constexpr int kColstatMaxSplits = 128;
constexpr int kRowLanes = 4;  // blockDim.y

template <typename T>
struct Vec;  // 16-byte vector of T
template <>
struct Vec<__half> {
  static constexpr int N = 8;
};
template <>
struct Vec<__nv_bfloat16> {
  static constexpr int N = 8;
};
template <>
struct Vec<float> {
  static constexpr int N = 4;
};

struct Partial {  // one per (split, column); interpretation depends on the kind
  float a;        // ABSMAX: max        ABSMEAN: unused      ARGMAX: max
  int b;          //                                          ARGMAX: first row attaining it
  double s1;      // ABSMEAN / MEAN3STD: sum |x|
  double s2;      // MEAN3STD: sum x^2
};

template <int KIND>
__device__ __forceinline__ void acc_init(Partial& p) {
  p.a = -1.0f;
  p.b = 0x7fffffff;
  p.s1 = 0.0;
  p.s2 = 0.0;
}
template <int KIND>
__device__ __forceinline__ void acc_add(Partial& p, float v, int row) {
  float av = fabsf(v);
  if (KIND == MQ_STAT_ABSMAX) {
    p.a = fmaxf(p.a, av);
  } else if (KIND == MQ_STAT_ABSMEAN) {
    p.s1 += (double)av;
  } else if (KIND == MQ_STAT_MEAN3STD) {
    p.s1 += (double)av;
    p.s2 += (double)av * (double)av;
  } else {  // ARGMAX: first row attaining the maximum
    if (av > p.a) {
      p.a = av;
      p.b = row;
    }
  }
}
template <int KIND>
__device__ __forceinline__ void acc_merge(Partial& p, const Partial& o) {
  if (KIND == MQ_STAT_ABSMAX) {
    p.a = fmaxf(p.a, o.a);
  } else if (KIND == MQ_STAT_ABSMEAN) {
    p.s1 += o.s1;
  } else if (KIND == MQ_STAT_MEAN3STD) {
    p.s1 += o.s1;
    p.s2 += o.s2;
  } else {
    if (o.a > p.a || (o.a == p.a && o.b < p.b)) {
      p.a = o.a;
      p.b = o.b;
    }
  }
}

// grid (col tiles of 32*VEC columns, splits); block (32, kRowLanes)
template <typename T, int KIND>
__global__ void __launch_bounds__(32 * kRowLanes)
colstat_partial_kernel(const T* __restrict__ x, int64_t M, int64_t K, int64_t ld, Partial* __restrict__ part,
                       bool vec_ok) {
  constexpr int VEC = Vec<T>::N;
  const int tx = threadIdx.x, ty = threadIdx.y;
  const int64_t col0 = ((int64_t)blockIdx.x * 32 + tx) * VEC;
  Partial acc[VEC];
#pragma unroll
  for (int v = 0; v < VEC; ++v) acc_init<KIND>(acc[v]);

  if (col0 < K) {
    const bool full = vec_ok && (col0 + VEC <= K);
    for (int64_t r = (int64_t)blockIdx.y * kRowLanes + ty; r < M; r += (int64_t)gridDim.y * kRowLanes) {
      const T* p = x + r * ld + col0;
      if (full) {
        uint4 raw = __ldg(reinterpret_cast<const uint4*>(p));
        const T* e = reinterpret_cast<const T*>(&raw);
#pragma unroll
        for (int v = 0; v < VEC; ++v) acc_add<KIND>(acc[v], Num<T>::to_f(e[v]), (int)r);
      } else {
#pragma unroll
        for (int v = 0; v < VEC; ++v)
          if (col0 + v < K) acc_add<KIND>(acc[v], Num<T>::to_f(p[v]), (int)r);
      }
    }
  }
  // combine the kRowLanes row lanes in a fixed order through shared memory
  __shared__ Partial sm[kRowLanes][32 * Vec<T>::N + 1];
#pragma unroll
  for (int v = 0; v < VEC; ++v) sm[ty][tx * VEC + v] = acc[v];
  __syncthreads();
  for (int c = ty * 32 + tx; c < 32 * VEC; c += 32 * kRowLanes) {
    int64_t col = (int64_t)blockIdx.x * 32 * VEC + c;
    if (col >= K) continue;
    Partial p = sm[0][c];
    for (int l = 1; l < kRowLanes; ++l) acc_merge<KIND>(p, sm[l][c]);
    part[(int64_t)blockIdx.y * K + col] = p;
  }
}

template <int KIND>
__global__ void colstat_final_kernel(const Partial* __restrict__ part, int splits, int64_t M, int64_t K,
                                     float* __restrict__ out) {
  int64_t col = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (col >= K) return;
  Partial p = part[col];
  for (int s = 1; s < splits; ++s) acc_merge<KIND>(p, part[(int64_t)s * K + col]);
  float r;
  if (KIND == MQ_STAT_ABSMAX) {
    r = p.a;
  } else if (KIND == MQ_STAT_ABSMEAN) {
    r = (float)(p.s1 / (double)M);
  } else if (KIND == MQ_STAT_MEAN3STD) {
    double mean = p.s1 / (double)M;
    double var = M > 1 ? (p.s2 - (double)M * mean * mean) / (double)(M - 1) : 0.0;
    if (var < 0.0) var = 0.0;
    r = (float)(mean + 3.0 * sqrt(var));
  } else {
    r = (float)p.b;
  }
  out[col] = r;
}

int colstat_splits(int64_t M) {
  int64_t s = (M + 63) / 64;
  if (s < 1) s = 1;
  if (s > kColstatMaxSplits) s = kColstatMaxSplits;
  return (int)s;
}

template <typename T, int KIND>
int launch_colstat(const void* x, int64_t M, int64_t K, int64_t ld, float* out, void* ws, cudaStream_t st) {
  constexpr int VEC = Vec<T>::N;
  int splits = colstat_splits(M);
  bool vec_ok = (reinterpret_cast<uintptr_t>(x) % 16 == 0) && (ld % VEC == 0);
  dim3 grid((unsigned)((K + 32 * VEC - 1) / (32 * VEC)), (unsigned)splits);
  dim3 block(32, kRowLanes);
  colstat_partial_kernel<T, KIND><<<grid, block, 0, st>>>(static_cast<const T*>(x), M, K, ld,
                                                           static_cast<Partial*>(ws), vec_ok);
  MQ_CHECK_CUDA(cudaGetLastError());
  colstat_final_kernel<KIND><<<(unsigned)((K + 255) / 256), 256, 0, st>>>(static_cast<const Partial*>(ws), splits, M,
                                                                          K, out);
  MQ_CHECK_CUDA(cudaGetLastError());
  return MQ_OK;
}

template <typename T>
int dispatch_kind(int kind, const void* x, int64_t M, int64_t K, int64_t ld, float* out, void* ws, cudaStream_t st) {
  switch (kind) {
    case MQ_STAT_ABSMAX: return launch_colstat<T, MQ_STAT_ABSMAX>(x, M, K, ld, out, ws, st);
    case MQ_STAT_ABSMEAN: return launch_colstat<T, MQ_STAT_ABSMEAN>(x, M, K, ld, out, ws, st);
    case MQ_STAT_MEAN3STD: return launch_colstat<T, MQ_STAT_MEAN3STD>(x, M, K, ld, out, ws, st);
    case MQ_STAT_ARGMAX: return launch_colstat<T, MQ_STAT_ARGMAX>(x, M, K, ld, out, ws, st);
  }
  mq_set_error("mq_colstat: unknown kind %d", kind);
  return MQ_ERR_ARG;
}

// ---------------------------------------------------------------- no-allocation variant (mq_colstat_static)
// Same two kernels with the split count passed in (it is capped by what fits the static scratch).
template <typename T, int KIND>
int launch_colstat_static(const void* x, int64_t M, int64_t K, int64_t ld, float* out, Partial* ws, int splits,
                          cudaStream_t st) {
  constexpr int VEC = Vec<T>::N;
  const bool vec_ok = (reinterpret_cast<uintptr_t>(x) % 16 == 0) && (ld % VEC == 0);
  dim3 grid((unsigned)((K + 32 * VEC - 1) / (32 * VEC)), (unsigned)splits);
  dim3 block(32, kRowLanes);
  colstat_partial_kernel<T, KIND><<<grid, block, 0, st>>>(static_cast<const T*>(x), M, K, ld, ws, vec_ok);
  MQ_CHECK_CUDA(cudaGetLastError());
  colstat_final_kernel<KIND><<<(unsigned)((K + 255) / 256), 256, 0, st>>>(ws, splits, M, K, out);
  MQ_CHECK_CUDA(cudaGetLastError());
  return MQ_OK;
}

template <typename T>
int dispatch_kind_static(int kind, const void* x, int64_t M, int64_t K, int64_t ld, float* out, Partial* ws, int splits,
                         cudaStream_t st) {
  switch (kind) {
    case MQ_STAT_ABSMAX: return launch_colstat_static<T, MQ_STAT_ABSMAX>(x, M, K, ld, out, ws, splits, st);
    case MQ_STAT_ABSMEAN: return launch_colstat_static<T, MQ_STAT_ABSMEAN>(x, M, K, ld, out, ws, splits, st);
    case MQ_STAT_MEAN3STD: return launch_colstat_static<T, MQ_STAT_MEAN3STD>(x, M, K, ld, out, ws, splits, st);
    case MQ_STAT_ARGMAX: return launch_colstat_static<T, MQ_STAT_ARGMAX>(x, M, K, ld, out, ws, splits, st);
  }
  mq_set_error("mq_colstat_static: unknown kind %d", kind);
  return MQ_ERR_ARG;
}

// Device address of g_colstat_static_scratch on the CURRENT device, resolved once per device (set-up) and kept.
constexpr int kMaxDevices = 64;
int static_scratch(Partial** out) {
  static std::mutex mu;
  static void* ptrs[kMaxDevices] = {};
  int dev = -1;
  MQ_CHECK_CUDA(cudaGetDevice(&dev));
  MQ_CHECK_ARG(dev >= 0 && dev < kMaxDevices, "mq_colstat_static: device ordinal %d outside [0, %d)", dev, kMaxDevices);
  std::lock_guard<std::mutex> lock(mu);
  if (ptrs[dev] == nullptr) {
    void* p = nullptr;
    MQ_CHECK_CUDA(cudaGetSymbolAddress(&p, g_colstat_static_scratch));
    MQ_CHECK_ARG(p != nullptr, "mq_colstat_static: cudaGetSymbolAddress returned a null address");
    ptrs[dev] = p;
  }
  *out = static_cast<Partial*>(ptrs[dev]);
  return MQ_OK;
}

// ---------------------------------------------------------------- K0b: stable argsort by ranking
// rank_i = #{ j : key_j < key_i  or (key_j == key_i and j < i) }  -> out[rank_i] = id_i.  O(n^2) compares, fully
// parallel and stable by construction; n <= 65535 keys (28 672 for Llama-2-70B) is ~1e9 compares at most.
constexpr int kSortTile = 2048;

// total order on the keys: NaN compares false with everything, which would give every NaN key the same rank and
// leave output slots unwritten.  Keys are compared through the usual order-preserving map float -> uint32 with NaN
// canonicalised to the positive quiet NaN (above +inf, where torch.argsort puts it) and -0 to +0 (equal keys tie by
// index); descending order is the complement.
__device__ __forceinline__ uint32_t sort_canon(float k, int descending) {
  k += 0.0f;
  uint32_t u = (k != k) ? 0x7fc00000u : __float_as_uint(k);
  u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
  return descending ? ~u : u;
}

__global__ void __launch_bounds__(256)
argsort_rank_kernel(const float* __restrict__ key, const int32_t* __restrict__ subset, int n, int descending,
                    int32_t* __restrict__ out) {
  __shared__ uint32_t sk[kSortTile];
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const bool live = i < n;
  int id = 0;
  uint32_t ki = 0u;
  if (live) {
    id = subset ? subset[i] : i;
    ki = sort_canon(key[id], descending);
  }
  int rank = 0;
  for (int base = 0; base < n; base += kSortTile) {
    int cnt = min(kSortTile, n - base);
    __syncthreads();
    for (int t = threadIdx.x; t < cnt; t += blockDim.x) {
      int jd = subset ? subset[base + t] : base + t;
      sk[t] = sort_canon(key[jd], descending);
    }
    __syncthreads();
    if (live) {
      // j < i part of the tile: ties count; j > i part: only strictly smaller keys count
      int split = min(max(i - base, 0), cnt);
      int r = 0;
      for (int t = 0; t < split; ++t) r += (sk[t] <= ki) ? 1 : 0;
      for (int t = split; t < cnt; ++t) r += (sk[t] < ki) ? 1 : 0;
      rank += r;
    }
  }
  if (live) out[rank] = id;
}

}  // namespace

extern "C" {

size_t mq_colstat_workspace_bytes(int64_t K, int kind) {
  (void)kind;
  if (K < 0) return 0;
  return (size_t)kColstatMaxSplits * (size_t)K * sizeof(Partial);
}

int mq_colstat(const void* x, int64_t M, int64_t K, int64_t ld, int dtype, int kind, float* out_key, void* workspace,
               size_t workspace_bytes, void* stream) {
  MQ_CHECK_ARG(x && out_key && workspace, "mq_colstat: null pointer");
  MQ_CHECK_ARG(M >= 1 && K >= 1 && ld >= K, "mq_colstat: bad shape M=%lld K=%lld ld=%lld", (long long)M, (long long)K,
               (long long)ld);
  MQ_CHECK_ARG(M < (1ll << 31), "mq_colstat: M too large");
  if (workspace_bytes < (size_t)colstat_splits(M) * (size_t)K * sizeof(Partial)) {
    mq_set_error("mq_colstat: workspace too small");
    return MQ_ERR_WORKSPACE;
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (dtype) {
    case MQ_F16: return dispatch_kind<__half>(kind, x, M, K, ld, out_key, workspace, st);
    case MQ_BF16: return dispatch_kind<__nv_bfloat16>(kind, x, M, K, ld, out_key, workspace, st);
    case MQ_F32: return dispatch_kind<float>(kind, x, M, K, ld, out_key, workspace, st);
  }
  mq_set_error("mq_colstat: unknown dtype %d", dtype);
  return MQ_ERR_ARG;
}

size_t mq_colstat_static_max_bytes(void) { return (size_t)MQ_COLSTAT_STATIC_BYTES; }

int mq_colstat_static(const void* x, int64_t M, int64_t K, int64_t ld, int dtype, int kind, float* out_key,
                      void* stream) {
  MQ_CHECK_ARG(x && out_key, "mq_colstat_static: null pointer");
  MQ_CHECK_ARG(M >= 1 && K >= 1 && ld >= K, "mq_colstat_static: bad shape M=%lld K=%lld ld=%lld", (long long)M,
               (long long)K, (long long)ld);
  MQ_CHECK_ARG(M < (1ll << 31), "mq_colstat_static: M too large");
  const size_t per_split = (size_t)K * sizeof(Partial);
  if (per_split > (size_t)MQ_COLSTAT_STATIC_BYTES) {
    mq_set_error("mq_colstat_static: K=%lld needs %zu bytes of scratch, the static scratch holds %zu (at most %zu "
                 "columns); use mq_colstat with a caller-provided workspace", (long long)K, per_split,
                 (size_t)MQ_COLSTAT_STATIC_BYTES, (size_t)MQ_COLSTAT_STATIC_BYTES / sizeof(Partial));
    return MQ_ERR_WORKSPACE;
  }
  int splits = colstat_splits(M);
  const int fit = (int)((size_t)MQ_COLSTAT_STATIC_BYTES / per_split);
  if (splits > fit) splits = fit;
  Partial* ws = nullptr;
  int rc = static_scratch(&ws);
  if (rc) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // nothing has to start at zero: the partial kernel writes every (split, column < K) cell the final kernel reads
  switch (dtype) {
    case MQ_F16: return dispatch_kind_static<__half>(kind, x, M, K, ld, out_key, ws, splits, st);
    case MQ_BF16: return dispatch_kind_static<__nv_bfloat16>(kind, x, M, K, ld, out_key, ws, splits, st);
    case MQ_F32: return dispatch_kind_static<float>(kind, x, M, K, ld, out_key, ws, splits, st);
  }
  mq_set_error("mq_colstat_static: unknown dtype %d", dtype);
  return MQ_ERR_ARG;
}

size_t mq_argsort_workspace_bytes(int64_t n) {
  (void)n;
  return 0;
}

int mq_argsort_stable(const float* key, const int32_t* subset, int64_t n, int descending, int32_t* out_perm,
                      void* workspace, size_t workspace_bytes, void* stream) {
  (void)workspace;
  (void)workspace_bytes;
  MQ_CHECK_ARG(key && out_perm, "mq_argsort_stable: null pointer");
  MQ_CHECK_ARG(n >= 0 && n <= 1 << 20, "mq_argsort_stable: n=%lld out of range", (long long)n);
  if (n == 0) return MQ_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  argsort_rank_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(key, subset, (int)n, descending ? 1 : 0, out_perm);
  MQ_CHECK_CUDA(cudaGetLastError());
  return MQ_OK;
}

}  // extern "C"
