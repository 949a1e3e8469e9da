// Host-side TMA plumbing (see tma_host.cuh).  No kernels here.
#include <atomic>
#include <mutex>
#include <unordered_map>

#include "common.cuh"
#include "tma_host.cuh"

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn) return fn;
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
      q != cudaDriverEntryPointSuccess || !p)
    return nullptr;
  fn = reinterpret_cast<EncodeTiledFn>(p);
  return fn;
}

// 2-D row-major tensor [rows, cols] of `esize`-byte elements with leading dimension ld (elements); box = 128 bytes
// of the inner dimension x box_rows rows, 128-byte swizzle.
//
// Descriptors are cached per (pointer, dtype, shape, ld, box): a CUtensorMap is a pure function of those values (it
// holds the global address, no allocation handle), so an entry stays valid when the allocator hands the same address
// out again.  The weight operand's three maps are encoded once per layer, the activation's once per distinct buffer
// address -- with torch's caching allocator that is a handful of addresses per model -- instead of six driver calls
// per launch.  Host-side only: nothing here touches the device or the stream, so a launch through
// mq_mixed_gemm can be captured into a CUDA graph.
struct MapKey {
  const void* ptr;
  int64_t rows, cols, ld;
  int dt, esize, box_rows, box_cols;
  bool operator==(const MapKey& o) const {
    return ptr == o.ptr && rows == o.rows && cols == o.cols && ld == o.ld && dt == o.dt && esize == o.esize &&
           box_rows == o.box_rows && box_cols == o.box_cols;
  }
};
struct MapKeyHash {
  size_t operator()(const MapKey& k) const {
    uint64_t h = reinterpret_cast<uintptr_t>(k.ptr) * 0x9E3779B97F4A7C15ull;
    auto mix = [&h](uint64_t v) { h = (h ^ v) * 0xFF51AFD7ED558CCDull; h ^= h >> 29; };
    mix((uint64_t)k.rows); mix((uint64_t)k.cols); mix((uint64_t)k.ld);
    mix(((uint64_t)k.dt << 48) | ((uint64_t)k.esize << 40) | ((uint64_t)k.box_rows << 16) | (uint64_t)k.box_cols);
    return (size_t)h;
  }
};
constexpr size_t kMapCacheMax = 8192;   // ~1.2 MB of descriptors; cleared wholesale when full
std::mutex g_map_mutex;
std::unordered_map<MapKey, CUtensorMap, MapKeyHash> g_map_cache;
std::atomic<long long> g_map_hits{0}, g_map_misses{0};

int mq_make_map(CUtensorMap* map, const void* ptr, CUtensorMapDataType dt, int esize, int64_t rows, int64_t cols,
                int64_t ld, int box_rows, int box_cols) {
  const MapKey key{ptr, rows, cols, ld, (int)dt, esize, box_rows, box_cols};
  {
    std::lock_guard<std::mutex> lk(g_map_mutex);
    auto it = g_map_cache.find(key);
    if (it != g_map_cache.end()) {
      *map = it->second;
      g_map_hits.fetch_add(1, std::memory_order_relaxed);
      return MQ_OK;
    }
  }
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) {
    mq_set_error("cuTensorMapEncodeTiled entry point not available");
    return MQ_ERR_CUDA;
  }
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * (cuuint64_t)esize};
  // box_cols == 0: operand tile (128-byte rows, 128-byte swizzle); otherwise a plain row vector of box_cols elements
  cuuint32_t box[2] = {(cuuint32_t)(box_cols ? box_cols : 128 / esize), (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(map, dt, 2, const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   box_cols ? CU_TENSOR_MAP_SWIZZLE_NONE : CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    mq_set_error("cuTensorMapEncodeTiled failed with CUresult %d (rows=%lld cols=%lld ld=%lld esize=%d)", (int)r,
                 (long long)rows, (long long)cols, (long long)ld, esize);
    return MQ_ERR_CUDA;
  }
  g_map_misses.fetch_add(1, std::memory_order_relaxed);
  std::lock_guard<std::mutex> lk(g_map_mutex);
  if (g_map_cache.size() >= kMapCacheMax) g_map_cache.clear();
  g_map_cache.emplace(key, *map);
  return MQ_OK;
}

// cudaFuncSetAttribute once per kernel instantiation and device (not per launch)
int mq_ensure_smem_attr(const void* kern, int bytes) {
  static std::mutex mu;
  static std::unordered_map<uint64_t, int> done;   // (function, device) -> bytes
  int dev = 0;
  MQ_CHECK_CUDA(cudaGetDevice(&dev));
  const uint64_t key = (uint64_t)reinterpret_cast<uintptr_t>(kern) * 64u + (uint64_t)(dev & 63);
  std::lock_guard<std::mutex> lk(mu);
  auto it = done.find(key);
  if (it != done.end() && it->second >= bytes) return MQ_OK;
  MQ_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  done[key] = bytes;
  return MQ_OK;
}

