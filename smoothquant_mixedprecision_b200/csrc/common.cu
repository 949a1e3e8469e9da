// libmq.so host-side plumbing: error convention, options, device queries.  No kernels here.
#include <stdarg.h>
#include <string.h>

#include <atomic>

#include "common.cuh"

static thread_local char g_err[512] = "";

void mq_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

struct MqOption {
  const char* name;
  std::atomic<int> value;
};
static MqOption g_options[] = {
    {"gemm_bn", {0}},       {"gemm_stages", {0}}, {"quant_rows_per_stage", {0}}, {"quant_warps", {0}},
    {"gemm_persistent", {0}}, {"quant_exact_div", {0}}, {"gemm_debug", {0}}, {"quant_force_generic", {0}},
    {"gemm_profile", {0}}, {"gemm_kernel", {0}}, {"gemm_pair", {0}}, {"gemm_magic_mma", {0}},
    {"quant_h2", {0}}, {"decode_bps", {0}}, {"decode_tpc", {0}}, {"decode_split", {0}},
    {"sm_reserve", {0}}, {"pdl", {0}},
};

int mq_get_option(const char* name, int dflt) {
  for (auto& o : g_options)
    if (strcmp(o.name, name) == 0) {
      int v = o.value.load(std::memory_order_relaxed);
      return v ? v : dflt;
    }
  return dflt;
}

// SMs the persistent kernels (prefill GEMM, row quantiser) may occupy: all of them minus option sm_reserve -- set when a
// collective's kernel has to run concurrently (column-parallel over NCCL: its CTAs otherwise queue behind a persistent
// GEMM that owns every SM).
int mq_worker_sms() {
  const int n = mq_sm_count(), r = mq_get_option("sm_reserve", 0);
  return r > 0 && r < n - 2 ? n - r : n;
}

int mq_sm_count() {
  static std::atomic<int> cached[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  int v = cached[dev].load(std::memory_order_relaxed);
  if (v) return v;
  if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
  cached[dev].store(v, std::memory_order_relaxed);
  return v;
}

extern "C" {

int mq_version(void) { return MQ_ABI_VERSION; }

const char* mq_last_error(void) { return g_err; }

int mq_device_info(int* sm_count, int* cc_major, int* cc_minor) {
  int dev = 0;
  MQ_CHECK_CUDA(cudaGetDevice(&dev));
  int v = 0;
  if (sm_count) {
    MQ_CHECK_CUDA(cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev));
    *sm_count = v;
  }
  if (cc_major) {
    MQ_CHECK_CUDA(cudaDeviceGetAttribute(&v, cudaDevAttrComputeCapabilityMajor, dev));
    *cc_major = v;
  }
  if (cc_minor) {
    MQ_CHECK_CUDA(cudaDeviceGetAttribute(&v, cudaDevAttrComputeCapabilityMinor, dev));
    *cc_minor = v;
  }
  return MQ_OK;
}

int mq_set_option(const char* name, int value) {
  if (!name) {
    mq_set_error("mq_set_option: null name");
    return MQ_ERR_ARG;
  }
  for (auto& o : g_options)
    if (strcmp(o.name, name) == 0) {
      o.value.store(value, std::memory_order_relaxed);
      return MQ_OK;
    }
  mq_set_error("mq_set_option: unknown option '%s'", name);
  return MQ_ERR_ARG;
}

}  // extern "C"
