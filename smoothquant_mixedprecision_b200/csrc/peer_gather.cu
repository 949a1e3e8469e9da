// All-gather of the column blocks of a column-parallel layer over peer-mapped ("symmetric") memory, without SMs.
//
// No reference counterpart (the reference is single-GPU).  What it replaces inside this package is the
// NCCL all-gather of parallel.gather_columns_blocked: NCCL's gather is a kernel, and the prefill GEMM is a persistent
// kernel that owns every SM, so the two serialise (profiles/r02_tp2_nccl.json: 56 of 61 ms of communication exposed
// at tp2).  Here the transfer is done by the copy engines over NVLink and the cross-GPU ordering by stream memory
// operations (cuStreamWriteValue32 / cuStreamWaitValue32) on flag words in peer-mapped memory -- nothing in this file
// launches a kernel, so a gather runs underneath the GEMMs of the other row chunk.
//
// Memory (set up by the host side, parallel.PeerGather, through torch's symmetric-memory allocator = cuMem mappings
// with peer access):
//   data   every rank holds a buffer [world][M][S]; rank r's GEMM writes block r of ITS OWN buffer
//   flags  every rank holds uint32 ready[world], pulled[world]
//            ready[p]  (written by rank p)  = sequence number of the last gather whose block p is complete in p's buffer
//            pulled[p] (written by rank p)  = sequence number of the last gather whose block p has copied out of mine
// Protocol of gather number seq on rank r, all on one stream (ordered after r's GEMM by the caller):
//   1. write seq to ready[r] of every peer
//   2. for every peer p: wait until my ready[p] >= seq, copy block p of p's buffer into block p of mine (DMA over
//      NVLink), write seq to pulled[r] of p
// and before r's GEMM overwrites block r of a buffer last used by gather `seq`, mq_peer_wait_pulled makes its stream
// wait until pulled[p] >= seq for every peer.  Sequence numbers are compared cyclically (GEQ of the driver).
#include <cuda.h>

#include "common.cuh"

namespace {

typedef CUresult (*MemOp32Fn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);

MemOp32Fn entry(const char* name) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
    return nullptr;
  return reinterpret_cast<MemOp32Fn>(p);
}
MemOp32Fn write32() {
  static MemOp32Fn fn = entry("cuStreamWriteValue32");
  return fn;
}
MemOp32Fn wait32() {
  static MemOp32Fn fn = entry("cuStreamWaitValue32");
  return fn;
}

#define MQ_CHECK_DRV(expr)                                                          \
  do {                                                                              \
    CUresult r__ = (expr);                                                          \
    if (r__ != CUDA_SUCCESS) {                                                      \
      mq_set_error("%s failed with CUresult %d (%s:%d)", #expr, (int)r__, __FILE__, __LINE__); \
      return MQ_ERR_CUDA;                                                           \
    }                                                                               \
  } while (0)

int check_common(const char* fn, const uint64_t* flags, int rank, int world) {
  MQ_CHECK_ARG(flags, "%s: null flag table", fn);
  MQ_CHECK_ARG(world >= 1 && world <= 64 && rank >= 0 && rank < world, "%s: rank %d of %d", fn, rank, world);
  if (!write32() || !wait32()) {
    mq_set_error("%s: cuStreamWriteValue32 / cuStreamWaitValue32 entry points not available", fn);
    return MQ_ERR_UNSUPPORTED;
  }
  return MQ_OK;
}

}  // namespace

extern "C" {

int mq_peer_gather(const uint64_t* bufs, const uint64_t* flags, int rank, int world, size_t block_bytes, uint32_t seq,
                   const uint64_t* streams, int n_streams) {
  int rc = check_common("mq_peer_gather", flags, rank, world);
  if (rc) return rc;
  MQ_CHECK_ARG(bufs && block_bytes > 0 && block_bytes % 16 == 0, "mq_peer_gather: bad buffer table / block size");
  MQ_CHECK_ARG(streams && n_streams >= 1 && n_streams <= 16, "mq_peer_gather: 1..16 streams, got %d", n_streams);
  const uint64_t w4 = (uint64_t)world * 4;
  CUstream st0 = reinterpret_cast<CUstream>(streams[0]);
  for (int s = 1; s < world; ++s) {                       // ready[rank] of every peer
    const int p = (rank + s) % world;
    MQ_CHECK_DRV(write32()(st0, (CUdeviceptr)(flags[p] + 4ull * rank), seq, CU_STREAM_WRITE_VALUE_DEFAULT));
  }
  for (int s = 1; s < world; ++s) {                       // staggered: rank r starts with its left neighbour
    const int p = (rank + world - s) % world;
    CUstream st = reinterpret_cast<CUstream>(streams[(s - 1) % n_streams]);   // lanes = copy engines in parallel
    MQ_CHECK_DRV(wait32()(st, (CUdeviceptr)(flags[rank] + 4ull * p), seq, CU_STREAM_WAIT_VALUE_GEQ));
    const uint64_t off = (uint64_t)p * block_bytes;
    MQ_CHECK_CUDA(cudaMemcpyAsync(reinterpret_cast<void*>(bufs[rank] + off), reinterpret_cast<const void*>(bufs[p] + off),
                                  block_bytes, cudaMemcpyDeviceToDevice, reinterpret_cast<cudaStream_t>(st)));
    MQ_CHECK_DRV(write32()(st, (CUdeviceptr)(flags[p] + w4 + 4ull * rank), seq, CU_STREAM_WRITE_VALUE_DEFAULT));
  }
  return MQ_OK;
}

// ---- push variant ------------------------------------------------------------------------------------------
// Flags (uint32 words of every rank's flag block): [0, w) ready, [w, 2w) pulled (pull protocol above);
// [2w, 3w) arrived[q] = last gather whose block q has landed in MY buffer (written by q after its copy; q's copies
// run in stream order, so the word is monotonic),
// [(4 + id) w, (5 + id) w) consumed[id][q] = last gather OF RING BUFFER `id` that q has finished reading (written by q
// from its compute stream).  One word per ring buffer: the gathers of one buffer are consumed in order, those of
// different buffers need not be, and a single word per peer would not be monotonic.
// Gather seq on rank r (`prev` = the gather that last used this ring buffer, 0 if none):
//   lanes:   wait consumed[p] >= prev for every peer p, copy my block into block r of every peer's buffer -- or ONE copy
//            to the multicast address of the buffers when mc_buf != 0 (the switch replicates it) -- then write
//            arrived[r] = seq into every peer's flag block
//   consumer (mq_peer_wait_arrived, on the stream that reads the gathered buffer): wait arrived[q] >= seq for all q
//   before a rank reuses a ring buffer it publishes consumed[r] = prev to every peer from its compute stream
//   (mq_peer_publish_consumed): everything queued on that stream before -- the readers of gather prev -- is complete.
int mq_peer_push(const uint64_t* bufs, const uint64_t* flags, uint64_t mc_buf, int rank, int world, size_t block_bytes,
                 uint32_t seq, uint32_t prev, int buf_id, const uint64_t* streams, int n_streams) {
  int rc = check_common("mq_peer_push", flags, rank, world);
  if (rc) return rc;
  MQ_CHECK_ARG(bufs && block_bytes > 0 && block_bytes % 16 == 0, "mq_peer_push: bad buffer table / block size");
  MQ_CHECK_ARG(streams && n_streams >= 1 && n_streams <= 16, "mq_peer_push: 1..16 streams, got %d", n_streams);
  MQ_CHECK_ARG(buf_id >= 0 && buf_id < 1024, "mq_peer_push: ring buffer id %d", buf_id);
  const uint64_t w4 = (uint64_t)world * 4, off = (uint64_t)rank * block_bytes;
  const uint64_t cons = (uint64_t)(4 + buf_id) * w4;      // byte offset of consumed[buf_id][*] in a flag block
  const void* src = reinterpret_cast<const void*>(bufs[rank] + off);
  if (mc_buf) {
    CUstream st = reinterpret_cast<CUstream>(streams[0]);
    if (prev)
      for (int s = 1; s < world; ++s) {
        const int p = (rank + s) % world;
        MQ_CHECK_DRV(wait32()(st, (CUdeviceptr)(flags[rank] + cons + 4ull * p), prev, CU_STREAM_WAIT_VALUE_GEQ));
      }
    MQ_CHECK_CUDA(cudaMemcpyAsync(reinterpret_cast<void*>(mc_buf + off), src, block_bytes, cudaMemcpyDeviceToDevice,
                                  reinterpret_cast<cudaStream_t>(st)));
    for (int s = 1; s < world; ++s) {
      const int p = (rank + s) % world;
      MQ_CHECK_DRV(write32()(st, (CUdeviceptr)(flags[p] + 2 * w4 + 4ull * rank), seq, CU_STREAM_WRITE_VALUE_DEFAULT));
    }
    return MQ_OK;
  }
  for (int s = 1; s < world; ++s) {
    const int p = (rank + s) % world;                     // staggered: rank r starts with its right neighbour
    CUstream st = reinterpret_cast<CUstream>(streams[(s - 1) % n_streams]);
    if (prev) MQ_CHECK_DRV(wait32()(st, (CUdeviceptr)(flags[rank] + cons + 4ull * p), prev, CU_STREAM_WAIT_VALUE_GEQ));
    MQ_CHECK_CUDA(cudaMemcpyAsync(reinterpret_cast<void*>(bufs[p] + off), src, block_bytes, cudaMemcpyDeviceToDevice,
                                  reinterpret_cast<cudaStream_t>(st)));
    MQ_CHECK_DRV(write32()(st, (CUdeviceptr)(flags[p] + 2 * w4 + 4ull * rank), seq, CU_STREAM_WRITE_VALUE_DEFAULT));
  }
  return MQ_OK;
}

int mq_peer_wait_arrived(const uint64_t* flags, int rank, int world, uint32_t seq, void* stream) {
  int rc = check_common("mq_peer_wait_arrived", flags, rank, world);
  if (rc) return rc;
  CUstream st = static_cast<CUstream>(stream);
  const uint64_t w4 = (uint64_t)world * 4;
  for (int s = 1; s < world; ++s) {
    const int p = (rank + s) % world;
    MQ_CHECK_DRV(wait32()(st, (CUdeviceptr)(flags[rank] + 2 * w4 + 4ull * p), seq, CU_STREAM_WAIT_VALUE_GEQ));
  }
  return MQ_OK;
}

int mq_peer_publish_consumed(const uint64_t* flags, int rank, int world, uint32_t seq, int buf_id, void* stream) {
  int rc = check_common("mq_peer_publish_consumed", flags, rank, world);
  if (rc) return rc;
  MQ_CHECK_ARG(buf_id >= 0 && buf_id < 1024, "mq_peer_publish_consumed: ring buffer id %d", buf_id);
  CUstream st = static_cast<CUstream>(stream);
  const uint64_t w4 = (uint64_t)world * 4, cons = (uint64_t)(4 + buf_id) * w4;
  for (int s = 1; s < world; ++s) {
    const int p = (rank + s) % world;
    MQ_CHECK_DRV(write32()(st, (CUdeviceptr)(flags[p] + cons + 4ull * rank), seq, CU_STREAM_WRITE_VALUE_DEFAULT));
  }
  return MQ_OK;
}

int mq_peer_wait_pulled(const uint64_t* flags, int rank, int world, uint32_t seq, void* stream) {
  int rc = check_common("mq_peer_wait_pulled", flags, rank, world);
  if (rc) return rc;
  CUstream st = static_cast<CUstream>(stream);
  const uint64_t w4 = (uint64_t)world * 4;
  for (int s = 1; s < world; ++s) {
    const int p = (rank + s) % world;
    MQ_CHECK_DRV(wait32()(st, (CUdeviceptr)(flags[rank] + w4 + 4ull * p), seq, CU_STREAM_WAIT_VALUE_GEQ));
  }
  return MQ_OK;
}

}  // extern "C"
