// Host arrays through the GPU plane: chunked H2D || fused kernel || D2H pipeline owned by the library.
//
// The reference's collectives take host (Java heap) arrays (PCS:1733, TCS:1915). The B200 equivalent
// of "array in, result in the same array" is a pinned host tensor that is streamed through the GPU:
// PCIe is the bound (2 GiB in per step for 8 x 64 Mi floats), so everything else must hide behind it.
// One thread of the process issues the whole schedule asynchronously:
//
//   copy-in stream : for chunk c: wait until buffer c % 3 is free, T x cudaMemcpyAsync H2D, event
//   kernel stream  : wait copy-in(c); fused thread allreduce of the T staged chunks into staged buffer 0
//                    (fanout off: the fan-out to the T host arrays is done by the copy engine); event
//   copy-out stream: wait kernel(c); cudaMemcpyAsync D2H of the result chunk into every thread's host
//                    array (or only the destination thread's); event (frees the buffer)
//
// PCIe is full duplex and the GPU has separate copy engines per direction, so in steady state the
// step costs max(H2D bytes, D2H bytes) / PCIe rate; the kernels (tens of microseconds per chunk) and the
// first/last chunk are the only exposed parts.
#include <algorithm>

#include "../core/partition.h"
#include "cuda_check.h"
#include "gpu_plane.h"

namespace mp4j {

namespace {
constexpr int kHpBufs = 3;
size_t align_down(size_t v, size_t a) { return v / a * a; }
}  // namespace

struct GpuPlane::HostPipe {
  cudaStream_t s_in = nullptr, s_k = nullptr, s_out = nullptr;
  cudaEvent_t ev_in[kHpBufs] = {}, ev_k[kHpBufs] = {}, ev_out[kHpBufs] = {}, ev_entry = nullptr;
  char* dev = nullptr;  // kHpBufs x (L x T) x chunk_bytes
  size_t dev_bytes = 0;
  size_t chunk_bytes = 0;
  int LT = 0;
};

void GpuPlane::hpipe_release() {
  if (!hpipe_) return;
  HostPipe* h = hpipe_;
  if (h->dev) cudaFree(h->dev);
  for (int i = 0; i < kHpBufs; ++i) {
    if (h->ev_in[i]) cudaEventDestroy(h->ev_in[i]);
    if (h->ev_k[i]) cudaEventDestroy(h->ev_k[i]);
    if (h->ev_out[i]) cudaEventDestroy(h->ev_out[i]);
  }
  if (h->ev_entry) cudaEventDestroy(h->ev_entry);
  if (h->s_in) cudaStreamDestroy(h->s_in);
  if (h->s_k) cudaStreamDestroy(h->s_k);
  if (h->s_out) cudaStreamDestroy(h->s_out);
  delete h;
  hpipe_ = nullptr;
}

void GpuPlane::thread_allreduce_host(void* const* hbufs, int T, int dtype, int op, int64_t from, int64_t to,
                                     int fanout, int dst, cudaStream_t s) {
  check_from_to(from, to);
  if (T < 1 || T > MP4J_MAX_THREADS) throw Error("thread count out of range");
  if (dst < 0 || dst >= T) throw Error("dst thread out of range");
  if (!plane_op_ok(dtype, op) || op == MP4J_COPY) throw Error("operator is not defined for this dtype");
  const int p = world();
  const int L = local_ranks();
  if (from == to) return;
  if (T == 1 && p == 1) return;  // nothing to combine
  const size_t es = dtype_size(dtype);
  MP4J_CUDA(cudaSetDevice(device_));

  // ---- set-up (first call, or a different T): streams, events, staging
  const int LT = L * T;
  // per-buffer chunk: the configured size, but the whole staging area stays under 384 MiB
  size_t cb = std::min<size_t>(opt_.host_chunk_bytes, (384ull << 20) / ((size_t)kHpBufs * LT));
  cb = std::max<size_t>(align_down(cb, 4096), 4096);
  if (!hpipe_) hpipe_ = new HostPipe();
  HostPipe* h = hpipe_;
  if (!h->s_in) {
    MP4J_CUDA(cudaStreamCreateWithFlags(&h->s_in, cudaStreamNonBlocking));
    MP4J_CUDA(cudaStreamCreateWithFlags(&h->s_k, cudaStreamNonBlocking));
    MP4J_CUDA(cudaStreamCreateWithFlags(&h->s_out, cudaStreamNonBlocking));
    for (int i = 0; i < kHpBufs; ++i) {
      MP4J_CUDA(cudaEventCreateWithFlags(&h->ev_in[i], cudaEventDisableTiming));
      MP4J_CUDA(cudaEventCreateWithFlags(&h->ev_k[i], cudaEventDisableTiming));
      MP4J_CUDA(cudaEventCreateWithFlags(&h->ev_out[i], cudaEventDisableTiming));
    }
    MP4J_CUDA(cudaEventCreateWithFlags(&h->ev_entry, cudaEventDisableTiming));
  }
  const size_t need = (size_t)kHpBufs * LT * cb;
  if (h->dev_bytes < need || h->chunk_bytes != cb || h->LT != LT) {
    MP4J_CUDA(cudaStreamSynchronize(h->s_out));
    if (h->dev) MP4J_CUDA(cudaFree(h->dev));
    h->dev = nullptr;
    MP4J_CUDA(cudaMalloc((void**)&h->dev, need));
    h->dev_bytes = need;
    h->chunk_bytes = cb;
    h->LT = LT;
  }
  auto dbuf = [&](int b, int i, int t) { return h->dev + ((size_t)b * LT + (size_t)i * T + t) * cb; };

  // everything the caller enqueued on `s` precedes the pipeline
  MP4J_CUDA(cudaEventRecord(h->ev_entry, s));
  MP4J_CUDA(cudaStreamWaitEvent(h->s_in, h->ev_entry, 0));
  MP4J_CUDA(cudaStreamWaitEvent(h->s_out, h->ev_entry, 0));

  const int64_t ce = (int64_t)(cb / es);  // elements per chunk
  const int64_t n = to - from;
  const int64_t nchunks = (n + ce - 1) / ce;
  void* ptrs[kMaxRanks * MP4J_MAX_THREADS];
  for (int64_t c = 0; c < nchunks; ++c) {
    const int b = (int)(c % kHpBufs);
    const int64_t e0 = from + c * ce;
    const int64_t ne = std::min<int64_t>(ce, to - e0);
    const size_t bytes = (size_t)ne * es;
    // ---- copy in
    if (c >= kHpBufs) MP4J_CUDA(cudaStreamWaitEvent(h->s_in, h->ev_out[b], 0));
    for (int i = 0; i < L; ++i)
      for (int t = 0; t < T; ++t)
        MP4J_CUDA(cudaMemcpyAsync(dbuf(b, i, t), (const char*)hbufs[(size_t)i * T + t] + (size_t)e0 * es, bytes,
                                  cudaMemcpyHostToDevice, h->s_in));
    MP4J_CUDA(cudaEventRecord(h->ev_in[b], h->s_in));
    // ---- combine on the device: result in staged buffer (i, 0)
    MP4J_CUDA(cudaStreamWaitEvent(h->s_k, h->ev_in[b], 0));
    if (T == 1) {
      for (int i = 0; i < L; ++i) ptrs[i] = dbuf(b, i, 0);
      allreduce(ptrs, dtype, op, 0, ne, MP4J_ALGO_AUTO, h->s_k);
    } else {
      for (int i = 0; i < LT; ++i) ptrs[i] = dbuf(b, i / T, i % T);
      thread_allreduce(ptrs, T, dtype, op, 0, ne, /*fanout=*/0, /*dst=*/0, /*local_only=*/p == 1, 0, h->s_k);
    }
    MP4J_CUDA(cudaEventRecord(h->ev_k[b], h->s_k));
    // ---- copy out: the copy engine fans the result out to the host arrays
    MP4J_CUDA(cudaStreamWaitEvent(h->s_out, h->ev_k[b], 0));
    for (int i = 0; i < L; ++i)
      for (int t = 0; t < T; ++t)
        if (fanout || t == dst)
          MP4J_CUDA(cudaMemcpyAsync((char*)hbufs[(size_t)i * T + t] + (size_t)e0 * es, dbuf(b, i, 0), bytes,
                                    cudaMemcpyDeviceToHost, h->s_out));
    MP4J_CUDA(cudaEventRecord(h->ev_out[b], h->s_out));
  }
  const char* kalgo = last_.algo;
  MP4J_CUDA(cudaStreamSynchronize(h->s_out));
  check_error();
  last_algo_store_ = std::string("host-pipeline[") + std::to_string(nchunks) + "x" + std::to_string(cb) + "B]+" + kalgo;
  last_ = CollStat{"thread_allreduce_host", last_algo_store_.c_str(), (size_t)n * es, last_.grid, false};
}

}  // namespace mp4j
