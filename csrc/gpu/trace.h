// Per-collective timing and NVTX ranges for the GPU data plane (the reference
// only logs wall-clock deltas from its check classes, J/check/checkdouble/ProcessAllReduceCheck.java:64-66).
//
// Off by default (MP4J_TRACE=1 or mp4j_gpu_trace(g, 1) turns it on): every collective entry point is
// then bracketed by a CUDA event pair on the caller's stream and an NVTX range; the pairs are
// resolved lazily when the statistics are read, so tracing never synchronises the stream itself.
#pragma once
#include <cuda_runtime.h>

#include <map>
#include <mutex>
#include <string>
#include <vector>

namespace mp4j {

class GpuPlane;

class CollTrace {
 public:
  CollTrace();
  ~CollTrace();
  void enable(bool on) { on_ = on; }
  bool enabled() const { return on_; }
  // called by TraceScope
  cudaEvent_t begin(cudaStream_t s);
  void end(cudaEvent_t e0, cudaStream_t s, const char* name, const char* algo, size_t bytes);
  // JSON array, one object per (collective, algorithm): calls, bytes, total/min/max us, GB/s (algorithmic)
  std::string stats_json();
  void reset();

 private:
  struct Pending {
    std::string key;
    size_t bytes;
    cudaEvent_t e0, e1;
  };
  struct Agg {
    unsigned long long calls = 0, bytes = 0;
    double us = 0.0, min_us = 1e300, max_us = 0.0;
  };
  cudaEvent_t get_event();
  void resolve();
  bool on_ = false;
  std::mutex mu_;
  std::vector<cudaEvent_t> pool_;
  std::vector<Pending> pending_;
  std::map<std::string, Agg> agg_;
};

// RAII bracket used by the extern "C" layer around every collective call.
class TraceScope {
 public:
  TraceScope(GpuPlane* g, cudaStream_t s, const char* nvtx_name);
  ~TraceScope();

 private:
  GpuPlane* g_;
  cudaStream_t s_;
  cudaEvent_t e0_ = nullptr;
  bool nvtx_ = false;
};

}  // namespace mp4j
