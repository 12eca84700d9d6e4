#include "trace.h"

#include <nvtx3/nvToolsExt.h>
#include <stdio.h>
#include <stdlib.h>

#include "cuda_check.h"
#include "gpu_plane.h"

namespace mp4j {

CollTrace::CollTrace() {
  const char* e = getenv("MP4J_TRACE");
  on_ = e && *e && *e != '0';
}

CollTrace::~CollTrace() {
  for (auto& p : pending_) {
    cudaEventDestroy(p.e0);
    cudaEventDestroy(p.e1);
  }
  for (cudaEvent_t e : pool_) cudaEventDestroy(e);
}

cudaEvent_t CollTrace::get_event() {
  if (!pool_.empty()) {
    cudaEvent_t e = pool_.back();
    pool_.pop_back();
    return e;
  }
  cudaEvent_t e;
  MP4J_CUDA(cudaEventCreate(&e));
  return e;
}

cudaEvent_t CollTrace::begin(cudaStream_t s) {
  std::lock_guard<std::mutex> lk(mu_);
  cudaEvent_t e = get_event();
  MP4J_CUDA(cudaEventRecord(e, s));
  return e;
}

void CollTrace::end(cudaEvent_t e0, cudaStream_t s, const char* name, const char* algo, size_t bytes) {
  std::lock_guard<std::mutex> lk(mu_);
  cudaEvent_t e1 = get_event();
  cudaEventRecord(e1, s);
  pending_.push_back(Pending{std::string(name) + "/" + algo, bytes, e0, e1});
  if (pending_.size() >= 4096) resolve();  // bound the number of live events
}

void CollTrace::resolve() {
  for (auto& p : pending_) {
    float ms = 0.f;
    if (cudaEventSynchronize(p.e1) == cudaSuccess && cudaEventElapsedTime(&ms, p.e0, p.e1) == cudaSuccess) {
      Agg& a = agg_[p.key];
      const double us = (double)ms * 1e3;
      a.calls += 1;
      a.bytes += p.bytes;
      a.us += us;
      a.min_us = us < a.min_us ? us : a.min_us;
      a.max_us = us > a.max_us ? us : a.max_us;
    }
    pool_.push_back(p.e0);
    pool_.push_back(p.e1);
  }
  pending_.clear();
}

std::string CollTrace::stats_json() {
  std::lock_guard<std::mutex> lk(mu_);
  resolve();
  std::string out = "[";
  bool first = true;
  for (const auto& kv : agg_) {
    const Agg& a = kv.second;
    const size_t slash = kv.first.find('/');
    char buf[512];
    snprintf(buf, sizeof buf,
             "%s{\"collective\": \"%s\", \"algo\": \"%s\", \"calls\": %llu, \"bytes\": %llu, \"total_us\": %.2f, "
             "\"min_us\": %.2f, \"max_us\": %.2f, \"avg_us\": %.2f, \"algbw_GBps\": %.2f}",
             first ? "" : ", ", kv.first.substr(0, slash).c_str(), kv.first.substr(slash + 1).c_str(), a.calls, a.bytes,
             a.us, a.min_us, a.max_us, a.us / (double)a.calls, a.us > 0 ? (double)a.bytes / a.us / 1e3 : 0.0);
    out += buf;
    first = false;
  }
  return out + "]";
}

void CollTrace::reset() {
  std::lock_guard<std::mutex> lk(mu_);
  resolve();
  agg_.clear();
}

TraceScope::TraceScope(GpuPlane* g, cudaStream_t s, const char* nvtx_name) : g_(g), s_(s) {
  if (!g_->trace().enabled()) return;
  nvtxRangePushA(nvtx_name);  // no-op unless a tool injected an NVTX library
  nvtx_ = true;
  e0_ = g_->trace().begin(s_);
}

TraceScope::~TraceScope() {
  if (e0_) {
    const CollStat& st = g_->last();
    try {
      g_->trace().end(e0_, s_, st.name, st.algo, st.bytes);
    } catch (...) {
    }
  }
  if (nvtx_) nvtxRangePop();
}

}  // namespace mp4j
