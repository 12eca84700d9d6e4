// extern "C" surface of the GPU data plane (include/mp4j_gpu.h).
#include <string.h>

#include <string>
#include <vector>

#include "../../include/mp4j.h"
#include "../capi/capi_util.h"
#include "../host/process_comm.h"
#include "../host/thread_comm.h"
#include "custom_op.h"
#include "gpu_plane.h"

using namespace mp4j;

struct mp4j_gpu {
  GpuPlane* g;
  bool owned_by_comm;
};

static void gpu_dtor(mp4j_comm* c) {
  auto* h = (mp4j_gpu*)c->gpu;
  if (!h) return;
  delete h->g;
  delete h;
  c->gpu = nullptr;
}

static GpuOptions make_opts(size_t heap_bytes, size_t staging_bytes) {
  GpuOptions o;
  o.load_env();
  if (heap_bytes) o.heap_bytes = heap_bytes;
  if (staging_bytes) o.staging_bytes = staging_bytes;
  return o;
}

static void copy_out(const std::string& s, char* buf, size_t n) {
  if (!n) return;
  size_t k = s.size() < n - 1 ? s.size() : n - 1;
  memcpy(buf, s.data(), k);
  buf[k] = 0;
}

extern "C" {

int mp4j_gpu_init(mp4j_comm* c, int device, size_t heap_bytes, size_t staging_bytes, mp4j_gpu** out) {
  MP4J_API_BEGIN
  if (!c->gpu) {
    auto* h = new mp4j_gpu{nullptr, true};
    try {
      h->g = GpuPlane::create_real(c->pc, device, make_opts(heap_bytes, staging_bytes));
    } catch (...) {
      delete h;
      throw;
    }
    c->gpu = h;
    c->gpu_dtor = gpu_dtor;
  }
  *out = (mp4j_gpu*)c->gpu;
  MP4J_API_END
}

int mp4j_gpu_emu_create(int world, int device, size_t heap_bytes, size_t staging_bytes, mp4j_gpu** out) {
  MP4J_API_BEGIN
  auto* h = new mp4j_gpu{nullptr, false};
  try {
    h->g = GpuPlane::create_emulated(world, device, make_opts(heap_bytes, staging_bytes));
  } catch (...) {
    delete h;
    throw;
  }
  *out = h;
  MP4J_API_END
}

void mp4j_gpu_destroy(mp4j_gpu* g) {
  if (!g || g->owned_by_comm) return;
  delete g->g;
  delete g;
}

int mp4j_gpu_world(mp4j_gpu* g) { return g->g->world(); }
int mp4j_gpu_rank(mp4j_gpu* g) { return g->g->rank(); }
int mp4j_gpu_has_multicast(mp4j_gpu* g) { return g->g->has_multicast() ? 1 : 0; }

int mp4j_gpu_describe(mp4j_gpu* g, char* buf, size_t n) {
  MP4J_API_BEGIN
  copy_out(g->g->describe(), buf, n);
  MP4J_API_END
}

int mp4j_gpu_last(mp4j_gpu* g, char* buf, size_t n) {
  MP4J_API_BEGIN
  const CollStat& s = g->g->last();
  copy_out(std::string(s.name) + " algo=" + s.algo + " bytes=" + std::to_string(s.bytes) +
               " grid=" + std::to_string(s.grid) + " staged=" + (s.staged ? "1" : "0"),
           buf, n);
  MP4J_API_END
}

int mp4j_gpu_set_option(mp4j_gpu* g, const char* name, long long value) {
  MP4J_API_BEGIN
  if (!g->g->set_option(name, value)) throw Error(std::string("unknown GPU plane option: ") + name);
  MP4J_API_END
}

int mp4j_gpu_alloc(mp4j_gpu* g, size_t bytes, size_t* off) {
  MP4J_API_BEGIN
  *off = g->g->alloc(bytes);
  MP4J_API_END
}
int mp4j_gpu_free(mp4j_gpu* g, size_t off) {
  MP4J_API_BEGIN
  g->g->free(off);
  MP4J_API_END
}
void* mp4j_gpu_ptr(mp4j_gpu* g, int rank, size_t off) { return g->g->ptr(rank, off); }
int mp4j_gpu_check(mp4j_gpu* g) {
  MP4J_API_BEGIN
  g->g->check_error();
  MP4J_API_END
}

int mp4j_gpu_allreduce(mp4j_gpu* g, void* const* arrs, int dtype, int op, int64_t from, int64_t to, int algo,
                       void* stream) {
  MP4J_API_BEGIN
  TraceScope trace_scope(g->g, (cudaStream_t)stream, "mp4j.allreduce");
  g->g->allreduce(arrs, dtype, op, from, to, algo, (cudaStream_t)stream);
  MP4J_API_END
}
int mp4j_gpu_trace(mp4j_gpu* g, int on) {
  MP4J_API_BEGIN
  g->g->trace().enable(on != 0);
  MP4J_API_END
}

int mp4j_gpu_stats(mp4j_gpu* g, char* buf, size_t n, int reset) {
  MP4J_API_BEGIN
  copy_out(g->g->trace().stats_json(), buf, n);
  if (reset) g->g->trace().reset();
  MP4J_API_END
}

int mp4j_gpu_custom_op(mp4j_gpu* g, int dtype, const char* expr, int* op_out) {
  MP4J_API_BEGIN
  *op_out = custom_op_register(dtype, expr, g->g->world());
  MP4J_API_END
}

int mp4j_rtc_compile(int dtype, const char* expr, int world, long long* cubin_bytes) {
  MP4J_API_BEGIN
  size_t n = 0;
  custom_op_compile_only(dtype, expr, world, &n);
  if (cubin_bytes) *cubin_bytes = (long long)n;
  MP4J_API_END
}

int mp4j_gpu_allreduce_wire(mp4j_gpu* g, void* const* arrs, int64_t from, int64_t to, int wire, void* stream) {
  MP4J_API_BEGIN
  TraceScope trace_scope(g->g, (cudaStream_t)stream, "mp4j.allreduce_wire");
  g->g->allreduce_wire((float* const*)arrs, from, to, wire, (cudaStream_t)stream);
  MP4J_API_END
}
int mp4j_gpu_reduce_scatter(mp4j_gpu* g, void* const* arrs, int dtype, int op, int64_t from,
                            const int64_t* counts, int algo, void* stream) {
  MP4J_API_BEGIN
  TraceScope trace_scope(g->g, (cudaStream_t)stream, "mp4j.reduce_scatter");
  g->g->reduce_scatter(arrs, dtype, op, from, counts, algo, (cudaStream_t)stream);
  MP4J_API_END
}
int mp4j_gpu_allgather(mp4j_gpu* g, void* const* arrs, int dtype, const int64_t* froms, const int64_t* tos,
                       int algo, void* stream) {
  MP4J_API_BEGIN
  TraceScope trace_scope(g->g, (cudaStream_t)stream, "mp4j.allgather");
  g->g->allgather(arrs, dtype, froms, tos, algo, (cudaStream_t)stream);
  MP4J_API_END
}
int mp4j_gpu_broadcast(mp4j_gpu* g, void* const* arrs, int dtype, int64_t from, int64_t to, int root, int algo,
                       void* stream) {
  MP4J_API_BEGIN
  TraceScope trace_scope(g->g, (cudaStream_t)stream, "mp4j.broadcast");
  g->g->broadcast(arrs, dtype, from, to, root, algo, (cudaStream_t)stream);
  MP4J_API_END
}
int mp4j_gpu_gather(mp4j_gpu* g, void* const* arrs, int dtype, const int64_t* froms, const int64_t* tos,
                    int root, void* stream) {
  MP4J_API_BEGIN
  TraceScope trace_scope(g->g, (cudaStream_t)stream, "mp4j.gather");
  g->g->gather(arrs, dtype, froms, tos, root, (cudaStream_t)stream);
  MP4J_API_END
}
int mp4j_gpu_scatter(mp4j_gpu* g, void* const* arrs, int dtype, const int64_t* froms, const int64_t* tos,
                     int root, void* stream) {
  MP4J_API_BEGIN
  TraceScope trace_scope(g->g, (cudaStream_t)stream, "mp4j.scatter");
  g->g->scatter(arrs, dtype, froms, tos, root, (cudaStream_t)stream);
  MP4J_API_END
}
int mp4j_gpu_reduce(mp4j_gpu* g, void* const* arrs, int dtype, int op, int64_t from, int64_t to, int root,
                    int algo, void* stream) {
  MP4J_API_BEGIN
  TraceScope trace_scope(g->g, (cudaStream_t)stream, "mp4j.reduce");
  g->g->reduce(arrs, dtype, op, from, to, root, algo, (cudaStream_t)stream);
  MP4J_API_END
}
int mp4j_gpu_thread_allreduce(mp4j_gpu* g, void* const* bufs, int T, int dtype, int op, int64_t from,
                              int64_t to, int fanout, int dst, int local_only, int thread_algo, int algo,
                              void* stream) {
  MP4J_API_BEGIN
  TraceScope trace_scope(g->g, (cudaStream_t)stream, "mp4j.thread_allreduce");
  g->g->thread_allreduce(bufs, T, dtype, op, from, to, fanout, dst, local_only != 0,
                         (thread_algo & 0xff) | (algo << 16), (cudaStream_t)stream);
  MP4J_API_END
}

// ---- ThreadCommSlave on device (or pinned host) buffers, called by ALL T threads of a process with their own
// buffer: the rendezvous, the election of the launching thread and the propagation of its error to every thread
// happen natively (ThreadComm::delegate), so that every binding gets them from the C ABI instead of rebuilding them.
static std::vector<void*> thread_bufs(const std::vector<uint64_t>& args) {
  std::vector<void*> bufs(args.size());
  for (size_t t = 0; t < args.size(); ++t) {
    bufs[t] = (void*)(uintptr_t)args[t];
    // a binding whose own argument check failed on one thread still joins the rendezvous (with no buffer), so
    // that the other T - 1 threads get an error instead of waiting for it forever
    if (!bufs[t]) throw Error("thread " + std::to_string(t) + " passed no usable buffer to a thread-level collective");
  }
  return bufs;
}

int mp4j_thread_gpu_allreduce(mp4j_tgroup* tg, mp4j_gpu* g, int tid, void* buf, int dtype, int op, int64_t from,
                              int64_t to, int fanout, int dst, int local_only, int thread_algo, int algo,
                              void* stream) {
  MP4J_API_BEGIN
  tg->tc->delegate(tid, 0, (uint64_t)(uintptr_t)buf, [&](const std::vector<uint64_t>& args) {
    std::vector<void*> bufs = thread_bufs(args);
    if (mp4j_gpu_thread_allreduce(g, bufs.data(), (int)bufs.size(), dtype, op, from, to, fanout, dst, local_only,
                                  thread_algo, algo, stream) != 0)
      throw Error(g_last_error);
  });
  MP4J_API_END
}

int mp4j_thread_gpu_allreduce_host(mp4j_tgroup* tg, mp4j_gpu* g, int tid, void* hbuf, int dtype, int op, int64_t from,
                                   int64_t to, int fanout, int dst, void* stream) {
  MP4J_API_BEGIN
  tg->tc->delegate(tid, 0, (uint64_t)(uintptr_t)hbuf, [&](const std::vector<uint64_t>& args) {
    std::vector<void*> bufs = thread_bufs(args);
    if (mp4j_gpu_thread_allreduce_host(g, bufs.data(), (int)bufs.size(), dtype, op, from, to, fanout, dst, stream) != 0)
      throw Error(g_last_error);
  });
  MP4J_API_END
}

int mp4j_thread_gpu_segments(mp4j_tgroup* tg, mp4j_gpu* g, int tid, int launch_tid, void* buf, int dtype, int op,
                             const int64_t* segs, int nseg, const char* name, void* stream) {
  MP4J_API_BEGIN
  tg->tc->delegate(tid, launch_tid, (uint64_t)(uintptr_t)buf, [&](const std::vector<uint64_t>& args) {
    std::vector<void*> bufs = thread_bufs(args);
    if (mp4j_gpu_thread_segments(g, bufs.data(), (int)bufs.size(), dtype, op, segs, nseg, name, stream) != 0)
      throw Error(g_last_error);
  });
  MP4J_API_END
}

int mp4j_gpu_thread_segments(mp4j_gpu* g, void* const* bufs, int T, int dtype, int op, const int64_t* segs, int nseg,
                             const char* name, void* stream) {
  MP4J_API_BEGIN
  // the name is kept by the plane for last(): map it onto static strings
  static const char* const kNames[] = {"thread_gather", "thread_allgather", "thread_broadcast", "thread_scatter",
                                       "thread_reduce_scatter", "thread_reduce"};
  const char* nm = "thread_segments";
  for (const char* k : kNames)
    if (name && strcmp(name, k) == 0) nm = k;
  TraceScope trace_scope(g->g, (cudaStream_t)stream, "mp4j.thread_segments");
  g->g->thread_segments(bufs, T, dtype, op, segs, nseg, nm, (cudaStream_t)stream);
  MP4J_API_END
}

int mp4j_gpu_thread_allreduce_host(mp4j_gpu* g, void* const* hbufs, int T, int dtype, int op, int64_t from,
                                   int64_t to, int fanout, int dst, void* stream) {
  MP4J_API_BEGIN
  g->g->thread_allreduce_host(hbufs, T, dtype, op, from, to, fanout, dst, (cudaStream_t)stream);
  MP4J_API_END
}

int mp4j_gpu_map_collective(mp4j_gpu* g, int mode, const int64_t* const* keys, const void* const* vals,
                            const int64_t* n, int64_t max_n, int64_t out_cap, int dtype, int op, int partition,
                            int64_t** out_keys, void** out_vals, uint64_t** info, void* stream) {
  MP4J_API_BEGIN
  const int L = g->g->emulated() ? g->g->world() : 1;
  long long nn[MP4J_MAX_RANKS];
  for (int i = 0; i < L; ++i) nn[i] = n[i];
  TraceScope trace_scope(g->g, (cudaStream_t)stream, "mp4j.map_collective");
  g->g->map_collective(mode, (const long long* const*)keys, vals, nn, max_n, out_cap, dtype, op, partition,
                       (long long**)out_keys, out_vals, (unsigned long long**)info, (cudaStream_t)stream);
  MP4J_API_END
}

int mp4j_gpu_lr_xtr_reduce_scatter(mp4j_gpu* g, const void* const* xt, const void* const* rt, int F, int C, int64_t K,
                                   float** shard_out, int* feats_per_rank, void* stream) {
  MP4J_API_BEGIN
  TraceScope trace_scope(g->g, (cudaStream_t)stream, "mp4j.lr_xtr_reduce_scatter");
  g->g->lr_xtr_reduce_scatter(xt, rt, F, C, K, shard_out, feats_per_rank, (cudaStream_t)stream);
  MP4J_API_END
}

int mp4j_gpu_gbdt_hist_reduce_scatter(mp4j_gpu* g, const uint8_t* const* bins, const int64_t* pitch,
                                      const float* const* grad, const float* const* hess, const int64_t* rows, int F,
                                      float** shard_out, int* feats_per_rank, void* stream) {
  MP4J_API_BEGIN
  const int L = g->g->emulated() ? g->g->world() : 1;
  long long pp[MP4J_MAX_RANKS], rr[MP4J_MAX_RANKS];
  for (int i = 0; i < L; ++i) {
    pp[i] = pitch[i];
    rr[i] = rows[i];
  }
  TraceScope trace_scope(g->g, (cudaStream_t)stream, "mp4j.gbdt_hist_reduce_scatter");
  g->g->gbdt_hist_reduce_scatter(bins, pp, grad, hess, rr, F, shard_out, feats_per_rank, (cudaStream_t)stream);
  MP4J_API_END
}

int mp4j_gpu_map_reduce_local(mp4j_gpu* g, int local_rank, const int64_t* keys, const void* vals, int64_t n,
                              int dtype, int op, int64_t* out_keys, void* out_vals, int64_t out_cap,
                              uint64_t* out_count, void* stream) {
  MP4J_API_BEGIN
  g->g->map_reduce_local(local_rank, (const long long*)keys, vals, n, dtype, op, (long long*)out_keys, out_vals,
                         out_cap, (unsigned long long*)out_count, (cudaStream_t)stream);
  MP4J_API_END
}

}  // extern "C"
