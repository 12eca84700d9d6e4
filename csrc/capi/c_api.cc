// extern "C" surface over the host core (include/mp4j.h). The GPU half lives in csrc/gpu/gpu_api.cu.
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "../../include/mp4j.h"
#include "../control/control.h"
#include "../host/process_comm.h"
#include "../host/thread_comm.h"
#include "capi_util.h"

using namespace mp4j;

namespace mp4j {
thread_local std::string g_last_error;
}

struct mp4j_master {
  Master* m;
};

extern "C" {

const char* mp4j_last_error(void) { return g_last_error.c_str(); }
const char* mp4j_version(void) { return "mp4j-b200 0.1.0 (sm_100a)"; }
void mp4j_free(void* p) { free(p); }

static char* dup_bytes(const std::string& s) {
  char* p = (char*)malloc(s.size() + 1);
  if (!p) throw Error("out of memory");
  memcpy(p, s.data(), s.size());
  p[s.size()] = 0;
  return p;
}

// ---------------------------------------------------------------- master
int mp4j_master_create(int nslaves, int port, int connect_timeout_s, int heartbeat_timeout_s,
                       int exchange_timeout_s, int watchdog_period_ms, const char* log_dir,
                       const char* work_dir, int quiet, mp4j_master** out) {
  MP4J_API_BEGIN
  MasterOptions o;
  if (connect_timeout_s > 0) o.connect_timeout_s = connect_timeout_s;
  if (heartbeat_timeout_s > 0) o.heartbeat_timeout_s = heartbeat_timeout_s;
  if (exchange_timeout_s > 0) o.exchange_timeout_s = exchange_timeout_s;
  if (watchdog_period_ms > 0) o.watchdog_period_ms = watchdog_period_ms;
  if (log_dir) o.log_dir = log_dir;
  if (work_dir && *work_dir) o.work_dir = work_dir;
  o.quiet = quiet != 0;
  *out = new mp4j_master{new Master(nslaves, port, o)};
  MP4J_API_END
}
int mp4j_master_start(mp4j_master* m) {
  MP4J_API_BEGIN
  m->m->start();
  MP4J_API_END
}
int mp4j_master_port(mp4j_master* m) { return m->m->port(); }
int mp4j_master_wait(mp4j_master* m, int timeout_ms, int* code) {
  try {
    if (timeout_ms < 0) {
      int c = m->m->wait();
      if (code) *code = c;
      return 0;
    }
    return m->m->wait_for(timeout_ms, code) ? 0 : 1;
  } catch (const std::exception& e) {
    g_last_error = e.what();
    return MP4J_ERR;
  }
}
int mp4j_master_shutdown(mp4j_master* m) {
  MP4J_API_BEGIN
  m->m->shutdown();
  MP4J_API_END
}
int mp4j_master_logs(mp4j_master* m, char** out, size_t* len) {
  MP4J_API_BEGIN
  std::string all;
  for (auto& l : m->m->logs()) {
    all += l;
    all += '\n';
  }
  *out = dup_bytes(all);
  *len = all.size();
  MP4J_API_END
}
void mp4j_master_destroy(mp4j_master* m) {
  if (!m) return;
  delete m->m;
  delete m;
}

// ---------------------------------------------------------------- communicator
int mp4j_comm_create(const char* login, const char* master_host, int master_port, int rank_hint,
                     const char* my_host, mp4j_comm** out) {
  MP4J_API_BEGIN
  auto* c = new mp4j_comm();
  try {
    c->pc = new ProcessComm(login ? login : "", master_host, master_port, rank_hint, my_host ? my_host : "");
  } catch (...) {
    delete c;
    throw;
  }
  *out = c;
  MP4J_API_END
}
void mp4j_comm_destroy(mp4j_comm* c) {
  if (!c) return;
  if (c->gpu_dtor) c->gpu_dtor(c);
  delete c->pc;
  delete c;
}
int mp4j_comm_rank(mp4j_comm* c) { return c->pc->rank(); }
int mp4j_comm_size(mp4j_comm* c) { return c->pc->size(); }
const char* mp4j_comm_host_transport(mp4j_comm* c) { return c->pc->host_transport(); }
int mp4j_comm_barrier(mp4j_comm* c) {
  MP4J_API_BEGIN
  c->pc->barrier();
  MP4J_API_END
}
int mp4j_comm_log(mp4j_comm* c, int level, const char* msg) {
  MP4J_API_BEGIN
  c->pc->log((LogLevel)level, msg);
  MP4J_API_END
}
int mp4j_comm_close(mp4j_comm* c, int code) {
  MP4J_API_BEGIN
  c->pc->close(code);
  MP4J_API_END
}
int mp4j_comm_exchange(mp4j_comm* c, int* other) {
  MP4J_API_BEGIN
  *other = c->pc->master().exchange(c->pc->rank());
  MP4J_API_END
}
int mp4j_comm_write_file(mp4j_comm* c, const char* content, const char* fname) {
  MP4J_API_BEGIN
  c->pc->master().write_file(content, fname);
  MP4J_API_END
}
int mp4j_comm_shutdown_master(mp4j_comm* c) {
  MP4J_API_BEGIN
  c->pc->master().shutdown_master();
  MP4J_API_END
}
int mp4j_comm_gather_all(mp4j_comm* c, const void* data, size_t n, char** out, size_t* out_len) {
  MP4J_API_BEGIN
  auto parts = c->pc->master().gather_all(c->pc->rank(), data, n);
  std::string all;
  for (auto& p : parts) {
    uint64_t len = p.size();
    all.append((const char*)&len, 8);
    all += p;
  }
  *out = dup_bytes(all);
  *out_len = all.size();
  MP4J_API_END
}

int mp4j_host_reduce_scatter(mp4j_comm* c, void* arr, int dtype, int op, int64_t from,
                             const int64_t* counts) {
  MP4J_API_BEGIN
  c->pc->reduce_scatter(arr, dtype, op, from, counts);
  MP4J_API_END
}
int mp4j_host_allgather(mp4j_comm* c, void* arr, int dtype, const int64_t* froms, const int64_t* tos) {
  MP4J_API_BEGIN
  c->pc->allgather(arr, dtype, froms, tos);
  MP4J_API_END
}
int mp4j_host_gather(mp4j_comm* c, void* arr, int dtype, const int64_t* froms, const int64_t* tos,
                     int root) {
  MP4J_API_BEGIN
  c->pc->gather(arr, dtype, froms, tos, root);
  MP4J_API_END
}
int mp4j_host_scatter(mp4j_comm* c, void* arr, int dtype, const int64_t* froms, const int64_t* tos,
                      int root) {
  MP4J_API_BEGIN
  c->pc->scatter(arr, dtype, froms, tos, root);
  MP4J_API_END
}
int mp4j_host_broadcast(mp4j_comm* c, void* arr, int dtype, int64_t from, int64_t to, int root) {
  MP4J_API_BEGIN
  c->pc->broadcast(arr, dtype, from, to, root);
  MP4J_API_END
}
int mp4j_host_reduce(mp4j_comm* c, void* arr, int dtype, int op, int64_t from, int64_t to, int root) {
  MP4J_API_BEGIN
  c->pc->reduce(arr, dtype, op, from, to, root);
  MP4J_API_END
}
int mp4j_host_allreduce(mp4j_comm* c, void* arr, int dtype, int op, int64_t from, int64_t to) {
  MP4J_API_BEGIN
  c->pc->allreduce(arr, dtype, op, from, to);
  MP4J_API_END
}
int mp4j_host_allreduce_rpc(mp4j_comm* c, void* arr, int dtype, int op, int64_t n) {
  MP4J_API_BEGIN
  c->pc->allreduce_rpc(arr, dtype, op, n);
  MP4J_API_END
}
int mp4j_host_reduce_into(int dtype, int op, void* dst, const void* src, int64_t n) {
  MP4J_API_BEGIN
  host_reduce_into(dtype, op, dst, src, (size_t)n);
  MP4J_API_END
}

int mp4j_host_map_reduce(mp4j_comm* c, int mode, int root, const char* keys, size_t keys_len, const void* vals,
                         size_t n, int dtype, int op, char** out_keys, size_t* out_keys_len, char** out_vals,
                         size_t* out_n) {
  MP4J_API_BEGIN
  std::string ok, ov;
  c->pc->map_reduce(mode, root, keys, keys_len, vals, n, dtype, op, &ok, &ov);
  *out_keys = dup_bytes(ok);
  *out_keys_len = ok.size();
  *out_vals = dup_bytes(ov);
  *out_n = ov.size() / dtype_size(dtype);
  MP4J_API_END
}

int mp4j_host_map_reduce_scatter(mp4j_comm* c, const char* keys, size_t keys_len, const void* vals, const int64_t* counts,
                                 int dtype, int op, char** out_keys, size_t* out_keys_len, char** out_vals,
                                 size_t* out_n) {
  MP4J_API_BEGIN
  std::string ok, ov;
  c->pc->map_reduce_scatter(keys, keys_len, vals, counts, dtype, op, &ok, &ov);
  *out_keys = dup_bytes(ok);
  *out_keys_len = ok.size();
  *out_vals = dup_bytes(ov);
  *out_n = ov.size() / dtype_size(dtype);
  MP4J_API_END
}

int mp4j_comm_send_bytes(mp4j_comm* c, int dest, const void* data, size_t n) {
  MP4J_API_BEGIN
  c->pc->send_bytes(dest, data, n);
  MP4J_API_END
}
int mp4j_comm_recv_bytes(mp4j_comm* c, int src, char** out, size_t* out_len) {
  MP4J_API_BEGIN
  std::string s = c->pc->recv_bytes(src);
  *out = dup_bytes(s);
  *out_len = s.size();
  MP4J_API_END
}
int mp4j_comm_sendrecv_bytes(mp4j_comm* c, int dest, const void* data, size_t n, int src, char** out,
                             size_t* out_len) {
  MP4J_API_BEGIN
  std::string s = c->pc->sendrecv_bytes(dest, data, n, src);
  *out = dup_bytes(s);
  *out_len = s.size();
  MP4J_API_END
}

int mp4j_host_blob_collective(mp4j_comm* c, int kind, int root, mp4j_blob_get get, mp4j_blob_put put, void* user) {
  MP4J_API_BEGIN
  switch (kind) {
    case MP4J_REDUCE_SCATTER: c->pc->blob_reduce_scatter(get, put, user); break;
    case MP4J_ALL_GATHER: c->pc->blob_allgather(get, put, user); break;
    case MP4J_GATHER: c->pc->blob_gather(root, get, put, user); break;
    case MP4J_SCATTER: c->pc->blob_scatter(root, get, put, user); break;
    default: throw Error("blob collective: kind must be reduce-scatter, allgather, gather or scatter");
  }
  MP4J_API_END
}

// ---------------------------------------------------------------- thread group
int mp4j_tgroup_create(int nthreads, mp4j_tgroup** out) {
  MP4J_API_BEGIN
  if (nthreads < 1) throw Error("threadNum must be >= 1");
  *out = new mp4j_tgroup{new ThreadComm(nthreads)};
  MP4J_API_END
}
void mp4j_tgroup_destroy(mp4j_tgroup* g) {
  if (!g) return;
  delete g->tc;
  delete g;
}
int mp4j_tgroup_barrier(mp4j_tgroup* g) {
  MP4J_API_BEGIN
  g->tc->group().barrier();
  MP4J_API_END
}
int mp4j_tgroup_set_slot(mp4j_tgroup* g, int tid, int which, uint64_t v) {
  MP4J_API_BEGIN
  if (tid < 0 || tid >= g->tc->threads() || which < 0 || which >= ThreadGroup::kSlots) throw Error("bad slot");
  g->tc->group().set_slot(tid, which, v);
  MP4J_API_END
}
uint64_t mp4j_tgroup_get_slot(mp4j_tgroup* g, int tid, int which) { return g->tc->group().get_slot(tid, which); }

// ThreadCommSlave on host arrays: thin wrappers over mp4j::ThreadComm (csrc/host/thread_comm.{h,cc}).
int mp4j_thread_allreduce(mp4j_tgroup* g, mp4j_comm* c, int tid, void* arr, int dtype, int op, int64_t from,
                          int64_t to) {
  MP4J_API_BEGIN
  g->tc->allreduce(*c->pc, tid, arr, dtype, op, from, to);
  MP4J_API_END
}
int mp4j_thread_collective(mp4j_tgroup* g, mp4j_comm* c, int tid, int kind, void* arr, int dtype, int op, int64_t from,
                           int64_t to, const int64_t* lo, const int64_t* hi, const int64_t* counts, int root_rank,
                           int root_tid) {
  MP4J_API_BEGIN
  g->tc->collective(*c->pc, tid, kind, arr, dtype, op, from, to, lo, hi, counts, root_rank, root_tid);
  MP4J_API_END
}

// ---------------------------------------------------------------- helpers
double mp4j_composite_double(float val, int32_t loc) { return composite_double(val, loc); }
float mp4j_composite_double_val(double c) { return composite_double_val(c); }
int32_t mp4j_composite_double_loc(double c) { return composite_double_loc(c); }
int64_t mp4j_composite_long(int32_t val, int32_t loc) { return composite_long(val, loc); }
int32_t mp4j_composite_long_val(int64_t c) { return composite_long_val(c); }
int32_t mp4j_composite_long_loc(int64_t c) { return composite_long_loc(c); }

// 64-bit keys for the GPU map kernels from NUL-separated UTF-8 strings: low 32 bits = String.hashCode (so that
// partition "java32" places a key on the rank the reference would, PCS:2063-2071), high 32 bits = FNV-1a of the
// bytes (an independent hash: two strings collide only if both 32-bit hashes agree). INT64_MIN, the kernels'
// free-slot marker, is never produced.
int mp4j_map_keys64(const char* blob, size_t blob_len, int64_t n, int64_t* out) {
  size_t pos = 0;
  for (int64_t i = 0; i < n; ++i) {
    size_t end = pos;
    while (end < blob_len && blob[end] != 0) ++end;
    uint32_t f = 2166136261u;
    for (size_t j = pos; j < end; ++j) f = (f ^ (unsigned char)blob[j]) * 16777619u;
    const uint32_t lo = (uint32_t)mp4j_java_string_hash(blob + pos, end - pos);
    uint64_t k = ((uint64_t)f << 32) | lo;
    if (k == 0x8000000000000000ull) k ^= 1ull << 32;
    out[i] = (int64_t)k;
    pos = end + 1;
  }
  return 0;
}

int32_t mp4j_java_string_hash(const char* s, size_t n) {
  // decode UTF-8 into UTF-16 code units on the fly; h = 31*h + unit (32-bit wrap)
  uint32_t h = 0;
  size_t i = 0;
  while (i < n) {
    uint32_t cp;
    unsigned char b = (unsigned char)s[i];
    if (b < 0x80) { cp = b; i += 1; }
    else if ((b >> 5) == 6 && i + 1 < n) { cp = ((b & 0x1f) << 6) | (s[i + 1] & 0x3f); i += 2; }
    else if ((b >> 4) == 14 && i + 2 < n) {
      cp = ((b & 0x0f) << 12) | ((s[i + 1] & 0x3f) << 6) | (s[i + 2] & 0x3f); i += 3;
    } else if ((b >> 3) == 30 && i + 3 < n) {
      cp = ((b & 0x07) << 18) | ((s[i + 1] & 0x3f) << 12) | ((s[i + 2] & 0x3f) << 6) | (s[i + 3] & 0x3f);
      i += 4;
    } else { cp = b; i += 1; }
    if (cp >= 0x10000) {
      cp -= 0x10000;
      h = 31 * h + (0xD800 + (cp >> 10));
      h = 31 * h + (0xDC00 + (cp & 0x3ff));
    } else {
      h = 31 * h + cp;
    }
  }
  return (int32_t)h;
}

}  // extern "C"
