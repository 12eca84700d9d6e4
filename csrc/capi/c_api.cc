// extern "C" surface over the host core (include/mp4j.h). The GPU half lives in csrc/gpu/gpu_api.cu.
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "../../include/mp4j.h"
#include "../control/control.h"
#include "../host/process_comm.h"
#include "../host/thread_group.h"
#include "capi_util.h"

using namespace mp4j;

namespace mp4j {
thread_local std::string g_last_error;
}

struct mp4j_master {
  Master* m;
};
struct mp4j_tgroup {
  ThreadGroup* g;
  std::string error;  // message of a failed delegate step, read by the other threads after a barrier
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
  *out = new mp4j_tgroup{new ThreadGroup(nthreads), std::string()};
  MP4J_API_END
}
void mp4j_tgroup_destroy(mp4j_tgroup* g) {
  if (!g) return;
  delete g->g;
  delete g;
}
int mp4j_tgroup_barrier(mp4j_tgroup* g) {
  MP4J_API_BEGIN
  g->g->barrier();
  MP4J_API_END
}
int mp4j_tgroup_set_slot(mp4j_tgroup* g, int tid, int which, uint64_t v) {
  MP4J_API_BEGIN
  if (tid < 0 || tid >= g->g->size() || which < 0 || which >= ThreadGroup::kSlots) throw Error("bad slot");
  g->g->set_slot(tid, which, v);
  MP4J_API_END
}
uint64_t mp4j_tgroup_get_slot(mp4j_tgroup* g, int tid, int which) { return g->g->get_slot(tid, which); }

// ThreadCommSlave.allreduceArray on host arrays (TCS:1915-1963), all T threads call it:
//   every thread reduces ITS 1/T slice of the range over the T arrays into thread 0's array
//   (thread-id order, cache-sized blocks) | thread 0 runs the process allreduce | every thread copies
//   the result into its own array. The reference does lg T serial full-length passes on pairs of
//   threads and one thread copies for everybody; here both thread-level steps use all T cores.
int mp4j_thread_allreduce(mp4j_tgroup* g, mp4j_comm* c, int tid, void* arr, int dtype, int op, int64_t from,
                          int64_t to) {
  MP4J_API_BEGIN
  ThreadGroup& tg = *g->g;
  const int T = tg.size();
  if (tid < 0 || tid >= T) throw Error("threadId out of range");
  check_from_to(from, to);
  if (!op_valid_for(dtype, op) || op == MP4J_COPY) throw Error("invalid operator for dtype");
  const size_t es = dtype_size(dtype);
  tg.set_slot(tid, 0, (uint64_t)(uintptr_t)arr);
  if (tid == 0) tg.set_slot(0, 1, 0);  // error flag of this call
  tg.barrier();
  char* acc = (char*)(uintptr_t)tg.get_slot(0, 0);
  const int64_t n = to - from, per = (n + T - 1) / T;
  const int64_t lo = from + std::min<int64_t>(n, (int64_t)tid * per), hi = from + std::min<int64_t>(n, (int64_t)(tid + 1) * per);
  const int64_t blk = (int64_t)(32768 / es);
  for (int64_t b0 = lo; b0 < hi; b0 += blk) {
    const int64_t b1 = std::min(hi, b0 + blk);
    for (int t = 1; t < T; ++t)
      host_reduce_into(dtype, op, acc + b0 * es, (const char*)(uintptr_t)tg.get_slot(t, 0) + b0 * es, (size_t)(b1 - b0));
  }
  tg.barrier();
  if (tid == 0) {
    try {
      c->pc->allreduce(acc, dtype, op, from, to);
    } catch (const std::exception& e) {
      tg.set_slot(0, 1, 1);
      g->error = e.what();
    }
  }
  tg.barrier();
  const bool failed = tg.get_slot(0, 1) != 0;
  if (!failed && tid != 0 && n > 0) memcpy((char*)arr + from * es, acc + from * es, (size_t)n * es);
  std::string err = failed ? g->error : std::string();
  tg.barrier();  // thread 0's array stays untouched until everybody has copied
  if (failed) throw Error(err);
  MP4J_API_END
}

// The other six ThreadCommSlave array collectives on host arrays (TCS:448-1430), same pattern: the T
// threads combine into the delegate's array IN PARALLEL (each thread its own slice or segment), the
// delegate (root thread on the root process, thread 0 elsewhere: TCS:486) runs the process collective,
// every thread fetches what it is owed in parallel. `kind` is an mp4j_collective; lo / hi are the 2-D
// ranges flattened as [slaveNum][threadNum] (gather, scatter, allgather), counts likewise (reduce-scatter).
int mp4j_thread_collective(mp4j_tgroup* g, mp4j_comm* c, int tid, int kind, void* arr, int dtype, int op, int64_t from,
                           int64_t to, const int64_t* lo, const int64_t* hi, const int64_t* counts, int root_rank,
                           int root_tid) {
  MP4J_API_BEGIN
  ThreadGroup& tg = *g->g;
  ProcessComm& pc = *c->pc;
  const int T = tg.size(), p = pc.size(), r = pc.rank();
  if (tid < 0 || tid >= T) throw Error("threadId out of range");
  const size_t es = dtype_size(dtype);
  if (es == 0) throw Error("unknown dtype");
  const bool rooted = kind == MP4J_GATHER || kind == MP4J_SCATTER || kind == MP4J_BROADCAST || kind == MP4J_REDUCE;
  if (rooted && (root_tid < 0 || root_tid >= T || root_rank < 0 || root_rank >= p)) throw Error("root out of range");
  const bool reducing = kind == MP4J_REDUCE || kind == MP4J_REDUCE_SCATTER;
  if (reducing && (!op_valid_for(dtype, op) || op == MP4J_COPY)) throw Error("invalid operator for dtype");
  const int rr = rooted ? (r == root_rank ? root_tid : 0) : 0;  // the delegate thread
  // process-level ranges derived from the 2-D ones
  std::vector<int64_t> pf((size_t)p), pt((size_t)p), pcnt((size_t)p);
  int64_t total_to = to;
  if (kind == MP4J_GATHER || kind == MP4J_SCATTER || kind == MP4J_ALL_GATHER) {
    for (int q = 0; q < p; ++q) {
      pf[(size_t)q] = lo[(size_t)q * T];
      pt[(size_t)q] = hi[(size_t)q * T + T - 1];
    }
  } else if (kind == MP4J_REDUCE_SCATTER) {
    total_to = from;
    for (int q = 0; q < p; ++q) {
      pcnt[(size_t)q] = 0;
      for (int t = 0; t < T; ++t) pcnt[(size_t)q] += counts[(size_t)q * T + t];
      total_to += pcnt[(size_t)q];
    }
  }
  tg.set_slot(tid, 0, (uint64_t)(uintptr_t)arr);
  if (tid == rr) tg.set_slot(rr, 1, 0);  // error flag of this call
  tg.barrier();
  char* dst = (char*)(uintptr_t)tg.get_slot(rr, 0);
  char* mine = (char*)arr;
  auto copy = [&](char* d, const char* s, int64_t a, int64_t b) {
    if (b > a && d != s) memcpy(d + a * es, s + a * es, (size_t)(b - a) * es);
  };
  // ---- combine into the delegate's array
  if (reducing) {
    const int64_t a = from, b = kind == MP4J_REDUCE ? to : total_to;
    const int64_t n = b - a, per = (n + T - 1) / T;
    const int64_t s0 = a + std::min<int64_t>(n, (int64_t)tid * per), s1 = a + std::min<int64_t>(n, (int64_t)(tid + 1) * per);
    const int64_t blk = (int64_t)(32768 / es);
    for (int64_t b0 = s0; b0 < s1; b0 += blk) {
      const int64_t b1 = std::min(s1, b0 + blk);
      for (int t = 0; t < T; ++t)  // delegate first (it is the accumulator), then ascending thread id
        if (t != rr) host_reduce_into(dtype, op, dst + b0 * es, (const char*)(uintptr_t)tg.get_slot(t, 0) + b0 * es, (size_t)(b1 - b0));
    }
  } else if (kind == MP4J_GATHER || kind == MP4J_ALL_GATHER) {
    copy(dst, mine, lo[(size_t)r * T + tid], hi[(size_t)r * T + tid]);
  }
  tg.barrier();
  // ---- the delegate runs the process collective
  if (tid == rr) {
    try {
      switch (kind) {
        case MP4J_GATHER: pc.gather(dst, dtype, pf.data(), pt.data(), root_rank); break;
        case MP4J_SCATTER: pc.scatter(dst, dtype, pf.data(), pt.data(), root_rank); break;
        case MP4J_ALL_GATHER: pc.allgather(dst, dtype, pf.data(), pt.data()); break;
        case MP4J_BROADCAST: pc.broadcast(dst, dtype, from, to, root_rank); break;
        case MP4J_REDUCE_SCATTER: pc.reduce_scatter(dst, dtype, op, from, pcnt.data()); break;
        case MP4J_REDUCE: pc.reduce(dst, dtype, op, from, to, root_rank); break;
        default: throw Error("mp4j_thread_collective: use mp4j_thread_allreduce for allreduce");
      }
    } catch (const std::exception& e) {
      tg.set_slot(rr, 1, 1);
      g->error = e.what();
    }
  }
  tg.barrier();
  const bool failed = tg.get_slot(rr, 1) != 0;
  // ---- every thread fetches what it is owed
  if (!failed && tid != rr) {
    if (kind == MP4J_ALL_GATHER) {
      for (int q = 0; q < p; ++q) copy(mine, dst, pf[(size_t)q], pt[(size_t)q]);
    } else if (kind == MP4J_BROADCAST) {
      copy(mine, dst, from, to);
    } else if (kind == MP4J_SCATTER) {
      copy(mine, dst, lo[(size_t)r * T + tid], hi[(size_t)r * T + tid]);
    }
  }
  if (!failed && kind == MP4J_REDUCE_SCATTER) {
    int64_t off = from;
    for (int q = 0; q < r; ++q) off += pcnt[(size_t)q];
    for (int t = 0; t < tid; ++t) off += counts[(size_t)r * T + t];
    copy(mine, dst, off, off + counts[(size_t)r * T + tid]);
  }
  std::string err = failed ? g->error : std::string();
  tg.barrier();  // the delegate's array stays untouched until everybody has copied
  if (failed) throw Error(err);
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
