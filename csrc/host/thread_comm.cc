// ThreadComm: see thread_comm.h.
#include "thread_comm.h"

#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <functional>
#include <vector>

#include "../../include/mp4j_types.h"
#include "../core/error.h"
#include "../core/ops.h"
#include "../core/partition.h"

namespace mp4j {

// MP4J_THREAD_TEAM=0: the delegate thread runs the process step alone (the behaviour before teams existed)
static bool team_enabled() {
  static const bool on = !(getenv("MP4J_THREAD_TEAM") && atoi(getenv("MP4J_THREAD_TEAM")) == 0);
  return on;
}

// ThreadCommSlave.allreduceArray on host arrays (TCS:1915-1963), all T threads call it:
//   every thread reduces ITS 1/T slice of the range over the T arrays into thread 0's array
//   (thread-id order, cache-sized blocks) | thread 0 runs the process allreduce | every thread copies
//   the result into its own array. The reference does lg T serial full-length passes on pairs of
//   threads and one thread copies for everybody; here both thread-level steps use all T cores.
void ThreadComm::allreduce(ProcessComm& pc, int tid, void* arr, int dtype, int op, int64_t from, int64_t to) {
  ThreadGroup& tg = tg_;
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
  if (pc.size() == 1) {
    // One process: there is no process step between combining and fanning out, so every thread finishes its slice
    // in one pass - reduce a cache-sized block of the T arrays (thread-id order) into a local buffer and write it
    // straight into all T arrays: T reads + T writes per element instead of T reads + 1 write, then T - 1 reads +
    // T - 1 writes for the copy-out.
    thread_local std::vector<char> blockbuf;
    if (blockbuf.size() < 32768) blockbuf.resize(32768);
    char* o = blockbuf.data();
    for (int64_t b0 = lo; b0 < hi; b0 += blk) {
      const int64_t b1 = std::min(hi, b0 + blk);
      const size_t bb = (size_t)(b1 - b0) * es;
      memcpy(o, acc + b0 * es, bb);
      for (int t = 1; t < T; ++t)
        host_reduce_into(dtype, op, o, (const char*)(uintptr_t)tg.get_slot(t, 0) + b0 * es, (size_t)(b1 - b0));
      for (int t = 0; t < T; ++t) memcpy((char*)(uintptr_t)tg.get_slot(t, 0) + b0 * es, o, bb);
    }
    tg.barrier();  // nobody returns while another thread still writes into its array
    return;
  }
  for (int64_t b0 = lo; b0 < hi; b0 += blk) {
    const int64_t b1 = std::min(hi, b0 + blk);
    for (int t = 1; t < T; ++t)
      host_reduce_into(dtype, op, acc + b0 * es, (const char*)(uintptr_t)tg.get_slot(t, 0) + b0 * es, (size_t)(b1 - b0));
  }
  tg.barrier();
  {
    // the process step: on the shared-memory plane all T threads share it, on the TCP ring thread 0 runs it
    const std::function<void()> sync = [&tg] { tg.barrier(); };
    const ProcessComm::Team team{tid, T, &sync};
    ProcessComm::TeamScope scope(team_enabled() ? &team : nullptr);
    try {
      if (team_enabled() || tid == 0) pc.allreduce(acc, dtype, op, from, to);
    } catch (const std::exception& e) {
      if (tid == 0) {  // a team fails as one (every thread gets the same error); thread 0 records it
        tg.set_slot(0, 1, 1);
        error_ = e.what();
      }
    }
  }
  tg.barrier();
  const bool failed = tg.get_slot(0, 1) != 0;
  if (!failed && tid != 0 && n > 0) memcpy((char*)arr + from * es, acc + from * es, (size_t)n * es);
  std::string err = failed ? error_ : std::string();
  tg.barrier();  // thread 0's array stays untouched until everybody has copied
  if (failed) throw Error(err);
}

// The other six ThreadCommSlave array collectives on host arrays (TCS:448-1430), same pattern: the T
// threads combine into the delegate's array IN PARALLEL (each thread its own slice or segment), the
// delegate (root thread on the root process, thread 0 elsewhere: TCS:486) runs the process collective,
// every thread fetches what it is owed in parallel. `kind` is an mp4j_collective; lo / hi are the 2-D
// ranges flattened as [slaveNum][threadNum] (gather, scatter, allgather), counts likewise (reduce-scatter).
void ThreadComm::collective(ProcessComm& pc, int tid, int kind, void* arr, int dtype, int op, int64_t from,
                            int64_t to, const int64_t* lo, const int64_t* hi, const int64_t* counts, int root_rank,
                            int root_tid) {
  ThreadGroup& tg = tg_;
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
  // ---- the process collective: a team of all T threads whose thread 0 is the delegate (shared-memory plane: the
  // threads share copies and reduction; TCP: the delegate runs it, the others fall through to the barrier)
  {
    const std::function<void()> sync = [&tg] { tg.barrier(); };
    const ProcessComm::Team team{(tid - rr + T) % T, T, &sync};
    ProcessComm::TeamScope scope(team_enabled() ? &team : nullptr);
    try {
      if (team_enabled() || tid == rr) switch (kind) {
        case MP4J_GATHER: pc.gather(dst, dtype, pf.data(), pt.data(), root_rank); break;
        case MP4J_SCATTER: pc.scatter(dst, dtype, pf.data(), pt.data(), root_rank); break;
        case MP4J_ALL_GATHER: pc.allgather(dst, dtype, pf.data(), pt.data()); break;
        case MP4J_BROADCAST: pc.broadcast(dst, dtype, from, to, root_rank); break;
        case MP4J_REDUCE_SCATTER: pc.reduce_scatter(dst, dtype, op, from, pcnt.data()); break;
        case MP4J_REDUCE: pc.reduce(dst, dtype, op, from, to, root_rank); break;
        default: throw Error("ThreadComm::collective: use allreduce() for allreduce");
      }
    } catch (const std::exception& e) {
      if (tid == rr) {
        tg.set_slot(rr, 1, 1);
        error_ = e.what();
      }
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
  std::string err = failed ? error_ : std::string();
  tg.barrier();  // the delegate's array stays untouched until everybody has copied
  if (failed) throw Error(err);
}

}  // namespace mp4j
