// ThreadComm: the host-array side of ThreadCommSlave (TCS:448-1963) as a C++ class, used by every binding
// through the C ABI (include/mp4j.h: mp4j_thread_allreduce, mp4j_thread_collective) and directly by C++ callers.
// T threads of one process share one object; each collective is called by ALL T threads with their own
// array. Pattern of every collective: the threads combine into the delegate thread's array in parallel
// (each thread its own slice or segment), the delegate (root thread on the root process, thread 0 elsewhere:
// TCS:486) runs the process collective, every thread fetches what it is owed in parallel. The reference
// does lg T serial full-length passes over pairs of threads (Exchanger tree, TCS:259-303) and has one thread
// copy for everybody.
#pragma once
#include <stdint.h>

#include <exception>
#include <string>
#include <vector>

#include "../core/error.h"
#include "process_comm.h"
#include "thread_group.h"

namespace mp4j {

class ThreadComm {
 public:
  explicit ThreadComm(int nthreads) : tg_(nthreads) {}
  int threads() const { return tg_.size(); }
  ThreadGroup& group() { return tg_; }

  // allreduceArray (TCS:1915-1963)
  void allreduce(ProcessComm& pc, int tid, void* arr, int dtype, int op, int64_t from, int64_t to);
  // gather / scatter / allgather / broadcast / reduce-scatter / reduce (TCS:448-1430). `kind` is an
  // mp4j_collective; lo / hi are the 2-D ranges flattened as [slaveNum][threadNum] (gather, scatter,
  // allgather), counts likewise (reduce-scatter).
  void collective(ProcessComm& pc, int tid, int kind, void* arr, int dtype, int op, int64_t from, int64_t to,
                  const int64_t* lo, const int64_t* hi, const int64_t* counts, int root_rank, int root_tid);

  // The rendezvous every thread-level collective is built on, for steps that ONE thread performs for all (a fused
  // GPU kernel over the T device buffers, a host pipeline): every thread deposits `my_arg` (a pointer), thread
  // `root` runs fn(args of all T threads), everybody returns when it has finished; an exception thrown by fn is
  // rethrown as mp4j::Error on ALL T threads. Replaces the Exchanger tree + volatile metadata of the reference
  // (TCS:259-303) and is what the bindings' own delegate loops (Python: thread.py _delegate) do for objects.
  template <typename F>
  void delegate(int tid, int root, uint64_t my_arg, F&& fn) {
    const int T = tg_.size();
    if (tid < 0 || tid >= T || root < 0 || root >= T) throw Error("threadId out of range");
    tg_.set_slot(tid, 0, my_arg);
    if (tid == root) tg_.set_slot(root, 1, 0);  // error flag of this call
    tg_.barrier();
    if (tid == root) {
      try {
        std::vector<uint64_t> args((size_t)T);
        for (int t = 0; t < T; ++t) args[(size_t)t] = tg_.get_slot(t, 0);
        fn(args);
      } catch (const std::exception& e) {
        tg_.set_slot(root, 1, 1);
        error_ = e.what();
      }
    }
    tg_.barrier();
    const bool failed = tg_.get_slot(root, 1) != 0;
    const std::string err = failed ? error_ : std::string();
    tg_.barrier();  // error_ and the slots stay untouched until everybody has read them
    if (failed) throw Error(err);
  }

 private:
  ThreadGroup tg_;
  std::string error_;  // message of a failed delegate step, read by the other threads after a barrier
};

}  // namespace mp4j
