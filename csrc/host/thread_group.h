// ThreadGroup: host-side rendezvous of the T threads of one ThreadCommSlave.
// Replaces the reference's Exchanger/CyclicBarrier/volatile delegatingMetaData machinery
// (TCS:52-65,139-157,259-303): threads deposit their argument pointers in per-thread slots,
// meet at a spin-then-sleep barrier, and ONE delegate thread launches the fused kernel.
#pragma once
#include <stdint.h>

#include <atomic>
#include <condition_variable>
#include <mutex>
#include <vector>

namespace mp4j {

class ThreadGroup {
 public:
  explicit ThreadGroup(int nthreads);
  int size() const { return n_; }
  void barrier();
  void set_slot(int tid, int which, uint64_t v) { slots_[(size_t)tid * kSlots + which] = v; }
  uint64_t get_slot(int tid, int which) const { return slots_[(size_t)tid * kSlots + which]; }
  static constexpr int kSlots = 8;

 private:
  const int n_;
  std::atomic<int> count_{0};
  std::atomic<uint64_t> gen_{0};
  std::mutex mu_;
  std::condition_variable cv_;
  std::vector<uint64_t> slots_;
};

}  // namespace mp4j
