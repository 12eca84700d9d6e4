#include "thread_group.h"

#include <chrono>
#include <thread>

namespace mp4j {

ThreadGroup::ThreadGroup(int nthreads) : n_(nthreads), slots_((size_t)nthreads * kSlots, 0) {}

void ThreadGroup::barrier() {
  if (n_ <= 1) return;
  uint64_t gen = gen_.load(std::memory_order_acquire);
  if (count_.fetch_add(1, std::memory_order_acq_rel) + 1 == n_) {
    count_.store(0, std::memory_order_relaxed);
    {
      std::lock_guard<std::mutex> lk(mu_);
      gen_.store(gen + 1, std::memory_order_release);
    }
    cv_.notify_all();
    return;
  }
  // short spin (the common case: all threads arrive within microseconds), then sleep
  auto t0 = std::chrono::steady_clock::now();
  for (int i = 0;; ++i) {
    if (gen_.load(std::memory_order_acquire) != gen) return;
    if ((i & 63) == 63) {
      if (std::chrono::steady_clock::now() - t0 > std::chrono::microseconds(50)) break;
      std::this_thread::yield();
    }
  }
  std::unique_lock<std::mutex> lk(mu_);
  cv_.wait(lk, [&] { return gen_.load(std::memory_order_acquire) != gen; });
}

}  // namespace mp4j
