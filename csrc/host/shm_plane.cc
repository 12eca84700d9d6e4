#include "shm_plane.h"

#include <errno.h>
#include <fcntl.h>
#include <limits.h>
#include <signal.h>
#include <linux/futex.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <sys/statvfs.h>
#include <sys/syscall.h>
#include <time.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <random>

#include "../control/control.h"
#include "../core/error.h"
#include "../core/ops.h"

namespace mp4j {

namespace {

constexpr size_t kHeaderBytes = 4096;

size_t env_size(const char* name, size_t dflt) {
  const char* v = getenv(name);
  return (v && *v) ? (size_t)strtoull(v, nullptr, 10) : dflt;
}

std::string boot_id() {
  std::string id;
  if (FILE* f = fopen("/proc/sys/kernel/random/boot_id", "r")) {
    char buf[128] = {0};
    if (fgets(buf, sizeof buf, f)) id = buf;
    fclose(f);
  }
  while (!id.empty() && (id.back() == '\n' || id.back() == ' ')) id.pop_back();
  if (id.empty()) {
    char host[256] = {0};
    gethostname(host, sizeof host - 1);
    id = std::string("host:") + host;
  }
  // POSIX shared memory is per IPC/mount namespace: two containers on one kernel share the boot id
  // but not /dev/shm, which the mapping attempt below detects.
  return id;
}

int futex(std::atomic<uint32_t>* addr, int op, uint32_t val, const struct timespec* ts) {
  return (int)syscall(SYS_futex, (uint32_t*)addr, op, val, ts, nullptr, 0);
}

inline void cpu_relax() {
#if defined(__x86_64__) || defined(__i386__)
  __builtin_ia32_pause();
#else
  std::atomic_signal_fence(std::memory_order_seq_cst);
#endif
}

// true if the process no longer runs: gone, or a zombie waiting to be reaped
bool process_gone(int32_t pid) {
  if (kill((pid_t)pid, 0) != 0) return errno == ESRCH;
  char path[64], buf[512];
  snprintf(path, sizeof path, "/proc/%d/stat", (int)pid);
  FILE* f = fopen(path, "r");
  if (!f) return false;
  const size_t k = fread(buf, 1, sizeof buf - 1, f);
  fclose(f);
  buf[k] = 0;
  const char* p = strrchr(buf, ')');  // the state letter follows the parenthesised command name
  return p && p[1] == ' ' && (p[2] == 'Z' || p[2] == 'X');
}

std::vector<std::string> split_lines(const std::string& s) {
  std::vector<std::string> out;
  size_t b = 0;
  for (;;) {
    size_t e = s.find('\n', b);
    out.push_back(s.substr(b, e == std::string::npos ? std::string::npos : e - b));
    if (e == std::string::npos) break;
    b = e + 1;
  }
  return out;
}

}  // namespace

struct ShmPlane::Header {
  std::atomic<uint32_t> count;
  char pad1[60];
  std::atomic<uint32_t> gen;
  char pad2[60];
  std::atomic<int32_t> pid;  // owner of this segment: lets waiting peers notice a dead process
};

// boot id of the machine, plus MP4J_NODE_ID when set (tests: several "machines" on one box)
std::string ShmPlane::machine_key() {
  std::string k = boot_id();
  if (const char* e = getenv("MP4J_NODE_ID")) k += std::string("#") + e;
  return k;
}

std::unique_ptr<ShmPlane> ShmPlane::create(MasterClient& master, int job_rank, int job_n, const std::vector<int>* group) {
  // members of this plane: the whole job, or the ranks of one machine (hierarchical collectives). Every rank of the
  // job calls this together (the exchanges below go through the master and involve everybody); the early returns
  // depend only on numbers that are the same on every rank.
  std::vector<int> members;
  if (group) members = *group;
  else
    for (int q = 0; q < job_n; ++q) members.push_back(q);
  const int n = (int)members.size();
  int rank = -1;
  for (int i = 0; i < n; ++i)
    if (members[(size_t)i] == job_rank) rank = i;
  if (n < 2 || n > 32 || rank < 0) return nullptr;  // masks are 32-bit
  bool enabled = env_size("MP4J_SHM", 1) != 0;
  size_t window = env_size("MP4J_SHM_BYTES", 64ull << 20);
  window = std::max<size_t>(window, 1 << 16) / 8192 * 8192;  // two page-aligned halves
  const size_t seg_bytes = kHeaderBytes + window;
  {
    // The segments are sparse files on a tmpfs: writing beyond what /dev/shm can hold would raise
    // SIGBUS in the middle of a collective (containers often cap it at 64 MB). Use shared memory only if
    // all n windows fit with room to spare; otherwise every rank reports "no" and TCP is used.
    struct statvfs vfs;
    if (statvfs("/dev/shm", &vfs) != 0 ||
        (unsigned long long)vfs.f_bavail * vfs.f_frsize < (unsigned long long)n * seg_bytes + (32ull << 20))
      enabled = false;
  }
  std::unique_ptr<ShmPlane> p(new ShmPlane());
  p->rank_ = rank;
  p->n_ = n;
  p->seg_bytes_ = seg_bytes;
  p->half_bytes_ = window / 2;
  p->timeout_s_ = (double)env_size("MP4J_PEER_TIMEOUT_S", 600);
  p->seg_.assign(n, nullptr);
  {
    // spinning only pays when every rank keeps a core to itself (helpers: master, heartbeats, the caller's threads)
    const long cores = sysconf(_SC_NPROCESSORS_ONLN);
    p->spin_us_ = cores >= 2L * n ? 200.0 : (cores >= n + 2 ? 20.0 : 0.0);
    if (getenv("MP4J_SHM_SPIN_US")) p->spin_us_ = (double)env_size("MP4J_SHM_SPIN_US", 0);
  }

  // ---- my segment
  std::string name;
  if (enabled) {
    std::random_device rd;
    char buf[96];
    snprintf(buf, sizeof buf, "/mp4j-%d-%d-%08x", (int)getpid(), rank, (unsigned)rd());
    int fd = shm_open(buf, O_CREAT | O_EXCL | O_RDWR, 0600);
    if (fd >= 0) {
      if (ftruncate(fd, (off_t)seg_bytes) == 0) {
        void* m = mmap(nullptr, seg_bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
        if (m != MAP_FAILED) {
          p->seg_[rank] = (char*)m;
          memset(m, 0, kHeaderBytes);
          ((Header*)m)->pid.store((int32_t)getpid(), std::memory_order_release);
          name = buf;
        }
      }
      ::close(fd);
      if (name.empty()) shm_unlink(buf);
    }
  }
  // ---- exchange (boot id, name, window) and map the peers
  const std::string mine = machine_key() + "\n" + name + "\n" + std::to_string(window);
  std::vector<std::string> parts = master.gather_all(job_rank, mine.data(), mine.size());
  bool ok = enabled && !name.empty() && (int)parts.size() == job_n;
  std::vector<std::string> names(n);
  for (int q = 0; ok && q < n; ++q) {
    std::vector<std::string> f = split_lines(parts[(size_t)members[(size_t)q]]);
    if (f.size() != 3 || f[1].empty() || f[0] != split_lines(mine)[0] || f[2] != std::to_string(window)) ok = false;
    else names[q] = f[1];
  }
  for (int q = 0; ok && q < n; ++q) {
    if (q == rank) continue;
    int fd = shm_open(names[q].c_str(), O_RDWR, 0600);
    if (fd < 0) {
      ok = false;
      break;
    }
    void* m = mmap(nullptr, seg_bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    ::close(fd);
    if (m == MAP_FAILED) ok = false;
    else p->seg_[q] = (char*)m;
  }
  // ---- everybody must agree, otherwise ranks would run different algorithms
  const char verdict = ok ? '1' : '0';
  parts = master.gather_all(job_rank, &verdict, 1);
  bool all_ok = (int)parts.size() == job_n;
  for (const auto& v : parts)
    if (v.size() != 1 || v[0] != '1') all_ok = false;
  if (!name.empty()) shm_unlink(name.c_str());  // everyone who could map it has done so: no leftovers on exit
  if (!all_ok) return nullptr;                  // ~ShmPlane unmaps
  return p;
}

ShmPlane::~ShmPlane() {
  for (char* s : seg_)
    if (s) munmap(s, seg_bytes_);
}

void ShmPlane::barrier() {
  Header* h = (Header*)seg_[0];
  const uint32_t g = h->gen.load(std::memory_order_acquire);
  if (h->count.fetch_add(1, std::memory_order_acq_rel) + 1 == (uint32_t)n_) {
    h->count.store(0, std::memory_order_relaxed);
    h->gen.fetch_add(1, std::memory_order_release);
    futex(&h->gen, FUTEX_WAKE, INT_MAX, nullptr);
    return;
  }
  // Phases of one pass take tens to hundreds of microseconds and a futex sleep costs about as much
  // to wake from, so spin for up to ~200 us (less or not at all when cores are scarce) before sleeping.
  const auto t0 = std::chrono::steady_clock::now();
  for (int i = 0; spin_us_ > 0; ++i) {
    if (h->gen.load(std::memory_order_acquire) != g) return;
    cpu_relax();
    if ((i & 255) == 255 &&
        std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count() > spin_us_)
      break;
  }
  while (h->gen.load(std::memory_order_acquire) == g) {
    struct timespec ts = {0, 100 * 1000 * 1000};
    futex(&h->gen, FUTEX_WAIT, g, &ts);  // returns on wake, value change, timeout or signal: re-check
    // fail fast: a peer process that no longer exists will never arrive
    for (int q = 0; q < n_; ++q) {
      const int32_t pid = ((Header*)seg_[q])->pid.load(std::memory_order_acquire);
      if (pid > 0 && pid != (int32_t)getpid() && process_gone(pid) && h->gen.load(std::memory_order_acquire) == g)
        throw Error("rank " + std::to_string(q) + " (pid " + std::to_string(pid) + ") died during a collective");
    }
    if (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() > timeout_s_)
      throw Error("shared-memory barrier timed out: a peer left the collective (MP4J_PEER_TIMEOUT_S)");
  }
}

void ShmPlane::grs(void* arr, int dtype, int op, const int64_t* lo, const int64_t* hi, const uint32_t* src_mask,
                   const uint32_t* dst_mask) {
  grs_team(arr, dtype, op, lo, hi, src_mask, dst_mask, 0, 1, nullptr);
}

// The same pass executed by a TEAM of T threads of this process (ThreadCommSlave: the process step of a thread-level
// collective, which one delegate thread used to run alone while T - 1 cores waited). All T threads call with the same
// arguments; every copy and every reduction is cut into T parts, thread `tid` does part `tid`; thread 0 meets the
// other processes at the barriers, bracketed by team_sync (the thread barrier). A failed process barrier is
// delivered to all T threads.
void ShmPlane::grs_team(void* arr, int dtype, int op, const int64_t* lo, const int64_t* hi, const uint32_t* src_mask,
                        const uint32_t* dst_mask, int tid, int T, const std::function<void()>* team_sync) {
  auto tsync = [&] {
    if (T > 1) (*team_sync)();
  };
  auto xbarrier = [&] {
    if (T == 1) return barrier();
    tsync();
    if (tid == 0) {
      try {
        barrier();
        team_failed_ = false;
      } catch (const std::exception& e) {
        team_failed_ = true;
        team_msg_ = e.what();
      }
    }
    tsync();
    if (team_failed_) throw Error(team_msg_);
  };
  // this thread's part of [l, h)
  auto part = [&](int64_t* l, int64_t* h) {
    const int64_t len = *h - *l, a = *l + len * tid / T, b = *l + len * (tid + 1) / T;
    *l = a;
    *h = b;
  };
  const size_t es = dtype_size(dtype);
  if (es == 0) throw Error("unknown dtype");
  int64_t span_lo = INT64_MAX, span_hi = INT64_MIN;
  bool need_b2 = false;  // some destination reads a reduced slice out of its owner's window
  for (int q = 0; q < n_; ++q) {
    if (hi[q] <= lo[q]) continue;
    span_lo = std::min(span_lo, lo[q]);
    span_hi = std::max(span_hi, hi[q]);
    if (op != MP4J_COPY && dst_mask[q] != (1u << q)) need_b2 = true;
  }
  if (span_hi <= span_lo) return;
  char* base = (char*)arr;
  // Arrays larger than a window half go through it in several passes. A pass takes the k-th part of EVERY slice
  // (part sizes proportional to the slice lengths), not the next contiguous stretch of the array: a contiguous
  // stretch would hold the slices of one or two owners only, and the other ranks would sit out the pass (measured:
  // reduce-scatter of 80 MB over 4 ranks slower than the TCP ring).
  const int64_t chunk = (int64_t)(half_bytes_ / es);
  int64_t total = 0;
  for (int q = 0; q < n_; ++q) total += std::max<int64_t>(0, hi[q] - lo[q]);
  const int64_t kAlign = 8;  // parts start on 64-byte boundaries of the window (8-byte elements)
  const int64_t room = std::max<int64_t>(1, chunk - (int64_t)n_ * kAlign);
  const int64_t passes = (total + room - 1) / room;
  std::vector<int64_t> sub((size_t)n_), woff((size_t)n_);
  {
    int64_t off = 0;
    for (int q = 0; q < n_; ++q) {
      const int64_t len = std::max<int64_t>(0, hi[q] - lo[q]);
      sub[(size_t)q] = ((len + passes - 1) / passes + kAlign - 1) / kAlign * kAlign;
      woff[(size_t)q] = off;
      off += sub[(size_t)q];
    }
    if (off > chunk) throw Error("shm window arithmetic");  // cannot happen: sum <= total / passes + n * kAlign
  }
  std::vector<char> acc;  // private accumulator for slices whose result goes back into my window
  for (int64_t k = 0; k < passes; ++k) {
    tsync();  // thread 0 has flipped half_ (end of the previous pass / call) before anybody reads it
    const size_t hoff = kHeaderBytes + (size_t)half_ * half_bytes_;
    // part k of slice q: [pl(q), ph(q)); its place in everybody's window: woff[q]
    auto pl = [&](int q) { return std::min(hi[q], lo[q] + k * sub[(size_t)q]); };
    auto ph = [&](int q) { return std::min(hi[q], lo[q] + (k + 1) * sub[(size_t)q]); };
    auto win = [&](int r, int q, int64_t elem) {
      return seg_[r] + hoff + (size_t)(woff[(size_t)q] + (elem - pl(q))) * es;
    };
    // ---- phase 1: publish my contributions
    for (int q = 0; q < n_; ++q) {
      if (!((src_mask[q] >> rank_) & 1u)) continue;
      int64_t l = pl(q), h = ph(q);
      if (l >= h) continue;
      if (op == MP4J_COPY && dst_mask[q] == (1u << rank_)) continue;  // my own slice to myself
      if (op != MP4J_COPY && q == rank_) continue;  // the slice I reduce myself: read from my array in phase 2
      part(&l, &h);
      if (l < h) memcpy(win(rank_, q, l), base + l * es, (size_t)(h - l) * es);
    }
    xbarrier();
    // ---- phase 2: the owner reduces its slice over the source windows, ascending rank order
    if (op != MP4J_COPY) {
      const int q = rank_;
      int64_t l = pl(q), h = ph(q);
      part(&l, &h);
      if (l < h && src_mask[q] != 0) {
        const bool to_me = (dst_mask[q] >> q) & 1u;
        const bool to_others = (dst_mask[q] & ~(1u << q)) != 0;
        // Blocked: a 32 KiB accumulator stays in L1 while the p source streams pass over it. My own
        // contribution comes straight from my array (it was not copied in), the others from their
        // windows; the result goes to my array and, if others want it, to my window.
        const int64_t blk = (int64_t)(32768 / es);
        if (acc.size() < 32768) acc.resize(32768);
        char* o = acc.data();
        for (int64_t b0 = l; b0 < h; b0 += blk) {
          const int64_t b1 = std::min(h, b0 + blk);
          const size_t bb = (size_t)(b1 - b0) * es;
          bool first = true;
          for (int s = 0; s < n_; ++s) {
            if (!((src_mask[q] >> s) & 1u)) continue;
            const char* x = (s == rank_) ? base + b0 * es : win(s, q, b0);
            if (first) memcpy(o, x, bb);
            else host_reduce_into(dtype, op, o, x, (size_t)(b1 - b0));  // op(accumulated, next rank)
            first = false;
          }
          if (to_me) memcpy(base + b0 * es, o, bb);
          if (to_others) memcpy(win(rank_, q, b0), o, bb);
        }
      }
      if (need_b2) xbarrier();
    }
    // ---- phase 3: fetch the slices I am a destination of
    for (int q = 0; q < n_; ++q) {
      if (!((dst_mask[q] >> rank_) & 1u)) continue;
      int64_t l = pl(q), h = ph(q);
      if (l >= h || src_mask[q] == 0) continue;
      int from_rank;
      if (op != MP4J_COPY) {
        if (q == rank_) continue;  // written by phase 2
        from_rank = q;
      } else {
        from_rank = __builtin_ctz(src_mask[q]);
        if (from_rank == rank_) continue;  // my own data
      }
      part(&l, &h);
      if (l < h) memcpy(base + l * es, win(from_rank, q, l), (size_t)(h - l) * es);
    }
    if (tid == 0) half_ ^= 1;  // the other half serves the next chunk / call: no closing barrier needed
  }
}

}  // namespace mp4j
