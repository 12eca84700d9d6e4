// Native multi-rank stress test of the host control + data plane through the C ABI: one master and
// `p` ranks as threads of ONE process (each rank is a full communicator with its own TCP sockets).
// It is the "single-process multi-rank harness" the reference lacks (Mocks: none) and the
// target of the sanitizer builds: build.py --cpp-tests [thread|address] compiles the
// host sources together with this file under -fsanitize=thread / address and runs it.
//
// Value patterns follow the reference's check classes (J/check/checkdouble/Process*Check.java),
// widened to every operator, non-zero roots, from > 0 and empty segments.
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <algorithm>
#include <string>
#include <thread>
#include <vector>

#include "../../csrc/host/thread_comm.h"
#include "../../include/mp4j.h"

static std::atomic<int> g_fail{0};

#define CHECK(cond, ...)                                   \
  do {                                                     \
    if (!(cond)) {                                         \
      fprintf(stderr, "[rank %d] FAIL %s:%d: ", r, __FILE__, __LINE__); \
      fprintf(stderr, __VA_ARGS__);                        \
      fprintf(stderr, "\n");                               \
      g_fail.fetch_add(1);                                 \
    }                                                      \
  } while (0)
#define OK(call) CHECK((call) == 0, "%s -> %s", #call, mp4j_last_error())

template <typename T>
struct Code;
template <> struct Code<double> { static const int v = MP4J_F64; };
template <> struct Code<float> { static const int v = MP4J_F32; };
template <> struct Code<int64_t> { static const int v = MP4J_I64; };
template <> struct Code<int32_t> { static const int v = MP4J_I32; };
template <> struct Code<int16_t> { static const int v = MP4J_I16; };
template <> struct Code<int8_t> { static const int v = MP4J_I8; };

static void split(int64_t from, int64_t to, int p, std::vector<int64_t>* lo, std::vector<int64_t>* hi) {
  const int64_t avg = (to - from) / p;  // last rank takes the remainder (PCS:1741-1756)
  lo->resize(p);
  hi->resize(p);
  for (int q = 0; q < p; ++q) {
    (*lo)[q] = from + q * avg;
    (*hi)[q] = q == p - 1 ? to : from + (q + 1) * avg;
  }
}

template <typename T>
static void dense_suite(mp4j_comm* c, int r, int p, int64_t n) {
  const int dt = Code<T>::v;
  const int64_t from = n > 8 ? 3 : 0;
  std::vector<int64_t> lo, hi, cnt(p);
  split(from, n, p, &lo, &hi);
  for (int q = 0; q < p; ++q) cnt[q] = hi[q] - lo[q];
  // allreduce SUM / MAX / MIN: rank r contributes (r + 1) everywhere
  for (int op : {MP4J_SUM, MP4J_MAX, MP4J_MIN}) {
    std::vector<T> a(n, (T)(r + 1));
    OK(mp4j_host_allreduce(c, a.data(), dt, op, from, n));
    const T want = op == MP4J_SUM ? (T)(p * (p + 1) / 2) : (op == MP4J_MAX ? (T)p : (T)1);
    for (int64_t i = 0; i < n; ++i) {
      const T w = i < from ? (T)(r + 1) : want;
      if (a[i] != w) {
        CHECK(false, "allreduce dt=%d op=%d i=%lld got=%f want=%f", dt, op, (long long)i, (double)a[i], (double)w);
        break;
      }
    }
  }
  // reduce-scatter: arr[i] = owner(i) on every rank -> own slice = owner * p (ProcessReduceScatterCheck:56-85)
  {
    std::vector<T> a(n, (T)0);
    for (int q = 0; q < p; ++q)
      for (int64_t i = lo[q]; i < hi[q]; ++i) a[i] = (T)q;
    OK(mp4j_host_reduce_scatter(c, a.data(), dt, MP4J_SUM, from, cnt.data()));
    for (int64_t i = lo[r]; i < hi[r]; ++i)
      if (a[i] != (T)(r * p)) {
        CHECK(false, "reduce_scatter dt=%d i=%lld got=%f", dt, (long long)i, (double)a[i]);
        break;
      }
  }
  // allgather / gather / scatter with a non-zero root
  const int root = p - 1;
  {
    std::vector<T> a(n, (T)-1);
    for (int64_t i = lo[r]; i < hi[r]; ++i) a[i] = (T)r;
    OK(mp4j_host_allgather(c, a.data(), dt, lo.data(), hi.data()));
    for (int q = 0; q < p; ++q)
      for (int64_t i = lo[q]; i < hi[q]; ++i)
        if (a[i] != (T)q) {
          CHECK(false, "allgather dt=%d i=%lld", dt, (long long)i);
          q = p;
          break;
        }
  }
  {
    std::vector<T> a(n, (T)-1);
    for (int64_t i = lo[r]; i < hi[r]; ++i) a[i] = (T)r;
    OK(mp4j_host_gather(c, a.data(), dt, lo.data(), hi.data(), root));
    if (r == root)
      for (int q = 0; q < p; ++q)
        for (int64_t i = lo[q]; i < hi[q]; ++i)
          if (a[i] != (T)q) {
            CHECK(false, "gather dt=%d i=%lld", dt, (long long)i);
            q = p;
            break;
          }
  }
  {
    std::vector<T> a(n, (T)-1);
    if (r == root)
      for (int q = 0; q < p; ++q)
        for (int64_t i = lo[q]; i < hi[q]; ++i) a[i] = (T)q;
    OK(mp4j_host_scatter(c, a.data(), dt, lo.data(), hi.data(), root));
    for (int64_t i = lo[r]; i < hi[r]; ++i)
      if (a[i] != (T)r) {
        CHECK(false, "scatter dt=%d i=%lld", dt, (long long)i);
        break;
      }
  }
  {
    std::vector<T> a(n, r == root ? (T)1 : (T)-1);
    OK(mp4j_host_broadcast(c, a.data(), dt, from, n, root));
    for (int64_t i = from; i < n; ++i)
      if (a[i] != (T)1) {
        CHECK(false, "broadcast dt=%d i=%lld", dt, (long long)i);
        break;
      }
  }
  {
    std::vector<T> a(n, (T)1);
    OK(mp4j_host_reduce(c, a.data(), dt, MP4J_SUM, from, n, root));
    if (r == root)
      for (int64_t i = from; i < n; ++i)
        if (a[i] != (T)p) {
          CHECK(false, "reduce dt=%d i=%lld", dt, (long long)i);
          break;
        }
  }
  {  // RPC allreduce through the master mailbox (small arrays, PCS:1776)
    std::vector<T> a(7, (T)1);
    OK(mp4j_host_allreduce_rpc(c, a.data(), dt, MP4J_SUM, 7));
    CHECK(a[0] == (T)p && a[6] == (T)p, "allreduce_rpc dt=%d", dt);
  }
}

static void rank_main(int port, int r, int p, int rounds, int64_t n) {
  mp4j_comm* c = nullptr;
  if (mp4j_comm_create("stress", "127.0.0.1", port, r, "127.0.0.1", &c) != 0) {
    fprintf(stderr, "[rank %d] comm_create: %s\n", r, mp4j_last_error());
    g_fail.fetch_add(1);
    return;
  }
  CHECK(mp4j_comm_rank(c) == r && mp4j_comm_size(c) == p, "rank/size");
  for (int it = 0; it < rounds; ++it) {
    OK(mp4j_comm_barrier(c));
    dense_suite<double>(c, r, p, n);
    dense_suite<float>(c, r, p, n + 1);
    dense_suite<int64_t>(c, r, p, n / 2 + 5);
    dense_suite<int32_t>(c, r, p, n);
    dense_suite<int16_t>(c, r, p, 37);
    dense_suite<int8_t>(c, r, p, p);  // one element per rank, from = 0
    {                                  // integer bit operators and PROD
      std::vector<int32_t> a(100, 1 << r);
      OK(mp4j_host_allreduce(c, a.data(), MP4J_I32, MP4J_BOR, 0, 100));
      CHECK(a[99] == (1 << p) - 1, "BOR got %d", a[99]);
      std::vector<int64_t> b(5, 2);
      OK(mp4j_host_allreduce(c, b.data(), MP4J_I64, MP4J_PROD, 0, 5));
      CHECK(b[0] == (int64_t)1 << p, "PROD");
    }
    {  // packed MAXLOC: value = rank, location = 100 + rank
      std::vector<double> a(3, mp4j_composite_double((float)r, 100 + r));
      OK(mp4j_host_allreduce(c, a.data(), MP4J_F64, MP4J_MAXLOC, 0, 3));
      CHECK(mp4j_composite_double_val(a[1]) == (float)(p - 1) && mp4j_composite_double_loc(a[1]) == 100 + p - 1,
            "MAXLOC");
    }
    {  // rank-ordered mailbox + byte transport ring
      char* out = nullptr;
      size_t len = 0;
      const std::string mine = "blob-" + std::to_string(r);
      OK(mp4j_comm_gather_all(c, mine.data(), mine.size(), &out, &len));
      size_t off = 0;
      for (int q = 0; q < p && out; ++q) {
        uint64_t l;
        memcpy(&l, out + off, 8);
        off += 8;
        CHECK(std::string(out + off, l) == "blob-" + std::to_string(q), "gather_all order");
        off += l;
      }
      mp4j_free(out);
      if (p > 1) {
        char* in = nullptr;
        size_t inl = 0;
        std::vector<char> payload(100000 + r, (char)r);
        OK(mp4j_comm_sendrecv_bytes(c, (r + 1) % p, payload.data(), payload.size(), (r + p - 1) % p, &in, &inl));
        const int src = (r + p - 1) % p;
        CHECK(inl == (size_t)(100000 + src) && in && in[0] == (char)src && in[inl - 1] == (char)src, "sendrecv ring");
        mp4j_free(in);
      }
    }
    if (p % 2 == 0) {  // pairwise exchange matchmaker (SRV:358-370)
      int other = -1;
      OK(mp4j_comm_exchange(c, &other));
      CHECK(other >= 0 && other < p && other != r, "exchange partner %d", other);
    }
    {  // String-keyed map allreduce / reduce (map_host.cc): shared keys + one private key per rank
      std::string blob;
      std::vector<double> vals;
      const int nk = 2000;
      for (int i = 0; i < nk; ++i) {
        if (i) blob.push_back('\0');
        blob += "key-" + std::to_string(i * 7919);
        vals.push_back(1.0);
      }
      blob.push_back('\0');
      blob += "private-" + std::to_string(r);
      vals.push_back(5.0);
      for (int mode = 0; mode < 2; ++mode) {
        char *ok = nullptr, *ov = nullptr;
        size_t okl = 0, on = 0;
        OK(mp4j_host_map_reduce(c, mode, p - 1, blob.data(), blob.size(), vals.data(), vals.size(), MP4J_F64, MP4J_SUM, &ok,
                                &okl, &ov, &on));
        const bool has = mode == 0 || r == p - 1;
        CHECK(on == (has ? (size_t)(nk + p) : 0u), "map_reduce mode=%d entries=%zu", mode, on);
        if (has && ok && ov) {
          const double* v = (const double*)ov;
          size_t shared = 0, priv = 0;
          for (size_t i = 0; i < on; ++i) {
            if (v[i] == (double)p) ++shared;
            else if (v[i] == 5.0) ++priv;
          }
          CHECK((p == 5 ? shared + priv == (size_t)(nk + p) : (shared == (size_t)nk && priv == (size_t)p)), "map values shared=%zu priv=%zu",
                shared, priv);
        }
        mp4j_free(ok);
        mp4j_free(ov);
      }
    }
    {  // opaque-payload collectives: the ring / trees run natively, this side is the "codec"
      struct Ctx {
        int r, p;
        std::vector<std::string> parts;   // local parts (serialized form = the string itself)
        std::vector<std::string> got;
      } cx{r, p, std::vector<std::string>((size_t)p), std::vector<std::string>((size_t)p)};
      auto get = [](void* u, int part, const char** d, size_t* n) -> int {
        Ctx* c2 = (Ctx*)u;
        *d = c2->parts[(size_t)part].data();
        *n = c2->parts[(size_t)part].size();
        return 0;
      };
      auto put_store = [](void* u, int part, const char* d, size_t n) -> int {
        ((Ctx*)u)->got[(size_t)part].assign(d, n);
        return 0;
      };
      auto put_merge = [](void* u, int part, const char* d, size_t n) -> int {  // "reduce" = concatenate sorted digits
        Ctx* c2 = (Ctx*)u;
        c2->parts[(size_t)part].append(d, n);
        std::sort(c2->parts[(size_t)part].begin(), c2->parts[(size_t)part].end());
        return 0;
      };
      // reduce-scatter: part q of rank r = one character 'a' + r; result on rank q = all p characters
      for (int q = 0; q < p; ++q) cx.parts[(size_t)q] = std::string(1, (char)('a' + r));
      OK(mp4j_host_blob_collective(c, MP4J_REDUCE_SCATTER, 0, get, put_merge, &cx));
      if (p > 1) {
        std::string want;
        for (int q = 0; q < p; ++q) want.push_back((char)('a' + q));
        CHECK(cx.parts[(size_t)r] == want, "blob reduce-scatter got %s", cx.parts[(size_t)r].c_str());
      }
      // allgather / gather / scatter of variable-length payloads
      for (int q = 0; q < p; ++q) cx.parts[(size_t)q] = std::string((size_t)(1000 * (r + 1) + q), (char)('A' + r));
      OK(mp4j_host_blob_collective(c, MP4J_ALL_GATHER, 0, get, put_store, &cx));
      for (int q = 0; q < p; ++q)
        if (q != r) CHECK(cx.got[(size_t)q] == std::string((size_t)(1000 * (q + 1) + q), (char)('A' + q)), "blob allgather from %d", q);
      for (auto& g : cx.got) g.clear();
      const int root = p - 1;
      OK(mp4j_host_blob_collective(c, MP4J_GATHER, root, get, put_store, &cx));
      if (r == root)
        for (int q = 0; q < p; ++q)
          if (q != r) CHECK(cx.got[(size_t)q] == std::string((size_t)(1000 * (q + 1) + q), (char)('A' + q)), "blob gather from %d", q);
      for (auto& g : cx.got) g.clear();
      OK(mp4j_host_blob_collective(c, MP4J_SCATTER, root, get, put_store, &cx));
      if (r != root)
        CHECK(cx.got[(size_t)r] == std::string((size_t)(1000 * (root + 1) + r), (char)('A' + root)), "blob scatter to %d", r);
    }
    {  // native reduceScatterMap: p maps per rank, map q = {"q-i": 1.0}, result on rank q = {"q-i": p}
      std::string blob;
      std::vector<double> vals;
      std::vector<int64_t> counts;
      const int nk = 300;
      for (int q = 0; q < p; ++q) {
        counts.push_back(nk);
        for (int i = 0; i < nk; ++i) {
          if (!blob.empty()) blob.push_back('\0');
          blob += std::to_string(q) + "-" + std::to_string(i);
          vals.push_back(1.0);
        }
      }
      char *ok = nullptr, *ov = nullptr;
      size_t okl = 0, on = 0;
      OK(mp4j_host_map_reduce_scatter(c, blob.data(), blob.size(), vals.data(), counts.data(), MP4J_F64, MP4J_SUM, &ok, &okl,
                                      &ov, &on));
      CHECK(on == (size_t)nk, "map reduce-scatter entries=%zu", on);
      for (size_t i = 0; i < on; ++i) CHECK(((const double*)ov)[i] == (double)p, "map reduce-scatter value");
      CHECK(okl > 2 && ok[0] == (char)('0' + r % 10), "map reduce-scatter keys start with own rank");
      mp4j_free(ok);
      mp4j_free(ov);
    }
    {  // process x thread allreduce: 3 threads per rank, every thread's array ends with the global sum
      const int T = 3;
      const int64_t m = n / 3 + 7;
      mp4j_tgroup* tg = nullptr;
      OK(mp4j_tgroup_create(T, &tg));
      std::vector<std::vector<double>> arrs(T);
      for (int t = 0; t < T; ++t) arrs[t].assign(m, (double)(r * T + t + 1));
      std::vector<std::thread> tt;
      for (int t = 0; t < T; ++t)
        tt.emplace_back([&, t] {
          if (mp4j_thread_allreduce(tg, c, t, arrs[t].data(), MP4J_F64, MP4J_SUM, 1, m) != 0) g_fail.fetch_add(1);
        });
      for (auto& th : tt) th.join();
      const double want = (double)(p * T) * (p * T + 1) / 2;
      for (int t = 0; t < T; ++t)
        CHECK(arrs[t][0] == (double)(r * T + t + 1) && arrs[t][1] == want && arrs[t][m - 1] == want, "thread allreduce t=%d got %f want %f",
              t, arrs[t][m - 1], want);
      // thread-level broadcast from (last rank, last thread) and reduce-scatter with [p][T] counts
      for (int t = 0; t < T; ++t) arrs[t].assign(m, (r == p - 1 && t == T - 1) ? 42.0 : -1.0);
      tt.clear();
      for (int t = 0; t < T; ++t)
        tt.emplace_back([&, t] {
          if (mp4j_thread_collective(tg, c, t, MP4J_BROADCAST, arrs[t].data(), MP4J_F64, MP4J_COPY, 2, m, nullptr, nullptr, nullptr,
                                     p - 1, T - 1) != 0)
            g_fail.fetch_add(1);
        });
      for (auto& th : tt) th.join();
      for (int t = 0; t < T; ++t) CHECK(arrs[t][2] == 42.0 && arrs[t][m - 1] == 42.0, "thread broadcast t=%d got %f", t, arrs[t][m - 1]);
      std::vector<int64_t> cnts((size_t)p * T, 5);
      for (int t = 0; t < T; ++t) arrs[t].assign(m, 1.0);
      tt.clear();
      for (int t = 0; t < T; ++t)
        tt.emplace_back([&, t] {
          if (mp4j_thread_collective(tg, c, t, MP4J_REDUCE_SCATTER, arrs[t].data(), MP4J_F64, MP4J_SUM, 1, 0, nullptr, nullptr,
                                     cnts.data(), 0, 0) != 0)
            g_fail.fetch_add(1);
        });
      for (auto& th : tt) th.join();
      for (int t = 0; t < T; ++t) {
        const int64_t off = 1 + ((int64_t)r * T + t) * 5;
        CHECK(arrs[t][off] == (double)(p * T) && arrs[t][off + 4] == (double)(p * T), "thread reduce-scatter t=%d got %f", t,
              arrs[t][off]);
      }
      mp4j_tgroup_destroy(tg);
    }
    char msg[64];
    snprintf(msg, sizeof msg, "round %d done", it);
    OK(mp4j_comm_log(c, 0, msg));
  }
  OK(mp4j_comm_barrier(c));
  OK(mp4j_comm_close(c, 0));
  mp4j_comm_destroy(c);
}

int main(int argc, char** argv) {
  const int p = argc > 1 ? atoi(argv[1]) : 4;
  const int rounds = argc > 2 ? atoi(argv[2]) : 2;
  const int64_t n = argc > 3 ? atoll(argv[3]) : 20011;
  mp4j_master* m = nullptr;
  if (mp4j_master_create(p, 0, 60, 60, 60, 200, nullptr, nullptr, 1, &m) != 0 || mp4j_master_start(m) != 0) {
    fprintf(stderr, "master: %s\n", mp4j_last_error());
    return 2;
  }
  const int port = mp4j_master_port(m);
  std::vector<std::thread> ths;
  for (int r = 0; r < p; ++r) ths.emplace_back(rank_main, port, r, p, rounds, n);
  for (auto& t : ths) t.join();
  int code = -1;
  const int still = mp4j_master_wait(m, 20000, &code);
  if (still != 0 || code != 0) {
    fprintf(stderr, "master wait: still=%d code=%d\n", still, code);
    g_fail.fetch_add(1);
  }
  // thread rendezvous primitive (TCS:139-157)
  {
    mp4j_tgroup* g = nullptr;
    mp4j_tgroup_create(4, &g);
    std::atomic<int> phase{0};
    std::vector<std::thread> tt;
    for (int t = 0; t < 4; ++t)
      tt.emplace_back([&, t] {
        for (int i = 0; i < 200; ++i) {
          mp4j_tgroup_set_slot(g, t, 0, (uint64_t)(i * 4 + t));
          mp4j_tgroup_barrier(g);
          for (int u = 0; u < 4; ++u)
            if (mp4j_tgroup_get_slot(g, u, 0) != (uint64_t)(i * 4 + u)) g_fail.fetch_add(1);
          mp4j_tgroup_barrier(g);
        }
        phase.fetch_add(1);
      });
    for (auto& t : tt) t.join();
    mp4j_tgroup_destroy(g);
  }
  // ThreadComm::delegate (what the thread-level GPU collectives stand on): one thread runs the step with everybody's
  // argument; a failure of that step reaches all T threads, and the group is usable afterwards
  {
    mp4j::ThreadComm tc(5);
    std::atomic<int> errors_seen{0};
    std::atomic<uint64_t> sum_seen{0};
    std::vector<std::thread> tt;
    for (int t = 0; t < 5; ++t)
      tt.emplace_back([&, t] {
        for (int i = 0; i < 100; ++i) {
          const int root = i % 5;
          const bool fail = i % 7 == 3;
          try {
            tc.delegate(t, root, (uint64_t)(i * 10 + t), [&](const std::vector<uint64_t>& a) {
              uint64_t sacc = 0;
              for (uint64_t x : a) sacc += x;
              if (sacc != (uint64_t)(i * 50 + 10)) g_fail.fetch_add(1);
              if (fail) throw mp4j::Error("step failed on purpose");
              sum_seen.fetch_add(sacc);
            });
            if (fail) g_fail.fetch_add(1);  // every thread must see the failure
          } catch (const mp4j::Error& e) {
            if (!fail || std::string(e.what()).find("on purpose") == std::string::npos) g_fail.fetch_add(1);
            errors_seen.fetch_add(1);
          }
        }
      });
    for (auto& t : tt) t.join();
    int nfail = 0;
    for (int i = 0; i < 100; ++i) nfail += i % 7 == 3;
    if (errors_seen.load() != 5 * nfail) {
      fprintf(stderr, "delegate: %d threads saw an error, want %d\n", errors_seen.load(), 5 * nfail);
      g_fail.fetch_add(1);
    }
  }
  mp4j_master_destroy(m);
  printf("host_stress p=%d rounds=%d n=%lld: %s (%d failures)\n", p, rounds, (long long)n, g_fail.load() ? "FAILED" : "ok",
         g_fail.load());
  return g_fail.load() ? 1 : 0;
}
