#include "process_comm.h"

#include <stdlib.h>
#include <sys/socket.h>
#include <unistd.h>

#include <algorithm>
#include <exception>
#include <thread>

namespace mp4j {

namespace {
// The team the calling thread belongs to while it executes the process step of a thread-level collective
// (ProcessComm::TeamScope); null for an ordinary call.
thread_local const ProcessComm::Team* tl_team = nullptr;

// op_mu_ serialises the collectives of a communicator. A team is ONE collective: its thread 0 takes the lock, the
// followers do not (they would wait for thread 0, which waits for them at the team's first thread barrier).
struct OpGuard {
  std::mutex& mu;
  const bool own;
  explicit OpGuard(std::mutex& m) : mu(m), own(!(tl_team && tl_team->tid != 0)) {
    if (own) mu.lock();
  }
  ~OpGuard() {
    if (own) mu.unlock();
  }
};
bool team_follower() { return tl_team && tl_team->tid != 0; }

// Small messages on the TCP paths skip the ring / tree: every rank talks to every peer directly over the mesh (one
// network latency instead of p - 1 or 2(p - 1) dependent hops), and reductions run in rank order on the receiver -
// the TCP twin of the GPU plane's one-shot kernel. MP4J_TCP_ONESHOT_MAX_BYTES (0 = off), up to
// MP4J_TCP_ONESHOT_MAX_RANKS ranks (a full mesh costs 2(p - 1) sockets per rank; the ring needs two).
size_t env_sz(const char* name, size_t dflt) {
  const char* e = getenv(name);
  return e ? (size_t)strtoull(e, nullptr, 10) : dflt;
}
const size_t kOneShotMaxBytes = env_sz("MP4J_TCP_ONESHOT_MAX_BYTES", 32 << 10);
const size_t kOneShotMaxRanks = env_sz("MP4J_TCP_ONESHOT_MAX_RANKS", 32);

constexpr size_t kChunk = 1 << 20;  // pipeline granularity of reduce-on-receive
// ring steps moving at least this much send on a helper thread (MP4J_TCP_DUPLEX_MIN_BYTES; 0 = never)
const size_t kDuplexMin = [] {
  const char* e = getenv("MP4J_TCP_DUPLEX_MIN_BYTES");
  const size_t v = e ? (size_t)strtoull(e, nullptr, 10) : (size_t)(4 << 20);
  return v ? v : (size_t)-1;
}();

// One full-duplex ring step: `send_len` bytes leave through `send_fd` on a helper thread while `receive` runs on the
// calling thread. If either side fails the socket is shut down so that the other side wakes up (the job is over
// anyway: any failed collective ends it), and the first error is rethrown.
template <typename F>
void duplex_step(int send_fd, const void* send_buf, size_t send_len, F&& receive) {
  std::exception_ptr send_err;
  std::thread sender([&] {
    try {
      if (send_len) net::send_all(send_fd, send_buf, send_len);
    } catch (...) {
      send_err = std::current_exception();
    }
  });
  try {
    receive();
  } catch (...) {
    ::shutdown(send_fd, SHUT_RDWR);
    sender.join();
    throw;
  }
  sender.join();
  if (send_err) std::rethrow_exception(send_err);
}

int env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return v && *v ? atoi(v) : dflt;
}
}  // namespace

ProcessComm::ProcessComm(const std::string& login, const std::string& master_host, int master_port,
                         int rank_hint, const std::string& my_host) {
  master_.reset(new MasterClient(master_host, master_port));
  listen_fd_ = net::tcp_listen(0, &listen_port_);
  std::string host = my_host.empty() ? (getenv("MP4J_HOST") ? getenv("MP4J_HOST") : net::local_hostname())
                                     : my_host;
  std::string me = host + kAddrDelim + std::to_string(listen_port_);
  rank_ = master_->rendezvous(me, rank_hint, &addrs_);
  master_->open_local_log(rank_);
  n_ = (int)addrs_.size();
  for (auto& a : addrs_) {
    size_t k = a.rfind(kAddrDelim);
    hosts_.push_back(a.substr(0, k));
    ports_.push_back(atoi(a.c_str() + k + 3));
  }
  out_fds_.assign(n_, -1);
  in_fds_.assign(n_, -1);
  accept_thr_ = std::thread([this, lfd = listen_fd_] { accept_loop(lfd); });
  // kill script line: ssh login@host "kill -9 pid" (PCS:178-183)
  master_->kill_me(rank_, "ssh " + login + "@" + host + " \"kill -9 " + std::to_string((int)getpid()) + "\"");
  int period = env_int("MP4J_HEARTBEAT_MS", 15000);
  master_->start_heartbeat(rank_, std::min(period, 5000), period, 4, nullptr, nullptr);
  everyone_.resize((size_t)n_);
  for (int q = 0; q < n_; ++q) everyone_[(size_t)q] = q;
  shm_ = ShmPlane::create(*master_, rank_, n_);  // collective; null on every rank if unusable
  if (!shm_) setup_node_plane();
}

// Several machines with several ranks each (8 GPUs per node, one process per GPU): no job-wide shared-memory plane,
// but every machine can have its own. Used by the hierarchical allreduce: reduce-scatter inside the machine, TCP
// between the ranks that hold the same shard on every machine, allgather inside the machine. Needs the same number
// (>= 2) of ranks on every machine; MP4J_HIER=0 turns it off. Collective; every rank reaches the same verdict.
void ProcessComm::setup_node_plane() {
  if (n_ < 4 || env_int("MP4J_HIER", 1) == 0 || env_int("MP4J_SHM", 1) == 0) return;  // same env on every rank
  const std::string key = ShmPlane::machine_key();
  const std::vector<std::string> keys = master_->gather_all(rank_, key.data(), key.size());
  if ((int)keys.size() != n_) return;
  // machines in the order of their lowest rank; members ascending
  std::vector<std::string> order;
  std::vector<std::vector<int>> nodes;
  for (int q = 0; q < n_; ++q) {
    size_t j = 0;
    while (j < order.size() && order[j] != keys[(size_t)q]) ++j;
    if (j == order.size()) {
      order.push_back(keys[(size_t)q]);
      nodes.emplace_back();
    }
    nodes[j].push_back(q);
  }
  const size_t L = nodes[0].size();
  if (nodes.size() < 2 || L < 2 || L > 32) return;
  for (const auto& nd : nodes)
    if (nd.size() != L) return;
  int ni = -1, li = -1;
  for (size_t j = 0; j < nodes.size(); ++j)
    for (size_t i = 0; i < L; ++i)
      if (nodes[j][i] == rank_) ni = (int)j, li = (int)i;
  std::vector<int> group = nodes[(size_t)ni], rail;
  for (const auto& nd : nodes) rail.push_back(nd[(size_t)li]);
  node_shm_ = ShmPlane::create(*master_, rank_, n_, &group);  // collective over the job
  if (node_shm_) {
    node_group_ = group;
    rail_group_ = rail;
    node_index_ = ni;
    nodes_ = nodes;
  }
}

ProcessComm::~ProcessComm() {
  master_->stop_heartbeat();
  // wake the acceptor, let it leave, and only then release the descriptor (it holds its own copy of
  // the number: closing first would let the kernel hand the same number to somebody else)
  if (listen_fd_ >= 0) net::shutdown_fd(listen_fd_);
  if (accept_thr_.joinable()) accept_thr_.join();
  if (listen_fd_ >= 0) {
    net::close_fd(listen_fd_);
    listen_fd_ = -1;
  }
  for (int fd : out_fds_) net::close_fd(fd);
  for (int fd : in_fds_) net::close_fd(fd);
}

void ProcessComm::accept_loop(int listen_fd) {
  for (;;) {
    int fd = net::tcp_accept(listen_fd);
    if (fd < 0) return;
    int32_t src = -1;
    try {
      if (!net::recv_all(fd, &src, 4) || src < 0 || src >= n_) {
        net::close_fd(fd);
        continue;
      }
    } catch (const std::exception&) {
      net::close_fd(fd);
      continue;
    }
    std::lock_guard<std::mutex> lk(mesh_mu_);
    in_fds_[src] = fd;
    mesh_cv_.notify_all();
  }
}

int ProcessComm::out_fd(int dest) {
  {
    std::lock_guard<std::mutex> lk(mesh_mu_);
    if (out_fds_[dest] >= 0) return out_fds_[dest];
  }
  int fd = net::tcp_connect(hosts_[dest], ports_[dest], 200, 30);
  int32_t me = rank_;
  net::send_all(fd, &me, 4);
  std::lock_guard<std::mutex> lk(mesh_mu_);
  out_fds_[dest] = fd;
  return fd;
}

int ProcessComm::in_fd(int src) {
  std::unique_lock<std::mutex> lk(mesh_mu_);
  int tmo = env_int("MP4J_PEER_TIMEOUT_S", 600);
  if (!mesh_cv_.wait_for(lk, std::chrono::seconds(tmo), [&] { return in_fds_[src] >= 0; }))
    throw Error("timeout waiting for a data connection from rank " + std::to_string(src));
  return in_fds_[src];
}

void ProcessComm::barrier() {
  if (n_ == 1) return;
  master_->barrier();
}

void ProcessComm::log(LogLevel lvl, const std::string& msg) { master_->log(lvl, msg); }

void ProcessComm::close(int code) {
  if (closed_) return;
  closed_ = true;
  master_->stop_heartbeat();
  master_->close(rank_, code);
}

// ------------------------------------------------------------------ host collectives

ProcessComm::TeamScope::TeamScope(const Team* t) { tl_team = (t && t->T > 1) ? t : nullptr; }
ProcessComm::TeamScope::~TeamScope() { tl_team = nullptr; }

bool ProcessComm::tcp_oneshot(size_t bytes, int members) const {
  if (members < 0) members = n_;
  return !shm_ && kOneShotMaxBytes > 0 && bytes <= kOneShotMaxBytes && (size_t)members <= kOneShotMaxRanks;
}

// `bytes` of mine to every peer, every peer's `bytes` into all + q * bytes (sends first: messages this small fit
// the socket buffers, so no send waits for a receive)
void ProcessComm::exchange_small(const void* mine, size_t bytes, char* all) {
  if (bytes == 0) return;
  for (int s = 1; s < n_; ++s) net::send_all(out_fd((rank_ + s) % n_), mine, bytes);
  for (int s = 1; s < n_; ++s) {
    const int src = (rank_ - s + n_) % n_;
    if (!net::recv_all(in_fd(src), all + (size_t)src * bytes, bytes)) throw Error("peer closed the data connection");
  }
}

void ProcessComm::shm_run(void* arr, int dtype, int op, const int64_t* lo, const int64_t* hi, const uint32_t* src_mask,
                          const uint32_t* dst_mask) {
  if (tl_team) shm_->grs_team(arr, dtype, op, lo, hi, src_mask, dst_mask, tl_team->tid, tl_team->T, tl_team->sync);
  else shm_->grs(arr, dtype, op, lo, hi, src_mask, dst_mask);
}

// The TCP reduce-scatter and allgather over a GROUP of ranks (grp[i] = global rank of member i, `me` = my index):
// the whole job for the flat collectives, the ranks with my local index on every machine for the inter-machine step
// of the hierarchical allreduce. froms / tos are indexed by member. Small totals go directly between all pairs,
// otherwise the ring (big steps with the send on a helper thread).
void ProcessComm::tcp_reduce_scatter(char* b, int dtype, int op, const int64_t* froms, const int64_t* tos,
                                     const std::vector<int>& grp, int me) {
  const int g = (int)grp.size();
  const size_t es = dtype_size(dtype);
  if (g == 1) return;
  size_t total_bytes = 0;
  for (int q = 0; q < g; ++q) total_bytes += (size_t)(tos[q] - froms[q]) * es;
  if (tcp_oneshot(total_bytes, g)) {
    // small: slice q goes straight to its owner q; the owner folds the contributions in member order
    for (int s = 1; s < g; ++s) {
      const int q = (me + s) % g;
      const size_t nb = (size_t)(tos[q] - froms[q]) * es;
      if (nb) net::send_all(out_fd(grp[q]), b + froms[q] * es, nb);
    }
    const size_t mine = (size_t)(tos[me] - froms[me]) * es;
    if (mine == 0) return;
    std::vector<char> acc(mine), in(mine);
    for (int q = 0; q < g; ++q) {
      const char* x = b + froms[me] * es;
      if (q != me) {
        if (!net::recv_all(in_fd(grp[q]), in.data(), mine)) throw Error("peer closed the data connection");
        x = in.data();
      }
      if (q == 0) memcpy(acc.data(), x, mine);
      else host_reduce_into(dtype, op, acc.data(), x, mine / es);
    }
    memcpy(b + froms[me] * es, acc.data(), mine);
    return;
  }
  const int sfd = out_fd(grp[(me + 1) % g]), rfd = in_fd(grp[(me - 1 + g) % g]);
  if (scratch_.size() < kChunk) scratch_.resize(kChunk);
  for (int s = 0; s < g - 1; ++s) {
    const int sseg = ((me - 1 - s) % g + g) % g;
    const int rseg = ((me - 2 - s) % g + g) % g;
    const size_t sbytes = (size_t)(tos[sseg] - froms[sseg]) * es, rbytes = (size_t)(tos[rseg] - froms[rseg]) * es;
    char* sp = b + froms[sseg] * es;
    char* rp = b + froms[rseg] * es;
    size_t so = 0, ro = 0;
    const size_t chunk = kChunk / es * es;
    if (sbytes + rbytes >= kDuplexMin) {
      // big step: the outgoing segment (reduced in the previous step, not touched in this one) leaves on a helper
      // thread while this thread receives and reduces chunk by chunk - two cores move the bytes instead of one
      // poll loop doing send copy, receive copy and the reduction in turn
      duplex_step(sfd, sp, sbytes, [&] {
        for (; ro < rbytes; ro += chunk) {
          const size_t rl = std::min(chunk, rbytes - ro);
          if (!net::recv_all(rfd, scratch_.data(), rl)) throw Error("peer closed during reduce-scatter");
          host_reduce_into(dtype, op, rp + ro, scratch_.data(), rl / es);  // op(local, incoming) DOP:196
        }
      });
      continue;
    }
    while (so < sbytes || ro < rbytes) {
      const size_t sl = std::min(chunk, sbytes - so), rl = std::min(chunk, rbytes - ro);
      net::sendrecv(sfd, sp + so, sl, rfd, scratch_.data(), rl);
      if (rl) host_reduce_into(dtype, op, rp + ro, scratch_.data(), rl / es);  // op(local, incoming) DOP:196
      so += sl;
      ro += rl;
    }
  }
}

void ProcessComm::tcp_allgather(char* b, size_t es, const int64_t* froms, const int64_t* tos,
                                const std::vector<int>& grp, int me) {
  const int g = (int)grp.size();
  if (g == 1) return;
  size_t total = 0;
  for (int q = 0; q < g; ++q) total += (size_t)(tos[q] - froms[q]) * es;
  if (tcp_oneshot(total, g)) {  // small: my slice to every peer, every peer's slice straight into place
    const size_t mine = (size_t)(tos[me] - froms[me]) * es;
    if (mine)
      for (int s = 1; s < g; ++s) net::send_all(out_fd(grp[(me + s) % g]), b + froms[me] * es, mine);
    for (int s = 1; s < g; ++s) {
      const int q = (me - s + g) % g;
      const size_t nb = (size_t)(tos[q] - froms[q]) * es;
      if (nb && !net::recv_all(in_fd(grp[q]), b + froms[q] * es, nb)) throw Error("peer closed the data connection");
    }
    return;
  }
  const int sfd = out_fd(grp[(me + 1) % g]), rfd = in_fd(grp[(me - 1 + g) % g]);
  for (int s = 0; s < g - 1; ++s) {
    const int sseg = ((me - s) % g + g) % g;
    const int rseg = ((me - 1 - s) % g + g) % g;
    const size_t sbytes = (size_t)(tos[sseg] - froms[sseg]) * es, rbytes = (size_t)(tos[rseg] - froms[rseg]) * es;
    if (sbytes + rbytes >= kDuplexMin) {
      duplex_step(sfd, b + froms[sseg] * es, sbytes, [&] {
        if (rbytes && !net::recv_all(rfd, b + froms[rseg] * es, rbytes)) throw Error("peer closed during allgather");
      });
      continue;
    }
    net::sendrecv(sfd, b + froms[sseg] * es, sbytes, rfd, b + froms[rseg] * es, rbytes);
  }
}

void ProcessComm::reduce_scatter(void* arr, int dtype, int op, int64_t from, const int64_t* counts) {
  check_from_counts(from, counts, n_);
  if (n_ == 1) return;  // PCS:1179
  if (!op_valid_for(dtype, op) || op == MP4J_COPY) throw Error("invalid operator for dtype");
  OpGuard lk(op_mu_);
  const size_t es = dtype_size(dtype);
  std::vector<int64_t> froms(n_), tos(n_);
  froms_tos_from_counts(from, counts, n_, froms.data(), tos.data());
  if (shm_) {
    std::vector<uint32_t> sm(n_, all_mask()), dm(n_);
    for (int q = 0; q < n_; ++q) dm[q] = 1u << q;
    return shm_run(arr, dtype, op, froms.data(), tos.data(), sm.data(), dm.data());
  }
  if (team_follower()) return;  // the ring is thread 0's
  if (node_shm_ && !tcp_oneshot((size_t)(tos[n_ - 1] - froms[0]) * es)) {
    // hierarchical: for every machine j, its ranks' slices are reduced over MY machine in shared memory (local rank i
    // takes the slice of machine j's local rank i); then the partial sums travel between the machines: on the rail
    // of local index i, member j owns the slice of machine j's local rank i
    const int L = node_shm_->members(), li = node_shm_->index(), m = (int)nodes_.size();
    const uint32_t all_l = L >= 32 ? 0xffffffffu : ((1u << L) - 1u);
    std::vector<int64_t> lf(L), lt(L), rf(m), rt(m);
    std::vector<uint32_t> sm(L, all_l), dm(L);
    for (int q = 0; q < L; ++q) dm[q] = 1u << q;
    for (int j = 0; j < m; ++j) {
      for (int i = 0; i < L; ++i) lf[i] = froms[nodes_[j][i]], lt[i] = tos[nodes_[j][i]];
      node_shm_->grs(arr, dtype, op, lf.data(), lt.data(), sm.data(), dm.data());
      rf[j] = froms[nodes_[j][li]];
      rt[j] = tos[nodes_[j][li]];
    }
    tcp_reduce_scatter((char*)arr, dtype, op, rf.data(), rt.data(), rail_group_, node_index_);
    return;
  }
  tcp_reduce_scatter((char*)arr, dtype, op, froms.data(), tos.data(), everyone_, rank_);
}

void ProcessComm::allgather(void* arr, int dtype, const int64_t* froms, const int64_t* tos) {
  check_froms_tos(froms, tos, n_);
  if (n_ == 1) return;  // PCS:630
  OpGuard lk(op_mu_);
  const size_t es = dtype_size(dtype);
  if (shm_) {
    std::vector<uint32_t> sm(n_), dm(n_, all_mask());
    for (int q = 0; q < n_; ++q) sm[q] = 1u << q;
    return shm_run(arr, dtype, MP4J_COPY, froms, tos, sm.data(), dm.data());
  }
  if (team_follower()) return;  // the ring is thread 0's
  {
    size_t total = 0;
    for (int q = 0; q < n_; ++q) total += (size_t)(tos[q] - froms[q]) * es;
    if (node_shm_ && !tcp_oneshot(total)) {
      // hierarchical: first between the machines (on my rail I collect the slices of every machine's local rank
      // li), then inside the machine in shared memory, one pass per source machine
      const int L = node_shm_->members(), li = node_shm_->index(), m = (int)nodes_.size();
      const uint32_t all_l = L >= 32 ? 0xffffffffu : ((1u << L) - 1u);
      std::vector<int64_t> lf(L), lt(L), rf(m), rt(m);
      for (int j = 0; j < m; ++j) rf[j] = froms[nodes_[j][li]], rt[j] = tos[nodes_[j][li]];
      tcp_allgather((char*)arr, es, rf.data(), rt.data(), rail_group_, node_index_);
      std::vector<uint32_t> sm(L), dm(L, all_l);
      for (int q = 0; q < L; ++q) sm[q] = 1u << q;
      for (int j = 0; j < m; ++j) {
        for (int i = 0; i < L; ++i) lf[i] = froms[nodes_[j][i]], lt[i] = tos[nodes_[j][i]];
        node_shm_->grs(arr, dtype, MP4J_COPY, lf.data(), lt.data(), sm.data(), dm.data());
      }
      return;
    }
  }
  tcp_allgather((char*)arr, es, froms, tos, everyone_, rank_);
}

void ProcessComm::gather(void* arr, int dtype, const int64_t* froms, const int64_t* tos, int root) {
  check_froms_tos(froms, tos, n_);
  if (root < 0 || root >= n_) throw Error("root rank out of range");
  if (n_ == 1) return;  // PCS:545
  OpGuard lk(op_mu_);
  const size_t es = dtype_size(dtype);
  if (shm_) {
    std::vector<uint32_t> sm(n_), dm(n_, 1u << root);
    for (int q = 0; q < n_; ++q) sm[q] = 1u << q;
    return shm_run(arr, dtype, MP4J_COPY, froms, tos, sm.data(), dm.data());
  }
  if (team_follower()) return;  // the tree is thread 0's
  {
    size_t total = 0;
    for (int q = 0; q < n_; ++q) total += (size_t)(tos[q] - froms[q]) * es;
    if (tcp_oneshot(total)) {  // small: every rank hands its slice to the root directly (one hop, no relays)
      char* b = (char*)arr;
      if (rank_ != root) {
        const size_t nb = (size_t)(tos[rank_] - froms[rank_]) * es;
        if (nb) net::send_all(out_fd(root), b + froms[rank_] * es, nb);
        return;
      }
      for (int q = 0; q < n_; ++q) {
        const size_t nb = (size_t)(tos[q] - froms[q]) * es;
        if (q != root && nb && !net::recv_all(in_fd(q), b + froms[q] * es, nb)) throw Error("peer closed during gather");
      }
      return;
    }
  }
  char* base = (char*)arr;
  int vr = (rank_ - root + n_) % n_;
  auto real = [&](int v) { return (v + root) % n_; };
  for (int m = 1; m < n_; m <<= 1) {
    if (vr & m) {  // hand everything accumulated so far to the parent
      int fd = out_fd(real(vr - m));
      for (int v = vr; v < std::min(vr + m, n_); ++v) {
        int q = real(v);
        net::send_all(fd, base + froms[q] * es, (size_t)(tos[q] - froms[q]) * es);
      }
      break;
    } else if (vr + m < n_) {
      int fd = in_fd(real(vr + m));
      for (int v = vr + m; v < std::min(vr + 2 * m, n_); ++v) {
        int q = real(v);
        size_t nb = (size_t)(tos[q] - froms[q]) * es;
        if (nb && !net::recv_all(fd, base + froms[q] * es, nb)) throw Error("peer closed during gather");
      }
    }
  }
}

void ProcessComm::scatter(void* arr, int dtype, const int64_t* froms, const int64_t* tos, int root) {
  check_froms_tos(froms, tos, n_);
  if (root < 0 || root >= n_) throw Error("root rank out of range");
  if (n_ == 1) return;
  OpGuard lk(op_mu_);
  const size_t es = dtype_size(dtype);
  if (shm_) {
    std::vector<uint32_t> sm(n_, 1u << root), dm(n_);
    for (int q = 0; q < n_; ++q) dm[q] = 1u << q;
    return shm_run(arr, dtype, MP4J_COPY, froms, tos, sm.data(), dm.data());
  }
  if (team_follower()) return;  // the tree is thread 0's
  {
    size_t total = 0;
    for (int q = 0; q < n_; ++q) total += (size_t)(tos[q] - froms[q]) * es;
    if (tcp_oneshot(total)) {  // small: the root hands every rank its slice directly (one hop, no relays)
      char* b = (char*)arr;
      if (rank_ == root) {
        for (int s = 1; s < n_; ++s) {
          const int q = (rank_ + s) % n_;
          const size_t nb = (size_t)(tos[q] - froms[q]) * es;
          if (nb) net::send_all(out_fd(q), b + froms[q] * es, nb);
        }
        return;
      }
      const size_t nb = (size_t)(tos[rank_] - froms[rank_]) * es;
      if (nb && !net::recv_all(in_fd(root), b + froms[rank_] * es, nb)) throw Error("peer closed during scatter");
      return;
    }
  }
  char* base = (char*)arr;
  int vr = (rank_ - root + n_) % n_;
  auto real = [&](int v) { return (v + root) % n_; };
  int top = 1;
  while (top < n_) top <<= 1;
  top >>= 1;
  int start = top;
  if (vr != 0) {
    int lsb = vr & -vr;
    int fd = in_fd(real(vr - lsb));
    for (int v = vr; v < std::min(vr + lsb, n_); ++v) {
      int q = real(v);
      size_t nb = (size_t)(tos[q] - froms[q]) * es;
      if (nb && !net::recv_all(fd, base + froms[q] * es, nb)) throw Error("peer closed during scatter");
    }
    start = lsb >> 1;
  }
  for (int m = start; m >= 1; m >>= 1) {
    if (vr + m >= n_) continue;
    int fd = out_fd(real(vr + m));
    for (int v = vr + m; v < std::min(vr + 2 * m, n_); ++v) {
      int q = real(v);
      net::send_all(fd, base + froms[q] * es, (size_t)(tos[q] - froms[q]) * es);
    }
  }
}

void ProcessComm::broadcast(void* arr, int dtype, int64_t from, int64_t to, int root) {
  check_from_to(from, to);
  if (n_ == 1) return;  // PCS:751
  std::vector<int64_t> froms(n_), tos(n_);
  balanced_split(from, to, n_, froms.data(), tos.data());  // PCS:760-767
  if (shm_) {  // one pass: everybody copies out of the root's window
    if (root < 0 || root >= n_) throw Error("root rank out of range");
    OpGuard lk(op_mu_);
    std::vector<uint32_t> sm(n_, 1u << root), dm(n_, all_mask());
    return shm_run(arr, dtype, MP4J_COPY, froms.data(), tos.data(), sm.data(), dm.data());
  }
  const size_t es = dtype_size(dtype), bytes = (size_t)(to - from) * es;
  if (tcp_oneshot(bytes)) {
    if (root < 0 || root >= n_) throw Error("root rank out of range");
    OpGuard lk(op_mu_);
    if (team_follower() || bytes == 0) return;
    char* mine = (char*)arr + from * es;
    if (rank_ == root) {
      for (int s = 1; s < n_; ++s) net::send_all(out_fd((rank_ + s) % n_), mine, bytes);
    } else if (!net::recv_all(in_fd(root), mine, bytes)) {
      throw Error("peer closed the data connection");
    }
    return;
  }
  scatter(arr, dtype, froms.data(), tos.data(), root);
  allgather(arr, dtype, froms.data(), tos.data());
}

void ProcessComm::reduce(void* arr, int dtype, int op, int64_t from, int64_t to, int root) {
  check_from_to(from, to);
  if (n_ == 1) return;
  std::vector<int64_t> froms(n_), tos(n_), counts(n_);
  balanced_split(from, to, n_, froms.data(), tos.data());
  for (int i = 0; i < n_; ++i) counts[i] = tos[i] - froms[i];
  if (shm_) {  // one pass: owners reduce, the root collects
    if (root < 0 || root >= n_) throw Error("root rank out of range");
    if (!op_valid_for(dtype, op) || op == MP4J_COPY) throw Error("invalid operator for dtype");
    OpGuard lk(op_mu_);
    std::vector<uint32_t> sm(n_, all_mask()), dm(n_, 1u << root);
    return shm_run(arr, dtype, op, froms.data(), tos.data(), sm.data(), dm.data());
  }
  const size_t es = dtype_size(dtype), bytes = (size_t)(to - from) * es;
  if (tcp_oneshot(bytes)) {
    if (root < 0 || root >= n_) throw Error("root rank out of range");
    if (!op_valid_for(dtype, op) || op == MP4J_COPY) throw Error("invalid operator for dtype");
    OpGuard lk(op_mu_);
    if (team_follower() || bytes == 0) return;
    char* mine = (char*)arr + from * es;
    if (rank_ != root) {
      net::send_all(out_fd(root), mine, bytes);
      return;
    }
    std::vector<char> acc(bytes), in(bytes);
    for (int q = 0; q < n_; ++q) {  // rank order
      const char* x = mine;
      if (q != root) {
        if (!net::recv_all(in_fd(q), in.data(), bytes)) throw Error("peer closed the data connection");
        x = in.data();
      }
      if (q == 0) memcpy(acc.data(), x, bytes);
      else host_reduce_into(dtype, op, acc.data(), x, (size_t)(to - from));
    }
    memcpy(mine, acc.data(), bytes);
    return;
  }
  reduce_scatter(arr, dtype, op, from, counts.data());
  gather(arr, dtype, froms.data(), tos.data(), root);
}

void ProcessComm::allreduce(void* arr, int dtype, int op, int64_t from, int64_t to) {
  check_from_to(from, to);
  if (n_ == 1) return;  // PCS:1734
  std::vector<int64_t> froms(n_), tos(n_), counts(n_);
  balanced_split(from, to, n_, froms.data(), tos.data());  // PCS:1741-1756
  for (int i = 0; i < n_; ++i) counts[i] = tos[i] - froms[i];
  if (shm_) {  // one pass: owners reduce, everybody collects
    if (!op_valid_for(dtype, op) || op == MP4J_COPY) throw Error("invalid operator for dtype");
    OpGuard lk(op_mu_);
    std::vector<uint32_t> sm(n_, all_mask()), dm(n_, all_mask());
    return shm_run(arr, dtype, op, froms.data(), tos.data(), sm.data(), dm.data());
  }
  const size_t es = dtype_size(dtype), bytes = (size_t)(to - from) * es;
  if (node_shm_ && !tcp_oneshot(bytes)) {
    // hierarchical: shared memory inside the machine, TCP only between the ranks that own the same shard
    if (!op_valid_for(dtype, op) || op == MP4J_COPY) throw Error("invalid operator for dtype");
    OpGuard lk(op_mu_);
    if (team_follower()) return;
    const int L = node_shm_->members(), li = node_shm_->index(), m = (int)rail_group_.size();
    const uint32_t all_l = L >= 32 ? 0xffffffffu : ((1u << L) - 1u);
    std::vector<int64_t> lf(L), lt(L), rf(m), rt(m);
    balanced_split(from, to, L, lf.data(), lt.data());
    std::vector<uint32_t> sm(L, all_l), dm(L);
    for (int q = 0; q < L; ++q) dm[q] = 1u << q;
    node_shm_->grs(arr, dtype, op, lf.data(), lt.data(), sm.data(), dm.data());  // reduce-scatter inside the machine
    balanced_split(lf[li], lt[li], m, rf.data(), rt.data());
    tcp_reduce_scatter((char*)arr, dtype, op, rf.data(), rt.data(), rail_group_, node_index_);  // my shard, across
    tcp_allgather((char*)arr, es, rf.data(), rt.data(), rail_group_, node_index_);               // the machines
    for (int q = 0; q < L; ++q) sm[q] = 1u << q, dm[q] = all_l;
    node_shm_->grs(arr, dtype, MP4J_COPY, lf.data(), lt.data(), sm.data(), dm.data());  // allgather inside the machine
    return;
  }
  if (tcp_oneshot(bytes)) {
    if (!op_valid_for(dtype, op) || op == MP4J_COPY) throw Error("invalid operator for dtype");
    OpGuard lk(op_mu_);
    if (team_follower() || bytes == 0) return;
    char* mine = (char*)arr + from * es;
    std::vector<char> all((size_t)n_ * bytes);
    memcpy(all.data() + (size_t)rank_ * bytes, mine, bytes);
    exchange_small(mine, bytes, all.data());
    for (int q = 1; q < n_; ++q) host_reduce_into(dtype, op, all.data(), all.data() + (size_t)q * bytes, (size_t)(to - from));
    memcpy(mine, all.data(), bytes);  // rank order on every rank: identical bits everywhere
    return;
  }
  reduce_scatter(arr, dtype, op, from, counts.data());
  allgather(arr, dtype, froms.data(), tos.data());
}

void ProcessComm::allreduce_rpc(void* arr, int dtype, int op, int64_t n) {
  if (n_ == 1) return;
  if (!op_valid_for(dtype, op) || op == MP4J_COPY) throw Error("invalid operator for dtype");
  const size_t es = dtype_size(dtype);
  auto parts = master_->gather_all(rank_, arr, (size_t)n * es);
  for (auto& p : parts)
    if (p.size() != (size_t)n * es) throw Error("allreduce_rpc: ranks passed arrays of different length");
  memcpy(arr, parts[0].data(), (size_t)n * es);  // rank order => identical bits on every rank
  for (int q = 1; q < n_; ++q) host_reduce_into(dtype, op, arr, parts[q].data(), (size_t)n);
}

// ------------------------------------------------------------------ byte transport

void ProcessComm::send_bytes(int dest, const void* data, size_t n) {
  if (dest < 0 || dest >= n_ || dest == rank_) throw Error("send_bytes: bad destination");
  int fd = out_fd(dest);
  uint64_t len = n;
  net::send_all(fd, &len, 8);
  net::send_all(fd, data, n);
}

std::string ProcessComm::recv_bytes(int src) {
  if (src < 0 || src >= n_ || src == rank_) throw Error("recv_bytes: bad source");
  int fd = in_fd(src);
  uint64_t len = 0;
  if (!net::recv_all(fd, &len, 8)) throw Error("peer closed the data connection");
  std::string out(len, '\0');
  if (len && !net::recv_all(fd, &out[0], len)) throw Error("peer closed the data connection");
  return out;
}

std::string ProcessComm::sendrecv_bytes(int dest, const void* data, size_t n, int src) {
  if (dest < 0 || dest >= n_ || dest == rank_) throw Error("sendrecv_bytes: bad destination");
  if (src < 0 || src >= n_ || src == rank_) throw Error("sendrecv_bytes: bad source");
  int sfd = out_fd(dest), rfd = in_fd(src);
  uint64_t slen = n, rlen = 0;
  net::sendrecv(sfd, &slen, 8, rfd, &rlen, 8);
  std::string out(rlen, '\0');
  net::sendrecv(sfd, data, n, rfd, rlen ? &out[0] : nullptr, rlen);
  return out;
}

}  // namespace mp4j
