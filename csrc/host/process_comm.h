// ProcessComm: the process-level communicator core (parity: J/comm/ProcessCommSlave.java).
//  * control: MasterClient (rendezvous, barrier, logs, close, heartbeat)    PCS:143-210
//  * host data plane: persistent full-mesh TCP between slaves               PCS:391-425 (per-message
//    sockets in the reference; persistent here)
//  * host collectives on primitive arrays: ring reduce-scatter / ring allgather /
//    binomial-tree gather / binomial-tree scatter, and the three compositions
//    broadcast = scatter+allgather, reduce = reduce-scatter+gather,
//    allreduce = reduce-scatter+allgather                                   PCS:1329-1373, 694-737,
//                                                                          440-520, 1103-1159, 769-770,
//                                                                          1415-1416, 1758-1759
//  * byte-blob point-to-point used by the bindings for String/Object/map payloads.
// The GPU data plane (csrc/gpu) attaches to this object for handle brokerage.
#pragma once
#include <condition_variable>
#include <functional>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../control/control.h"
#include "../core/ops.h"
#include "../core/partition.h"
#include "shm_plane.h"

namespace mp4j {

class ProcessComm {
 public:
  // rank_hint < 0: reference rule (rank = index of "host###port" in the sorted address list).
  ProcessComm(const std::string& login, const std::string& master_host, int master_port,
              int rank_hint = -1, const std::string& my_host = "");
  ~ProcessComm();

  int rank() const { return rank_; }
  int size() const { return n_; }
  const std::vector<std::string>& addresses() const { return addrs_; }
  MasterClient& master() { return *master_; }
  bool closed() const { return closed_; }
  // "shm" when the ranks share a machine and exchange host arrays through shared-memory windows
  // (shm_plane.h), "tcp" for the ring / tree algorithms over the socket mesh.
  const char* host_transport() const { return shm_ ? "shm" : (node_shm_ ? "tcp+shm" : "tcp"); }

  void barrier();
  void log(LogLevel lvl, const std::string& msg);  // forwards verbatim; prefixes are the binding's job
  void close(int code);                            // PCS:234

  // ---- host collectives on primitive arrays (in place, element ranges) ----
  void reduce_scatter(void* arr, int dtype, int op, int64_t from, const int64_t* counts);
  void allgather(void* arr, int dtype, const int64_t* froms, const int64_t* tos);
  void gather(void* arr, int dtype, const int64_t* froms, const int64_t* tos, int root);
  void scatter(void* arr, int dtype, const int64_t* froms, const int64_t* tos, int root);
  void broadcast(void* arr, int dtype, int64_t from, int64_t to, int root);
  void reduce(void* arr, int dtype, int op, int64_t from, int64_t to, int root);
  void allreduce(void* arr, int dtype, int op, int64_t from, int64_t to);
  // The process step of a thread-level collective, executed by ALL T threads of the process: each thread opens a
  // TeamScope (tid 0 = the delegate) and calls the same array collective with the same arguments. On the
  // shared-memory plane the team shares the copies and the reduction (ShmPlane::grs_team); on the TCP paths thread
  // 0 runs the collective and the followers return at once (the caller's next thread barrier is where they wait).
  struct Team {
    int tid;
    int T;
    const std::function<void()>* sync;  // the team's thread barrier
  };
  class TeamScope {
   public:
    explicit TeamScope(const Team* t);
    ~TeamScope();
    TeamScope(const TeamScope&) = delete;
    TeamScope& operator=(const TeamScope&) = delete;
  };
  // small-message allreduce through the master mailbox (PCS:1776-1799, SRV:373-497)
  void allreduce_rpc(void* arr, int dtype, int op, int64_t n);

  // ---- maps of (String key, primitive value): allreduceMap (mode 0) / reduceMap (mode 1), map_host.cc.
  // keys: n NUL-separated UTF-8 strings; vals: n elements of dtype. Outputs use the same layout.
  void map_reduce(int mode, int root, const char* blob, size_t blob_len, const void* vals, size_t n, int dtype,
                  int op, std::string* out_blob, std::string* out_vals);
  // reduceScatterMap: p maps back to back in blob / vals (counts[q] entries each); result = this rank's merged map
  void map_reduce_scatter(const char* blob, size_t blob_len, const void* vals, const int64_t* counts, int dtype, int op,
                          std::string* out_blob, std::string* out_vals);

  // ---- byte-blob transport (length-prefixed) ----
  // Collectives on opaque payloads (blob_coll.cc): the ring / tree runs here, the binding supplies the codec.
  typedef int (*BlobGet)(void* user, int part, const char** data, size_t* n);
  typedef int (*BlobPut)(void* user, int part, const char* data, size_t n);
  void blob_reduce_scatter(BlobGet get, BlobPut put, void* user);   // p parts per rank; result = part[rank]
  void blob_allgather(BlobGet get, BlobPut put, void* user);        // part[rank] is mine; put(q) for the others
  void blob_gather(int root, BlobGet get, BlobPut put, void* user); // put(q) on the root only
  void blob_scatter(int root, BlobGet get, BlobPut put, void* user);  // root: get(q) for all q; others: put(rank)
  void send_bytes(int dest, const void* data, size_t n);
  std::string recv_bytes(int src);
  std::string sendrecv_bytes(int dest, const void* data, size_t n, int src);

 private:
  bool tcp_oneshot(size_t bytes, int members = -1) const;
  void tcp_reduce_scatter(char* b, int dtype, int op, const int64_t* froms, const int64_t* tos,
                          const std::vector<int>& grp, int me);
  void tcp_allgather(char* b, size_t es, const int64_t* froms, const int64_t* tos, const std::vector<int>& grp, int me);
  std::vector<int> everyone_;  // 0 .. n-1: the group of the flat collectives
  // hierarchical allreduce (several machines, several ranks on each): the plane of my machine, its members, and the
  // ranks with my local index on every machine (in machine order; node_index_ = my place among them)
  void setup_node_plane();
  std::unique_ptr<ShmPlane> node_shm_;
  std::vector<int> node_group_, rail_group_;
  std::vector<std::vector<int>> nodes_;  // members of every machine, machines in the order of their lowest rank
  int node_index_ = 0;
  void exchange_small(const void* mine, size_t bytes, char* all);
  void shm_run(void* arr, int dtype, int op, const int64_t* lo, const int64_t* hi, const uint32_t* src_mask,
               const uint32_t* dst_mask);
  uint32_t all_mask() const { return n_ >= 32 ? 0xffffffffu : ((1u << n_) - 1u); }
  int out_fd(int dest);
  int in_fd(int src);
  void accept_loop(int listen_fd);

  std::unique_ptr<MasterClient> master_;
  int rank_ = -1, n_ = 0;
  std::vector<std::string> addrs_;
  std::vector<std::string> hosts_;
  std::vector<int> ports_;
  int listen_fd_ = -1, listen_port_ = 0;
  std::thread accept_thr_;
  std::mutex mesh_mu_;
  std::condition_variable mesh_cv_;
  std::vector<int> out_fds_, in_fds_;
  std::mutex op_mu_;  // one collective at a time per communicator
  bool closed_ = false;
  std::vector<char> scratch_;
  std::unique_ptr<ShmPlane> shm_;  // intra-node data plane; null = TCP
};

}  // namespace mp4j
