// ShmPlane: intra-node host data plane for primitive arrays.
//
// The reference moves every host array through TCP sockets even between slaves of one machine
// (PCS:391-425: connect + Kryo write per message, p-1 serial ring steps, a master barrier RPC per
// phase). On an 8-GPU node all slaves share one host, so the host path here uses the same idea as
// the GPU plane: every rank owns a window in POSIX shared memory that all peers map, and every
// collective is ONE "gather-reduce-scatter" pass over those windows -
//     copy-in my contribution | barrier | reduce my slice over all windows (ascending rank order)
//     | barrier | copy-out the slices I am a destination of
// with futex barriers in shared memory instead of master round trips. The seven array collectives
// are mask choices of that pass, exactly as in csrc/gpu/dense.cu. Arrays larger than a window are
// streamed through the two halves of the window (double buffering saves the closing barrier).
//
// Used automatically when every rank of the communicator can map every window (same boot id);
// MP4J_SHM=0 forces the TCP ring/tree algorithms, which remain the multi-host path.
#pragma once
#include <stddef.h>
#include <stdint.h>

#include <functional>
#include <memory>
#include <string>
#include <vector>

namespace mp4j {

class MasterClient;

class ShmPlane {
 public:
  // Collective over all ranks. Returns nullptr (on every rank alike) if shared memory cannot be
  // used: different machines, MP4J_SHM=0 on any rank, or a mapping failure anywhere.
  // group = the job ranks that form this plane (ascending), null = the whole job. With a group the call is still
  // collective over the WHOLE job (every rank builds the plane of its own group at the same time), and the verdict
  // is global: one group that cannot map its members turns the planes of all groups off.
  static std::unique_ptr<ShmPlane> create(MasterClient& master, int rank, int n, const std::vector<int>* group = nullptr);
  static std::string machine_key();
  int members() const { return n_; }
  int index() const { return rank_; }
  ~ShmPlane();

  // One gather-reduce-scatter pass. Element ranges [lo[q], hi[q]) are owned by rank q; the owner
  // reduces the contributions of the ranks in src_mask[q] (ascending) and the result is delivered
  // into the arrays of the ranks in dst_mask[q]. op == MP4J_COPY: no reduction (src_mask has one bit).
  void grs(void* arr, int dtype, int op, const int64_t* lo, const int64_t* hi, const uint32_t* src_mask,
           const uint32_t* dst_mask);
  // grs executed by a team of T threads of this process (see shm_plane.cc); team_sync = their thread barrier
  void grs_team(void* arr, int dtype, int op, const int64_t* lo, const int64_t* hi, const uint32_t* src_mask,
                const uint32_t* dst_mask, int tid, int T, const std::function<void()>* team_sync);
  void barrier();
  size_t window_bytes() const { return half_bytes_ * 2; }

 private:
  ShmPlane() {}
  struct Header;
  int rank_ = 0, n_ = 0;
  size_t seg_bytes_ = 0, half_bytes_ = 0;
  std::vector<char*> seg_;  // mapping of every rank's segment
  int half_ = 0;            // which half of the windows the next chunk uses (same on every rank)
  double timeout_s_ = 600.0;
  double spin_us_ = 0.0;    // how long a barrier spins before it sleeps (0 when cores are scarce)
  bool team_failed_ = false;  // outcome of thread 0's process barrier, read by its team after the thread barrier
  std::string team_msg_;
};

}  // namespace mp4j
