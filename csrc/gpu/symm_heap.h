// SymmHeap: one equally-sized, equally-laid-out device allocation per rank, mapped into every
// rank's address space, plus (when the fabric supports it) one multicast mapping for NVLS.
//
// Set-up paths, chosen collectively at init:
//   1. CUDA VMM: cuMemCreate + POSIX-fd export; fds travel between the ranks of a node over
//      Unix-domain sockets (SCM_RIGHTS); peers cuMemMap them. Required for multicast.
//   2. fallback: cudaMalloc + cudaIpcGetMemHandle, handles brokered by the master mailbox.
//   3. emulation: all "ranks" are allocations of ONE process on ONE GPU (tests on a single GPU).
// The master's gather mailbox (SRV:373-514 equivalent) is the broker in all real modes - this is
// the "new duty" of the control plane named in SURVEY §2.1 C2.
#pragma once
#include <stddef.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "peer.cuh"

namespace mp4j {

class ProcessComm;

class SymmHeap {
 public:
  ~SymmHeap();
  // Collective over the communicator. Current device must already be `device`.
  void init_real(ProcessComm* pc, int device, size_t bytes, bool want_multicast);
  void init_emulated(int world, int device, size_t bytes);

  int world() const { return world_; }
  int rank() const { return rank_; }
  size_t bytes() const { return bytes_; }
  char* base(int r) const { return base_[r]; }
  char* local() const { return base_[emulated_ ? 0 : rank_]; }
  char* mc_base() const { return mc_base_; }
  bool emulated() const { return emulated_; }
  const char* mode() const { return emulated_ ? "emulated" : (vmm_ ? (mc_base_ ? "vmm+multicast" : "vmm") : "cuda-ipc"); }
  const std::string& notes() const { return notes_; }

 private:
  void destroy();
  bool try_vmm_local(size_t bytes);
  void exchange_fds(ProcessComm* pc, const std::vector<int>& send_fds, std::vector<std::vector<int>>* recv);

  int world_ = 0, rank_ = 0, device_ = 0;
  size_t bytes_ = 0;
  bool emulated_ = false, vmm_ = false;
  char* base_[kMaxRanks] = {};
  char* mc_base_ = nullptr;
  std::string notes_;
  // VMM bookkeeping (opaque 64-bit driver handles)
  unsigned long long local_handle_ = 0;
  unsigned long long peer_handles_[kMaxRanks] = {};
  unsigned long long mc_handle_ = 0;
  bool mc_bound_ = false;
  bool ipc_opened_[kMaxRanks] = {};
};

}  // namespace mp4j
