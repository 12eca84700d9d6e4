// Argument blocks of the dense collective kernels (shared by the kernels and the host launcher).
#pragma once
#include "peer.cuh"

namespace mp4j {

// One work item per rank of the "gather-reduce-scatter" kernel: reduce elements [lo, hi) over the
// ranks in src_mask (ascending rank order => identical bits everywhere) and store the result into
// the heaps of the ranks in dst_mask and/or a local pointer. All seven array collectives are
// instances of this shape (see dense.cu).
struct GrsItem {
  long long lo, hi;        // element range (bytes when the op is COPY)
  unsigned src_mask;       // which ranks' heaps are read
  unsigned dst_mask;       // which ranks' heaps are written
  void* dst_local;         // optional: element 0 of a local (non-symmetric) output array
  const void* src_local;   // optional: local input copied into this rank's heap first (one-shot)
  long long adj;           // bytes added to src_off / dst_off for THIS item (staged windows pack one sub-range per rank)
};

struct GrsArgs {
  PeerTable tab;
  unsigned long long src_off;  // byte offset in every heap of element 0 of the source array
  unsigned long long dst_off;  // byte offset in every heap of element 0 of the destination array
  int entry_barrier;
  int exit_barrier;
  int vec_ok;                  // local src/dst pointers are 16-byte congruent with the heap addresses
  int dead_rank_p1;            // test hook (option dead_rank): rank + 1 that returns at once, 0 = none
  unsigned long long layout_key;  // where this rank believes the array lives; all ranks must agree (0 = not checked)
  GrsItem item[kMaxRanks];     // indexed by rank in emulation, [0] otherwise
};

// Fused thread-level + inter-GPU allreduce (ThreadCommSlave.allreduceArray on CUDA tensors).
struct ThreadArgs {
  PeerTable tab;
  int T;                        // thread buffers per rank
  int fanout;                   // 1: result to all T buffers; 0: only to bufs[dst_index]
  int dst_index;
  int slots;                    // ring depth of the warp-specialised kernel
  int dbg;                      // diagnostics only (MP4J_WS_DBG): bit mask that disables parts of the pipeline
  int tail_split;               // pull kernel: a block's last chunk runs as this many shorter pipeline steps (>= 1)
  long long from, to;           // element range
  int push;                     // pull kernel: 1 = phase B pushes its result to every rank (hybrid), 0 = phase C pulls
  int fa_gpu;                   // register formulation: 1 = phase A publishes with a device-scope fence + relaxed flag
  unsigned long long stg_off;   // byte offset in every heap of the per-block staging slots
  unsigned long long slot_bytes;  // bytes per block slot (multiple of 16 * world)
  void* bufs[kMaxRanks][MP4J_MAX_THREADS];  // [rank][thread] in emulation, [0][thread] otherwise
};

}  // namespace mp4j
