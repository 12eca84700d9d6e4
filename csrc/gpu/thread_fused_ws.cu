// Warp-specialised fused ThreadCommSlave allreduce (world > 1): the production path of
// "T thread buffers per GPU x p GPUs" (TCS:1915-1963 on CUDA tensors).
//
// One persistent CTA per SM, three warp groups plus a signal warp that run CONCURRENTLY on a ring
// of staging slots:
//   group A (HBM read)   : slot[k] = reduce_t bufs[t][chunk k]          (256-bit loads, T static)
//   signal warp          : publishes "A(k) done" to every peer (the system-scope release costs ~us
//                          under load; keeping it out of group A keeps the HBM read stream running)
//   group B (NVLink)     : rank r reduces sub-slice r of slot[k] over all peers and pushes it back
//                          to every peer - multimem.ld_reduce + multimem.st (NVLS) or P2P ld/st.
//                          Each WARP owns every WB-th chunk, so WB chunks are in their NVLink round
//                          trips at once
//   group C (HBM write)  : bufs[t][chunk k] = slot[k] for all t         (256-bit stores)
// Dependencies are flags, never a CTA-wide or grid-wide barrier:
//   A(k) needs C(k - kSlots) done (local shared-memory counter)
//   B(k) needs A(k) of block b on EVERY rank   (flag array FA in the peers' signal pads)
//   C(k) needs B(k) of block b on EVERY rank   (flag array FB, indexed by ring slot: B finishes out of order)
// so the HBM read stream, the NVLink exchange and the HBM write stream of different chunks overlap
// inside every SM, and 1/T of the bytes cross NVLink. The chunk-synchronous kernel in
// thread_kernels.cu (phases in lock-step) is kept as the generic fallback (odd alignments, other
// thread counts / operators).
#include <stdlib.h>

#include <algorithm>

#include "cuda_check.h"
#include "dense.cuh"
#include "launch.h"

namespace mp4j {

namespace {

// warp split of the CTA between the three groups
template <int WA, int WB, int WC>
struct Split {
  static constexpr int kThreadsA = WA * 32, kThreadsB = WB * 32, kThreadsC = WC * 32;
  static constexpr int kWsThreads = kThreadsA + kThreadsB + kThreadsC + 32;  // + the signal warp
};
typedef Split<6, 4, 5> SplitDefault;  // 512 threads with the signal warp
constexpr int kMaxSlots = kWsMaxSlots;  // ring depth is a launch parameter (ThreadArgs::slots), 2..kMaxSlots

__device__ __forceinline__ void named_bar(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}


// diagnostics: stage timestamps of block 0 (dbg bit 128), read back by tools/ws_trace.py
#define WS_TRACE(ev, k)                                                                                  \
  do {                                                                                                   \
    if ((a.dbg & 128) && blockIdx.x == 0 && (k) < kWsTraceChunks)                                        \
      ((unsigned long long*)((char*)tab.sig[rank] + kSigWsTraceOff))[(k) * 8 + (ev)] = globaltimer_ns(); \
  } while (0)

template <typename T, int OP, bool NVLS, int NT, typename SP>
__global__ void __launch_bounds__(SP::kWsThreads, 1) thread_fused_ws_kernel(const __grid_constant__ ThreadArgs a) {
  constexpr int kThreadsA = SP::kThreadsA, kThreadsB = SP::kThreadsB, kThreadsC = SP::kThreadsC;
  __shared__ volatile unsigned c_done;  // chunks finished by group C (slot recycling)
  // arrivals of the phase-A warps per ring slot. Per SLOT, not one running total: a warp that is a chunk ahead
  // of a slow one must not complete the slow warp's chunk (a warp cannot lap the ring, so per-slot counts are exact)
  __shared__ unsigned a_cnt[kWsMaxSlots];
  const PeerTable& tab = a.tab;
  const int rank = my_rank(tab);
  const int world = tab.world;
  const int erow = tab.emulate ? rank : 0;
  const unsigned long long b0 = (unsigned long long)a.from * sizeof(T);
  const long long nbytes = (a.to - a.from) * (long long)sizeof(T);
  const long long ch = (long long)a.slot_bytes;
  const long long nchunks = (nbytes + ch - 1) / ch;
  // chunks of this block: blockIdx.x, blockIdx.x + G, ...
  const long long mine = nchunks > (long long)blockIdx.x ? (nchunks - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  uint32_t* ws_epoch = (uint32_t*)((char*)tab.sig[rank] + kSigWsEpochOff) + blockIdx.x;
  const uint32_t base = *(volatile uint32_t*)ws_epoch;
  const int kSlots = a.slots;
  const unsigned long long ring_off = a.stg_off + (unsigned long long)blockIdx.x * kSlots * a.slot_bytes;
  if (threadIdx.x == 0) {
    c_done = 0;
    for (int i = 0; i < kWsMaxSlots; ++i) a_cnt[i] = 0;
  }
  __syncthreads();

  const int tid = threadIdx.x;
  if (tid < kThreadsA) {
    // ================= group A: T-way reduce of the local thread buffers into the slot ring.
    // The warps free-run (no group barrier): each finishes its share of chunk k, bumps a shared
    // counter and moves on; publishing the chunk to the peers (a system-scope release, ~us) is the
    // signal warp's job, so the HBM read stream never waits for a fence.
    const int lane = tid & 31;
    for (long long k = 0; k < mine; ++k) {
      const long long c = blockIdx.x + k * gridDim.x;
      const long long off = c * ch;
      const int nv = (int)(min(ch, nbytes - off) / 32);
      if (k >= kSlots) {
        if (lane == 0)
          while (c_done < (unsigned)(k - kSlots + 1)) {
          }
        __syncwarp();
      }
      char* slot = tab.heap[rank] + ring_off + (unsigned long long)(k % kSlots) * a.slot_bytes;
      for (int v = tid; v < ((a.dbg & 4) ? 0 : nv); v += kThreadsA) {
        const unsigned long long o = b0 + off + (unsigned long long)v * 32;
        Vec32 x[NT];
#pragma unroll
        for (int t = 0; t < NT; ++t) x[t] = ld_stream32((const char*)a.bufs[erow][t] + o);
#pragma unroll
        for (int t = 1; t < NT; ++t) {
          x[0].lo = vec_reduce<T, OP>(x[0].lo, x[t].lo);
          x[0].hi = vec_reduce<T, OP>(x[0].hi, x[t].hi);
        }
        if (!(a.dbg & 32)) st_keep32(slot + (unsigned long long)v * 32, x[0]);
      }
      // fa_gpu: every storing thread fences its own slot stores to device scope. The slot is in THIS GPU's
      // memory and the peers read it through this GPU's L2, so that is what the flag must be ordered after;
      // the signal warp then needs no system-scope release (25-30 us per chunk at 8 GPUs under load).
      if (a.fa_gpu) __threadfence();
      __syncwarp();
      if (lane == 0) {
        __threadfence_block();  // the warp's slot stores are ordered before the count
        atomicAdd(&a_cnt[k % kSlots], 1u);
        if (tid == 0) WS_TRACE(0, k);
      }
    }
  } else if (tid >= kThreadsA + kThreadsB + kThreadsC) {
    // ================= signal warp: FA(k) to every peer once all A warps finished chunk k
    const int lane = tid & 31;
    constexpr unsigned WA = kThreadsA / 32;
    for (long long k = 0; k < mine; ++k) {
      if (lane < world) {
        while (*(volatile unsigned*)&a_cnt[k % kSlots] < WA * (unsigned)(k / kSlots + 1)) {
        }
        __threadfence_block();
        if (lane == 0) WS_TRACE(1, k);
        if (a.fa_gpu) {
          __threadfence();
          st_relaxed_sys((uint32_t*)((char*)tab.sig[lane] + kSigWsFaOff) + blockIdx.x * kMaxRanks + rank,
                         base + (uint32_t)k + 1u);
        } else {
          st_release_sys((uint32_t*)((char*)tab.sig[lane] + kSigWsFaOff) + blockIdx.x * kMaxRanks + rank,
                         base + (uint32_t)k + 1u);
        }
        if (lane == 0) WS_TRACE(2, k);
      }
    }
  } else if (tid < kThreadsA + kThreadsB) {
    // ================= group B: inter-GPU reduce-scatter + allgather of the slot over NVLink.
    // Every WARP owns the chunks k = warp, warp + WB, ... of this block, so WB chunks are in their
    // NVLink round trips at once: the per-chunk latency (flag hop + load RTT + store + flag hop, ~10 us)
    // is hidden by chunk-level parallelism instead of by large chunks, which keeps the ring small
    // (L2-resident) and the pipeline fill/drain short.
    constexpr int WB = kThreadsB / 32;
    const int bw = (tid - kThreadsA) >> 5, lane = tid & 31;
    for (long long k = bw; k < mine; k += WB) {
      const long long c = blockIdx.x + k * gridDim.x;
      const long long off = c * ch;
      const int nv16 = (int)(min(ch, nbytes - off) / 16);
      if (lane < world && (!(a.dbg & 2) || lane == rank))
        wait_flag(tab, (uint32_t*)((char*)tab.sig[rank] + kSigWsFaOff) + blockIdx.x * kMaxRanks + lane,
                  base + (uint32_t)k + 1u);
      __syncwarp();
      if (lane == 0) WS_TRACE(3, k);
      const int sl = (int)(k % kSlots);
      const unsigned long long so = ring_off + (unsigned long long)sl * a.slot_bytes;
      const int lo = (int)((long long)nv16 * rank / world), hi = (int)((long long)nv16 * (rank + 1) / world);
      // An NVLink round trip is ~2-3 us: keep kBatch independent 16-byte requests in flight per
      // thread (the asm volatile accessors are never reordered, so batching is explicit).
      constexpr int kBatch = NVLS ? 16 : 8;  // NVLS holds only accumulators: 16 requests per lane = one round trip at 8 ranks
      const bool loc = (a.dbg & 8) != 0;
      for (int v0 = lo + lane; v0 < ((a.dbg & 1) ? lo : hi); v0 += 32 * kBatch) {
        Vec16 acc[kBatch];
        if constexpr (NVLS) {
#pragma unroll
          for (int j = 0; j < kBatch; ++j) {
            const int v = v0 + j * 32;
            if (v < hi) acc[j] = Nvls<T, OP>::ld_reduce(tab.mc_heap + so + (unsigned long long)v * 16);
          }
#pragma unroll
          for (int j = 0; j < kBatch; ++j) {
            const int v = v0 + j * 32;
            if (v < hi) multimem_st(tab.mc_heap + so + (unsigned long long)v * 16, acc[j]);
          }
        } else {
          for (int q = 0; q < world; ++q) {  // ascending rank order => identical bits on every rank
            if (loc && q != rank) continue;
            Vec16 x[kBatch];
#pragma unroll
            for (int j = 0; j < kBatch; ++j) {
              const int v = v0 + j * 32;
              if (v < hi) x[j] = ld_coherent(tab.heap[q] + so + (unsigned long long)v * 16);
            }
#pragma unroll
            for (int j = 0; j < kBatch; ++j) acc[j] = (q == 0 || loc) ? x[j] : vec_reduce<T, OP>(acc[j], x[j]);
          }
          for (int q = 0; q < world; ++q) {
            if (loc && q != rank) continue;
#pragma unroll
            for (int j = 0; j < kBatch; ++j) {
              const int v = v0 + j * 32;
              if (v < hi) st_stream(tab.heap[q] + so + (unsigned long long)v * 16, acc[j]);
            }
          }
        }
      }
      __syncwarp();  // the warp's stores precede the release below
      if (lane == 0) WS_TRACE(4, k);
      if (lane < world)
        st_release_sys((uint32_t*)((char*)tab.sig[lane] + kSigWsFbOff) +
                           ((size_t)blockIdx.x * kWsMaxSlots + sl) * kMaxRanks + rank,
                       base + (uint32_t)k + 1u);
      if (lane == 0) WS_TRACE(5, k);
    }
  } else {
    // ================= group C: fan the finished slot out to the T thread buffers
    const int ct = tid - kThreadsA - kThreadsB;
    for (long long k = 0; k < mine; ++k) {
      const long long c = blockIdx.x + k * gridDim.x;
      const long long off = c * ch;
      const int nv = (int)(min(ch, nbytes - off) / 32);
      if (ct < world && (!(a.dbg & 2) || ct == rank))
        wait_flag(tab, (uint32_t*)((char*)tab.sig[rank] + kSigWsFbOff) +
                           ((size_t)blockIdx.x * kWsMaxSlots + (int)(k % kSlots)) * kMaxRanks + ct,
                  base + (uint32_t)k + 1u);
      named_bar(3, kThreadsC);
      if (ct == 0) WS_TRACE(6, k);
      const char* slot = tab.heap[rank] + ring_off + (unsigned long long)(k % kSlots) * a.slot_bytes;
      for (int v = ct; v < ((a.dbg & 4) ? 0 : nv); v += kThreadsC) {
        Vec32 x{};
        if (!(a.dbg & 16)) x = ld_coherent32(slot + (unsigned long long)v * 32);
        const unsigned long long o = b0 + off + (unsigned long long)v * 32;
        if (a.fanout) {
#pragma unroll
          for (int t = 0; t < NT; ++t) st_stream32((char*)a.bufs[erow][t] + o, x);
        } else {
          st_stream32((char*)a.bufs[erow][a.dst_index] + o, x);
        }
      }
      named_bar(3, kThreadsC);  // slot fully consumed
      if (ct == 0) c_done = (unsigned)(k + 1);
      if (ct == 0) WS_TRACE(7, k);
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) *(volatile uint32_t*)ws_epoch = base + (uint32_t)mine;
}

template <typename T, int OP, bool NVLS, int NT, typename SP>
void launch_ws_sp(const ThreadArgs& args, int grid, cudaStream_t stream) {
  auto kfn = thread_fused_ws_kernel<T, OP, NVLS, NT, SP>;
  if (args.tab.emulate) {
    void* params[] = {(void*)&args};
    MP4J_CUDA(cudaLaunchCooperativeKernel((const void*)kfn, dim3(grid, args.tab.world), dim3(SP::kWsThreads), params,
                                          0, stream));
  } else {
    kfn<<<grid, SP::kWsThreads, 0, stream>>>(args);
    MP4J_CUDA(cudaGetLastError());
  }
}

template <typename T, int OP, bool NVLS, int NT>
void launch_ws_nt(const ThreadArgs& args, int grid, cudaStream_t stream) {
  if constexpr (sizeof(T) == 4 && OP == MP4J_SUM && NT == 8) {  // experiment: alternative warp splits
    static const int split = getenv("MP4J_WS_SPLIT") ? atoi(getenv("MP4J_WS_SPLIT")) : 0;
    switch (split) {
      case 1: return launch_ws_sp<T, OP, NVLS, NT, Split<8, 2, 5>>(args, grid, stream);
      case 2: return launch_ws_sp<T, OP, NVLS, NT, Split<7, 3, 5>>(args, grid, stream);
      case 3: return launch_ws_sp<T, OP, NVLS, NT, Split<8, 3, 4>>(args, grid, stream);
      case 4: return launch_ws_sp<T, OP, NVLS, NT, Split<6, 5, 4>>(args, grid, stream);
    }
  }
  launch_ws_sp<T, OP, NVLS, NT, SplitDefault>(args, grid, stream);
}

template <typename T, int OP, bool NVLS>
void launch_ws_t(const ThreadArgs& args, int grid, cudaStream_t stream) {
  switch (args.T) {
    case 2: return launch_ws_nt<T, OP, NVLS, 2>(args, grid, stream);
    case 4: return launch_ws_nt<T, OP, NVLS, 4>(args, grid, stream);
    case 8: return launch_ws_nt<T, OP, NVLS, 8>(args, grid, stream);
  }
  throw Error("warp-specialised fused kernel is built for T in {2, 4, 8}");
}

template <typename T, int OP>
void launch_ws_op(const ThreadArgs& args, bool nvls, int grid, cudaStream_t stream) {
  if constexpr (Nvls<T, OP>::supported) {
    if (nvls) return launch_ws_t<T, OP, true>(args, grid, stream);
  }
  launch_ws_t<T, OP, false>(args, grid, stream);
}

template <typename T>
void launch_ws_type(int op, const ThreadArgs& args, bool nvls, int grid, cudaStream_t stream) {
  switch (op) {
    case MP4J_SUM: return launch_ws_op<T, MP4J_SUM>(args, nvls, grid, stream);
    case MP4J_MAX: return launch_ws_op<T, MP4J_MAX>(args, nvls, grid, stream);
    case MP4J_MIN: return launch_ws_op<T, MP4J_MIN>(args, nvls, grid, stream);
  }
  throw Error("warp-specialised fused kernel is built for SUM/MAX/MIN");
}

}  // namespace

int thread_ws_slots() { return kMaxSlots; }
int thread_ws_max_blocks() { return kWsMaxBlocks; }

void launch_thread_fused_ws(int dtype, int op, bool nvls, const ThreadArgs& args, int grid, cudaStream_t stream) {
  switch (dtype) {
    case MP4J_F64: return launch_ws_type<double>(op, args, nvls, grid, stream);
    case MP4J_F32: return launch_ws_type<float>(op, args, nvls, grid, stream);
    case MP4J_I64: return launch_ws_type<int64_t>(op, args, nvls, grid, stream);
    case MP4J_I32: return launch_ws_type<int32_t>(op, args, nvls, grid, stream);
  }
  throw Error("warp-specialised fused kernel is built for 4/8-byte types");
}

}  // namespace mp4j
