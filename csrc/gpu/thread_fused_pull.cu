// Fused ThreadCommSlave allreduce (world > 1), pull formulation: no remote data store and therefore no
// system-scope release fence anywhere in the kernel (TCS:1915-1963 on CUDA tensors).
//
// Measured on 2 x B200 with per-stage timestamps (tools/ws_sweep.py, dbg bit 128; profiles/README.md
// §1): in the push formulation (thread_fused_ws.cu) one chunk needs 60-75 us from "local reduce done"
// to "fanned out" - 5-15 us for EACH st.release.sys (a system-scope fence that follows remote stores
// waits for the fabric to acknowledge them and queues behind the fences of the other 147 CTAs) and
// ~30 us for phase B, where one warp walked its chunk in 8 dependent NVLink round trips. With a ring
// of 4 slots that is the whole latency budget, so every hiccup stalls the HBM streams; it is also the
// pipeline's fill/drain time.
//
// Here every cross-GPU transfer is a LOAD by the consumer:
//   group A  : slot[k] = reduce_t bufs[t][chunk k]                       (local stores)
//   signal   : fence.gpu, then one relaxed flag store per peer: FA(k)    (data is in this GPU's L2,
//              where the peers' NVLink reads are served; nothing remote to wait for)
//   group B  : the WHOLE group works on one chunk: rank r loads sub-slice r of every rank's slot
//              (multimem.ld_reduce, or all peers' loads in flight together), reduces in rank order and
//              writes the result back into ITS OWN slot; fence.gpu; relaxed flag stores: FB(k)
//   group C  : pulls sub-slice q of the chunk from rank q's slot for every q (8 independent 32-byte
//              loads per thread in flight) and fans it out to the T thread buffers; then tells every
//              owner "pulled": FC(k)
//   group A  : reuses slot[k % slots] once its own C and every peer's FC of chunk k - slots arrived.
// The same bytes cross NVLink as before (1/T of the data, twice), but a chunk is now two flag hops and
// two load round trips deep.
#include <stdlib.h>

#include <algorithm>

#include "cuda_check.h"
#include "dense.cuh"
#include "launch.h"

namespace mp4j {

namespace {

template <int WA, int WB, int WC>
struct SplitP {
  static constexpr int kThreadsA = WA * 32, kThreadsB = WB * 32, kThreadsC = WC * 32;
  static constexpr int kThreads = kThreadsA + kThreadsB + kThreadsC + 64;  // + the two signal warps (FA, FB)
};

__device__ __forceinline__ void named_bar(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
__device__ __forceinline__ void st_slot16(void* p, const Vec16& v) {  // result sub-slice: read by every peer soon
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z),
               "r"(v.w)
               : "memory");
}

// diagnostics: stage timestamps of block 0 (dbg bit 128), read back by tools/ws_sweep.py
#define WS_TRACE(ev, k)                                                                                  \
  do {                                                                                                   \
    if ((a.dbg & 128) && blockIdx.x == 0 && (k) < kWsTraceChunks)                                        \
      ((unsigned long long*)((char*)tab.sig[rank] + kSigWsTraceOff))[(k) * 8 + (ev)] = globaltimer_ns(); \
  } while (0)

// The pipeline steps of one block: its chunks blockIdx.x, blockIdx.x + G, ... in order, except that the LAST
// chunk is cut into `split` shorter steps - what is still in flight when the read stream ends (the drain
// of the pipeline) is then a quarter-size step instead of a whole chunk.
struct Steps {
  long long nbytes, ch, sub;   // range bytes, chunk bytes, sub-step bytes (multiple of 32 * world is not needed)
  long long first, stride;     // blockIdx.x, gridDim.x
  long long mine;              // chunks of this block
  long long nsteps;
  __device__ __forceinline__ void init(long long nbytes_, long long ch_, int split) {
    nbytes = nbytes_;
    ch = ch_;
    first = blockIdx.x;
    stride = gridDim.x;
    const long long nchunks = (nbytes + ch - 1) / ch;
    mine = nchunks > first ? (nchunks - first + stride - 1) / stride : 0;
    sub = split > 1 ? max((ch / split + 31) / 32 * 32, 32ll) : ch;
    nsteps = mine;
    if (mine > 0 && sub < ch) {
      const long long off = (first + (mine - 1) * stride) * ch;
      const long long len = min(ch, nbytes - off);
      nsteps = mine - 1 + (len + sub - 1) / sub;
    }
  }
  __device__ __forceinline__ void get(long long s, long long& off, int& nv32) const {
    if (s < mine - 1 || sub >= ch) {
      off = (first + s * stride) * ch;
      nv32 = (int)(min(ch, nbytes - off) / 32);
    } else {
      const long long o0 = (first + (mine - 1) * stride) * ch;
      const long long len = min(ch, nbytes - o0);
      const long long j = s - (mine - 1);
      off = o0 + j * sub;
      nv32 = (int)(min(sub, len - j * sub) / 32);
    }
  }
};

// phase B without the switch: W ranks, all loads of a batch in flight together, rank-ordered reduce
template <typename T, int OP, int W, int nthr, bool PUSH>
__device__ __forceinline__ void b_p2p(const PeerTable& tab, unsigned long long so, int rank, int lo, int hi, int btid) {
  constexpr int kB = W == 2 ? 6 : 8 / W;  // vectors per thread per round: 8-12 loads in flight (16 spill)
  for (int v0 = lo + btid; v0 < hi; v0 += nthr * kB) {
    Vec16 x[kB][W];
#pragma unroll
    for (int j = 0; j < kB; ++j) {
      const int v = v0 + j * nthr;
#pragma unroll
      for (int q = 0; q < W; ++q)
        if (v < hi) x[j][q] = ld_coherent(tab.heap[q] + so + (unsigned long long)v * 16);
    }
#pragma unroll
    for (int j = 0; j < kB; ++j) {
      const int v = v0 + j * nthr;
      if (v < hi) {
        Vec16 acc = x[j][0];
#pragma unroll
        for (int q = 1; q < W; ++q) acc = vec_reduce<T, OP>(acc, x[j][q]);  // ascending rank order
        if constexpr (PUSH) {
#pragma unroll
          for (int q = 0; q < W; ++q) st_slot16(tab.heap[q] + so + (unsigned long long)v * 16, acc);
        } else {
          st_slot16(tab.heap[rank] + so + (unsigned long long)v * 16, acc);
        }
      }
    }
  }
}

// PUSH = false: the pull formulation described at the top. PUSH = true ("hybrid", the choice beyond 2 ranks):
// phase B pushes its reduced sub-slice into every rank's slot (multimem.st: one store, the switch replicates -
// pulling the same bytes would double every GPU's NVLink egress at 8 ranks and made phase C, 7/8 remote
// loads, the bottleneck: 0.85 vs 0.82 ms), so phase C reads only local memory and no FC flags exist. The
// remote stores need ONE system-scope release per chunk; it is issued by a warp of its own (signal warp B), so
// its 15-20 us under load are latency of that chunk only, never a stall of the group that moves data.
template <typename T, int OP, bool NVLS, int NT, typename SP, bool PUSH>
__global__ void __launch_bounds__(SP::kThreads, 1) thread_fused_pull_kernel(const __grid_constant__ ThreadArgs a) {
  constexpr int kThreadsA = SP::kThreadsA, kThreadsB = SP::kThreadsB, kThreadsC = SP::kThreadsC;
  __shared__ volatile unsigned c_done;  // chunks whose slot group C has fully read
  // arrivals of the phase-A warps per ring slot. Per SLOT, not one running total: a warp that is a chunk ahead
  // of a slow one must not complete the slow warp's chunk (a warp cannot lap the ring, so per-slot counts are exact)
  __shared__ unsigned a_cnt[kWsMaxSlots];
  __shared__ unsigned b_cnt[kWsMaxSlots];  // same for the phase-B warps (PUSH: consumed by signal warp B)
  const PeerTable& tab = a.tab;
  const int rank = my_rank(tab);
  const int world = tab.world;
  const int erow = tab.emulate ? rank : 0;
  const unsigned long long b0 = (unsigned long long)a.from * sizeof(T);
  const long long nbytes = (a.to - a.from) * (long long)sizeof(T);
  Steps steps;
  steps.init(nbytes, (long long)a.slot_bytes, a.tail_split);
  const long long mine = steps.nsteps;  // pipeline steps of this block (k below is a step index)
  uint32_t* ws_epoch = (uint32_t*)((char*)tab.sig[rank] + kSigWsEpochOff) + blockIdx.x;
  const uint32_t base = *(volatile uint32_t*)ws_epoch;
  const int kSlots = a.slots;
  const unsigned long long ring_off = a.stg_off + (unsigned long long)blockIdx.x * kSlots * a.slot_bytes;
  // flag addresses: F?(rank that OWNS the pad, source rank)
  auto fa = [&](int owner, int src) {
    return (uint32_t*)((char*)tab.sig[owner] + kSigWsFaOff) + blockIdx.x * kMaxRanks + src;
  };
  auto fb = [&](int owner, int sl, int src) {
    return (uint32_t*)((char*)tab.sig[owner] + kSigWsFbOff) + ((size_t)blockIdx.x * kWsMaxSlots + sl) * kMaxRanks + src;
  };
  auto fc = [&](int owner, int sl, int src) {
    return (uint32_t*)((char*)tab.sig[owner] + kSigWsFcOff) + ((size_t)blockIdx.x * kWsMaxSlots + sl) * kMaxRanks + src;
  };
  if (threadIdx.x == 0) {
    c_done = 0;
    for (int i = 0; i < kWsMaxSlots; ++i) a_cnt[i] = b_cnt[i] = 0;
  }
  __syncthreads();

  const int tid = threadIdx.x;
  if (tid < kThreadsA) {
    // ================= group A: T-way reduce of the local thread buffers into the slot ring
    const int lane = tid & 31;
    for (long long k = 0; k < mine; ++k) {
      long long off;
      int nv;
      steps.get(k, off, nv);
      if (k >= kSlots) {
        // the slot is free once this rank's C and every peer's C have read chunk k - slots out of it
        if (lane == 0)
          while (c_done < (unsigned)(k - kSlots + 1)) {
          }
        if constexpr (!PUSH) {
          if (lane < world && !(a.dbg & 2))
            wait_flag(tab, fc(rank, (int)(k % kSlots), lane), base + (uint32_t)(k - kSlots) + 1u);
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
        st_keep32(slot + (unsigned long long)v * 32, x[0]);
      }
      // every storing thread fences its OWN stores to device scope (they must be in L2 before the flag
      // leaves; a fence issued later by the signal warp alone does not wait for another warp's stores)
      __threadfence();
      __syncwarp();
      if (lane == 0) {
        atomicAdd(&a_cnt[k % kSlots], 1u);
        if (tid == 0) WS_TRACE(0, k);
      }
    }
  } else if (tid >= kThreadsA + kThreadsB + kThreadsC + 32) {
    // ================= signal warp B (PUSH only): FB(k) once every B warp has issued its stores of chunk k.
    // The release is at system scope because the stores it orders went to other GPUs.
    if constexpr (PUSH) {
      const int lane = tid & 31;
      constexpr unsigned WB = kThreadsB / 32;
      for (long long k = 0; k < mine; ++k) {
        const int sl = (int)(k % kSlots);
        if (lane == 0) {
          while (*(volatile unsigned*)&b_cnt[sl] < WB * (unsigned)(k / kSlots + 1)) {
          }
          WS_TRACE(4, k);
        }
        __syncwarp();
        if (lane < world) st_release_sys(fb(lane, sl, rank), base + (uint32_t)k + 1u);
        if (lane == 0) WS_TRACE(5, k);
      }
    }
  } else if (tid >= kThreadsA + kThreadsB + kThreadsC) {
    // ================= signal warp A: FA(k) to every rank once all A warps finished chunk k.
    // The slot lives in THIS GPU's memory and the peers read it through this GPU's L2, so a
    // device-scope fence is what orders the data before the flag; the flag itself is a plain store.
    const int lane = tid & 31;
    constexpr unsigned WA = kThreadsA / 32;
    for (long long k = 0; k < mine; ++k) {
      if (lane == 0) {
        while (*(volatile unsigned*)&a_cnt[k % kSlots] < WA * (unsigned)(k / kSlots + 1)) {
        }
        WS_TRACE(1, k);
      }
      __syncwarp();
      __threadfence();
      if (lane < world) st_relaxed_sys(fa(lane, rank), base + (uint32_t)k + 1u);
      if (lane == 0) WS_TRACE(2, k);
    }
  } else if (tid < kThreadsA + kThreadsB) {
    // ================= group B: the whole group reduces sub-slice `rank` of chunk k over all ranks
    // (one or two NVLink round trips) and leaves the result in its own slot
    const int btid = tid - kThreadsA, bw = btid >> 5, lane = tid & 31;
    for (long long k = 0; k < mine; ++k) {
      long long off;
      int nv32;
      steps.get(k, off, nv32);
      if (lane < world && (!(a.dbg & 2) || lane == rank)) wait_flag(tab, fa(rank, lane), base + (uint32_t)k + 1u);
      __syncwarp();
      if (btid == 0) WS_TRACE(3, k);
      const int sl = (int)(k % kSlots);
      const unsigned long long so = ring_off + (unsigned long long)sl * a.slot_bytes;
      // sub-slice boundaries are 32-byte aligned so that group C's 32-byte loads have one owner each
      const int lo = 2 * (int)((long long)nv32 * rank / world), hi = 2 * (int)((long long)nv32 * (rank + 1) / world);
      if (!(a.dbg & 1)) {
        if constexpr (NVLS) {
          constexpr int kB = 8;
          for (int v0 = lo + btid; v0 < hi; v0 += kThreadsB * kB) {
            Vec16 acc[kB];
#pragma unroll
            for (int j = 0; j < kB; ++j) {
              const int v = v0 + j * kThreadsB;
              if (v < hi) acc[j] = Nvls<T, OP>::ld_reduce(tab.mc_heap + so + (unsigned long long)v * 16);
            }
#pragma unroll
            for (int j = 0; j < kB; ++j) {
              const int v = v0 + j * kThreadsB;
              if (v < hi) {
                if constexpr (PUSH) multimem_st(tab.mc_heap + so + (unsigned long long)v * 16, acc[j]);
                else st_slot16(tab.heap[rank] + so + (unsigned long long)v * 16, acc[j]);
              }
            }
          }
        } else {
          switch (world) {
            case 2: b_p2p<T, OP, 2, kThreadsB, PUSH>(tab, so, rank, lo, hi, btid); break;
            case 4: b_p2p<T, OP, 4, kThreadsB, PUSH>(tab, so, rank, lo, hi, btid); break;
            case 8: b_p2p<T, OP, 8, kThreadsB, PUSH>(tab, so, rank, lo, hi, btid); break;
            default:
              for (int v = lo + btid; v < hi; v += kThreadsB) {
                Vec16 acc = ld_coherent(tab.heap[0] + so + (unsigned long long)v * 16);
                for (int q = 1; q < world; ++q)
                  acc = vec_reduce<T, OP>(acc, ld_coherent(tab.heap[q] + so + (unsigned long long)v * 16));
                if constexpr (PUSH) {
                  for (int q = 0; q < world; ++q) st_slot16(tab.heap[q] + so + (unsigned long long)v * 16, acc);
                } else {
                  st_slot16(tab.heap[rank] + so + (unsigned long long)v * 16, acc);
                }
              }
          }
        }
      }
      if constexpr (PUSH) {
        // hand the chunk to signal warp B and go on with the next one
        __syncwarp();
        if (lane == 0) {
          __threadfence_block();  // the warp's stores are ordered before the count
          atomicAdd(&b_cnt[sl], 1u);
        }
      } else {
        __threadfence();          // each thread's result stores are in L2 before the barrier
        named_bar(2, kThreadsB);  // ... and the barrier orders them before the flag of warp 0
        if (bw == 0) {
          if (lane == 0) WS_TRACE(4, k);
          __threadfence();
          if (lane < world) st_relaxed_sys(fb(lane, sl, rank), base + (uint32_t)k + 1u);
          if (lane == 0) WS_TRACE(5, k);
        }
      }
    }
  } else {
    // ================= group C: pull every owner's sub-slice of the chunk and fan it out
    const int ct = tid - kThreadsA - kThreadsB;
    constexpr int kCB = 4;  // independent 32-byte loads per thread (an NVLink round trip is ~3 us)
    for (long long k = 0; k < mine; ++k) {
      long long off;
      int nv;
      steps.get(k, off, nv);
      const int sl = (int)(k % kSlots);
      if (ct < world && (!(a.dbg & 2) || ct == rank)) wait_flag(tab, fb(rank, sl, ct), base + (uint32_t)k + 1u);
      named_bar(3, kThreadsC);
      if (ct == 0) WS_TRACE(6, k);
      const unsigned long long so = ring_off + (unsigned long long)sl * a.slot_bytes;
      for (int v0 = ct; v0 < ((a.dbg & 4) ? 0 : nv); v0 += kThreadsC * kCB) {
        Vec32 x[kCB];
#pragma unroll
        for (int j = 0; j < kCB; ++j) {
          const int v = v0 + j * kThreadsC;
          if (v < nv) {
            int q = (int)((long long)v * world / nv);
            while (v >= (int)((long long)nv * (q + 1) / world)) ++q;  // owner of 32-byte vector v
            if (PUSH || (a.dbg & 8)) q = rank;
            x[j] = ld_coherent32(tab.heap[q] + so + (unsigned long long)v * 32);
          }
        }
#pragma unroll
        for (int j = 0; j < kCB; ++j) {
          const int v = v0 + j * kThreadsC;
          if (v < nv) {
            const unsigned long long o = b0 + off + (unsigned long long)v * 32;
            if (a.fanout) {
#pragma unroll
              for (int t = 0; t < NT; ++t) st_stream32((char*)a.bufs[erow][t] + o, x[j]);
            } else {
              st_stream32((char*)a.bufs[erow][a.dst_index] + o, x[j]);
            }
          }
        }
      }
      named_bar(3, kThreadsC);  // every load of the chunk has returned (its value was stored above)
      if constexpr (!PUSH) {
        if (ct < world) st_relaxed_sys(fc(ct, sl, rank), base + (uint32_t)k + 1u);
      }
      if (ct == 0) {
        c_done = (unsigned)(k + 1);
        WS_TRACE(7, k);
      }
    }
  }
  __syncthreads();
  // nobody may reuse the ring (next launch) while a peer still pulls from it
  if (!PUSH && tid < world && !(a.dbg & 2)) {
    const long long k0 = mine > kSlots ? mine - kSlots : 0;
    for (long long k = k0; k < mine; ++k) wait_flag(tab, fc(rank, (int)(k % kSlots), tid), base + (uint32_t)k + 1u);
  }
  if (threadIdx.x == 0) *(volatile uint32_t*)ws_epoch = base + (uint32_t)mine;
}

template <typename T, int OP, bool NVLS, int NT, bool PUSH>
void launch_pull_mode(const ThreadArgs& args, int grid, cudaStream_t stream) {
  typedef SplitP<6, 4, 4> SP;
  auto kfn = thread_fused_pull_kernel<T, OP, NVLS, NT, SP, PUSH>;
  if (args.tab.emulate) {
    void* params[] = {(void*)&args};
    MP4J_CUDA(cudaLaunchCooperativeKernel((const void*)kfn, dim3(grid, args.tab.world), dim3(SP::kThreads), params, 0,
                                          stream));
  } else {
    kfn<<<grid, SP::kThreads, 0, stream>>>(args);
    MP4J_CUDA(cudaGetLastError());
  }
}

template <typename T, int OP, bool NVLS, int NT>
void launch_pull_nt(const ThreadArgs& args, int grid, cudaStream_t stream) {
  if (args.push) return launch_pull_mode<T, OP, NVLS, NT, true>(args, grid, stream);
  launch_pull_mode<T, OP, NVLS, NT, false>(args, grid, stream);
}

template <typename T, int OP, bool NVLS>
void launch_pull_t(const ThreadArgs& args, int grid, cudaStream_t stream) {
  switch (args.T) {
    case 2: return launch_pull_nt<T, OP, NVLS, 2>(args, grid, stream);
    case 4: return launch_pull_nt<T, OP, NVLS, 4>(args, grid, stream);
    case 8: return launch_pull_nt<T, OP, NVLS, 8>(args, grid, stream);
  }
  throw Error("fused thread kernel (pull) is built for T in {2, 4, 8}");
}

template <typename T, int OP>
void launch_pull_op(const ThreadArgs& args, bool nvls, int grid, cudaStream_t stream) {
  if constexpr (Nvls<T, OP>::supported) {
    if (nvls) return launch_pull_t<T, OP, true>(args, grid, stream);
  }
  launch_pull_t<T, OP, false>(args, grid, stream);
}

template <typename T>
void launch_pull_type(int op, const ThreadArgs& args, bool nvls, int grid, cudaStream_t stream) {
  switch (op) {
    case MP4J_SUM: return launch_pull_op<T, MP4J_SUM>(args, nvls, grid, stream);
    case MP4J_MAX: return launch_pull_op<T, MP4J_MAX>(args, nvls, grid, stream);
    case MP4J_MIN: return launch_pull_op<T, MP4J_MIN>(args, nvls, grid, stream);
  }
  throw Error("fused thread kernel (pull) is built for SUM/MAX/MIN");
}

}  // namespace

// Same preconditions as launch_thread_fused_ws (32-byte aligned, T in {2,4,8}, SUM/MAX/MIN on 4/8-byte types).
void launch_thread_fused_pull(int dtype, int op, bool nvls, const ThreadArgs& args, int grid, cudaStream_t stream) {
  switch (dtype) {
    case MP4J_F64: return launch_pull_type<double>(op, args, nvls, grid, stream);
    case MP4J_F32: return launch_pull_type<float>(op, args, nvls, grid, stream);
    case MP4J_I64: return launch_pull_type<int64_t>(op, args, nvls, grid, stream);
    case MP4J_I32: return launch_pull_type<int32_t>(op, args, nvls, grid, stream);
  }
  throw Error("fused thread kernel (pull) is built for 4/8-byte types");
}

}  // namespace mp4j
