// Warp-specialised fused ThreadCommSlave allreduce (world > 1): the production path of
// "T thread buffers per GPU x p GPUs" (TCS:1915-1963 on CUDA tensors).
//
// One persistent CTA per SM, three warp groups that run CONCURRENTLY on a ring of staging slots:
//   group A (HBM read)   : slot[k] = reduce_t bufs[t][chunk k]          (256-bit loads, T static)
//   group B (NVLink)     : rank r reduces sub-slice r of slot[k] over all peers and pushes it back
//                          to every peer - multimem.ld_reduce + multimem.st (NVLS) or P2P ld/st
//   group C (HBM write)  : bufs[t][chunk k] = slot[k] for all t         (256-bit stores)
// Dependencies are flags, never a CTA-wide or grid-wide barrier:
//   A(k) needs C(k - kSlots) done (local shared-memory counter)
//   B(k) needs A(k) of block b on EVERY rank   (flag array FA in the peers' signal pads)
//   C(k) needs B(k) of block b on EVERY rank   (flag array FB)
// so the HBM read stream, the NVLink exchange and the HBM write stream of different chunks overlap
// inside every SM, and 1/T of the bytes cross NVLink. The chunk-synchronous kernel in
// thread_kernels.cu (phases in lock-step) is kept as the generic fallback (odd alignments, other
// thread counts / operators).
#include <algorithm>

#include "cuda_check.h"
#include "dense.cuh"
#include "launch.h"

namespace mp4j {

namespace {

constexpr int kWarpsA = 6, kWarpsB = 4, kWarpsC = 6;
constexpr int kThreadsA = kWarpsA * 32, kThreadsB = kWarpsB * 32, kThreadsC = kWarpsC * 32;
constexpr int kWsThreads = kThreadsA + kThreadsB + kThreadsC;  // 512
constexpr int kMaxSlots = 8;  // ring depth is a launch parameter (ThreadArgs::slots), 2..kMaxSlots

__device__ __forceinline__ void named_bar(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

template <typename T, int OP, bool NVLS, int NT>
__global__ void __launch_bounds__(kWsThreads, 1) thread_fused_ws_kernel(const __grid_constant__ ThreadArgs a) {
  __shared__ volatile unsigned c_done;  // chunks finished by group C (slot recycling)
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
  if (threadIdx.x == 0) c_done = 0;
  __syncthreads();

  const int tid = threadIdx.x;
  if (tid < kThreadsA) {
    // ================= group A: T-way reduce of the local thread buffers into the slot ring
    for (long long k = 0; k < mine; ++k) {
      const long long c = blockIdx.x + k * gridDim.x;
      const long long off = c * ch;
      const int nv = (int)(min(ch, nbytes - off) / 32);
      if (k >= kSlots) {
        if (tid == 0)
          while (c_done < (unsigned)(k - kSlots + 1)) {
          }
        named_bar(1, kThreadsA);
      }
      char* slot = tab.heap[rank] + ring_off + (unsigned long long)(k % kSlots) * a.slot_bytes;
      for (int v = tid; v < nv; v += kThreadsA) {
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
      named_bar(1, kThreadsA);  // all of A's stores precede the release below
      if (tid < world)
        st_release_sys((uint32_t*)((char*)tab.sig[tid] + kSigWsFaOff) + blockIdx.x * kMaxRanks + rank,
                       base + (uint32_t)k + 1u);
    }
  } else if (tid < kThreadsA + kThreadsB) {
    // ================= group B: inter-GPU reduce-scatter + allgather of the slot over NVLink
    const int bt = tid - kThreadsA;
    for (long long k = 0; k < mine; ++k) {
      const long long c = blockIdx.x + k * gridDim.x;
      const long long off = c * ch;
      const int nv16 = (int)(min(ch, nbytes - off) / 16);
      if (bt < world)
        wait_flag(tab, (uint32_t*)((char*)tab.sig[rank] + kSigWsFaOff) + blockIdx.x * kMaxRanks + bt,
                  base + (uint32_t)k + 1u);
      named_bar(2, kThreadsB);
      const unsigned long long so = ring_off + (unsigned long long)(k % kSlots) * a.slot_bytes;
      const int lo = (int)((long long)nv16 * rank / world), hi = (int)((long long)nv16 * (rank + 1) / world);
      // An NVLink round trip is ~2-3 us: keep kBatch independent 16-byte requests in flight per
      // thread (the asm volatile accessors are never reordered, so batching is explicit).
      constexpr int kBatch = 8;
      for (int v0 = lo + bt; v0 < hi; v0 += kThreadsB * kBatch) {
        Vec16 acc[kBatch];
        if constexpr (NVLS) {
#pragma unroll
          for (int j = 0; j < kBatch; ++j) {
            const int v = v0 + j * kThreadsB;
            if (v < hi) acc[j] = Nvls<T, OP>::ld_reduce(tab.mc_heap + so + (unsigned long long)v * 16);
          }
#pragma unroll
          for (int j = 0; j < kBatch; ++j) {
            const int v = v0 + j * kThreadsB;
            if (v < hi) multimem_st(tab.mc_heap + so + (unsigned long long)v * 16, acc[j]);
          }
        } else {
          for (int q = 0; q < world; ++q) {  // ascending rank order => identical bits on every rank
            Vec16 x[kBatch];
#pragma unroll
            for (int j = 0; j < kBatch; ++j) {
              const int v = v0 + j * kThreadsB;
              if (v < hi) x[j] = ld_coherent(tab.heap[q] + so + (unsigned long long)v * 16);
            }
#pragma unroll
            for (int j = 0; j < kBatch; ++j) acc[j] = (q == 0) ? x[j] : vec_reduce<T, OP>(acc[j], x[j]);
          }
          for (int q = 0; q < world; ++q) {
#pragma unroll
            for (int j = 0; j < kBatch; ++j) {
              const int v = v0 + j * kThreadsB;
              if (v < hi) st_stream(tab.heap[q] + so + (unsigned long long)v * 16, acc[j]);
            }
          }
        }
      }
      named_bar(2, kThreadsB);
      if (bt < world)
        st_release_sys((uint32_t*)((char*)tab.sig[bt] + kSigWsFbOff) + blockIdx.x * kMaxRanks + rank,
                       base + (uint32_t)k + 1u);
    }
  } else {
    // ================= group C: fan the finished slot out to the T thread buffers
    const int ct = tid - kThreadsA - kThreadsB;
    for (long long k = 0; k < mine; ++k) {
      const long long c = blockIdx.x + k * gridDim.x;
      const long long off = c * ch;
      const int nv = (int)(min(ch, nbytes - off) / 32);
      if (ct < world)
        wait_flag(tab, (uint32_t*)((char*)tab.sig[rank] + kSigWsFbOff) + blockIdx.x * kMaxRanks + ct,
                  base + (uint32_t)k + 1u);
      named_bar(3, kThreadsC);
      const char* slot = tab.heap[rank] + ring_off + (unsigned long long)(k % kSlots) * a.slot_bytes;
      for (int v = ct; v < nv; v += kThreadsC) {
        const Vec32 x = ld_coherent32(slot + (unsigned long long)v * 32);
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
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) *(volatile uint32_t*)ws_epoch = base + (uint32_t)mine;
}

template <typename T, int OP, bool NVLS, int NT>
void launch_ws_nt(const ThreadArgs& args, int grid, cudaStream_t stream) {
  auto kfn = thread_fused_ws_kernel<T, OP, NVLS, NT>;
  if (args.tab.emulate) {
    void* params[] = {(void*)&args};
    MP4J_CUDA(cudaLaunchCooperativeKernel((const void*)kfn, dim3(grid, args.tab.world), dim3(kWsThreads), params, 0,
                                          stream));
  } else {
    kfn<<<grid, kWsThreads, 0, stream>>>(args);
    MP4J_CUDA(cudaGetLastError());
  }
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
