// ThreadCommSlave on the GPU: the T "threads" of a process are T device buffers on one GPU.
//
// thread_allreduce_kernel fuses, in ONE launch (S13/S14/S6 of SURVEY §2.6; reference: lg T
// full-length Exchanger rounds TCS:259-303 + ring allreduce PCS:1733-1763 + T-1 copies DOP:362-383):
//   phase A  reduce the T local buffers in registers (thread-id order)      -> this block's staging slot
//   phase B  inter-GPU: rank r reduces sub-slice r of the slot over all peers (rank order, or in the
//            switch with multimem.ld_reduce) and pushes it to every peer's slot (or multimem.st)
//   phase C  fan the finished slot out to the T local buffers
// Blocks are persistent and chunk-synchronous: block b of every rank walks the same chunks, so all
// cross-GPU synchronisation is block-b-to-block-b flags; while one block waits on NVLink the other
// blocks of the SM stream HBM, which overlaps the two phases without a grid-wide barrier. Only
// 1/T of the data ever crosses NVLink. With world == 1 phases A and C collapse into one pass.
//
// thread_cluster_kernel is the thread-block-cluster / DSMEM formulation of the intra-GPU step
// (world == 1): each CTA of a cluster bulk-loads (TMA 1-D, cp.async.bulk + mbarrier) a tile of
// "its" thread buffers into shared memory, the cluster reduce-scatters the tile through
// distributed shared memory (CTA j sums sub-tile j from every CTA's smem) and each CTA fans its
// sub-tile out to all T buffers.
#include <cooperative_groups.h>

#include <algorithm>

#include "cuda_check.h"
#include "dense.cuh"
#include "launch.h"

namespace cg = cooperative_groups;

namespace mp4j {

namespace {

constexpr int kTThreads = 256;

// ---- access unit: 16-byte vector (fast path) or a single element (unaligned head/tail)
template <typename T, int OP, bool VEC>
struct Unit;
template <typename T, int OP>
struct Unit<T, OP, true> {
  using V = Vec16;
  static constexpr int BYTES = 16;
  __device__ __forceinline__ static V ld_in(const void* p) { return ld_stream(p); }
  __device__ __forceinline__ static V ld_co(const void* p) { return ld_coherent(p); }
  __device__ __forceinline__ static void st(void* p, const V& v) { st_stream(p, v); }
  __device__ __forceinline__ static V red(const V& a, const V& b) { return vec_reduce<T, OP>(a, b); }
};
template <typename T, int OP>
struct Unit<T, OP, false> {
  using V = T;
  static constexpr int BYTES = sizeof(T);
  __device__ __forceinline__ static V ld_in(const void* p) { return *(const volatile T*)p; }
  __device__ __forceinline__ static V ld_co(const void* p) { return *(const volatile T*)p; }
  __device__ __forceinline__ static void st(void* p, const V& v) { *(volatile T*)p = v; }
  __device__ __forceinline__ static V red(const V& a, const V& b) { return Reduce<T, OP>::apply(a, b); }
};

// acc = bufs[0][off] (+) bufs[1][off] (+) ... in thread-id order, loads issued 8 at a time
template <typename U>
__device__ __forceinline__ typename U::V reduce_threads(void* const* bufs, int nT, unsigned long long off) {
  typename U::V acc;
  for (int t0 = 0; t0 < nT; t0 += 8) {
    typename U::V x[8];
#pragma unroll
    for (int j = 0; j < 8; ++j)
      if (t0 + j < nT) x[j] = U::ld_in((const char*)bufs[t0 + j] + off);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (t0 + j < nT) acc = (t0 + j == 0) ? x[j] : U::red(acc, x[j]);
    }
  }
  return acc;
}

template <typename U>
__device__ __forceinline__ void fan_out(void* const* bufs, int nT, int fanout, int dst, unsigned long long off,
                                        const typename U::V& v) {
  if (fanout) {
    for (int t = 0; t < nT; ++t) U::st((char*)bufs[t] + off, v);
  } else {
    U::st((char*)bufs[dst] + off, v);
  }
}

template <typename T, int OP, bool NVLS, bool VEC>
__global__ void __launch_bounds__(kTThreads, 2) thread_allreduce_kernel(const __grid_constant__ ThreadArgs a) {
  using U = Unit<T, OP, VEC>;
  using V = typename U::V;
  constexpr int UB = U::BYTES;
  const PeerTable& tab = a.tab;
  const int rank = my_rank(tab);
  const int world = tab.world;
  void* const* bufs = a.bufs[tab.emulate ? rank : 0];
  const int nT = a.T;
  const unsigned long long b0 = (unsigned long long)a.from * sizeof(T);
  const long long nunits = (a.to - a.from) * (long long)sizeof(T) / UB;

  if (world == 1) {
    // intra-GPU only: one pass, T loads -> reduce -> T stores
    const long long gstride = (long long)gridDim.x * blockDim.x;
    for (long long u = (long long)blockIdx.x * blockDim.x + threadIdx.x; u < nunits; u += gstride) {
      const unsigned long long off = b0 + (unsigned long long)u * UB;
      V acc = reduce_threads<U>(bufs, nT, off);
      fan_out<U>(bufs, nT, a.fanout, a.dst_index, off, acc);
    }
    return;
  }

  BlockEpoch ep;
  ep.load(tab, rank);
  const long long units_per_slot = (long long)(a.slot_bytes / UB);
  const unsigned long long slot_off = a.stg_off + (unsigned long long)blockIdx.x * a.slot_bytes;
  char* my_slot = tab.heap[rank] + slot_off;

  for (long long c0 = (long long)blockIdx.x * units_per_slot; c0 < nunits;
       c0 += (long long)gridDim.x * units_per_slot) {
    const long long len = min(units_per_slot, nunits - c0);
    // ---- phase A: local T-way reduce into my slot
    for (long long u = threadIdx.x; u < len; u += blockDim.x) {
      V acc = reduce_threads<U>(bufs, nT, b0 + (unsigned long long)(c0 + u) * UB);
      U::st(my_slot + u * UB, acc);
    }
    if (!block_barrier_peers(tab, rank, ep)) break;  // peer missing: store nothing more
    // ---- phase B: inter-GPU reduce of sub-slice `rank`, pushed to every peer
    const long long lo = len * rank / world, hi = len * (rank + 1) / world;
    for (long long u = lo + threadIdx.x; u < hi; u += blockDim.x) {
      const unsigned long long o = slot_off + (unsigned long long)u * UB;
      if constexpr (NVLS && VEC) {
        Vec16 acc = Nvls<T, OP>::ld_reduce(tab.mc_heap + o);
        multimem_st(tab.mc_heap + o, acc);
      } else {
        V x[kMaxRanks];
#pragma unroll
        for (int q = 0; q < kMaxRanks; ++q)
          if (q < world) x[q] = U::ld_co(tab.heap[q] + o);
        V acc = x[0];
#pragma unroll
        for (int q = 1; q < kMaxRanks; ++q)
          if (q < world) acc = U::red(acc, x[q]);
#pragma unroll
        for (int q = 0; q < kMaxRanks; ++q)
          if (q < world) U::st(tab.heap[q] + o, acc);
      }
    }
    if (!block_barrier_peers(tab, rank, ep)) break;
    // ---- phase C: fan the finished chunk out to the T thread buffers
    for (long long u = threadIdx.x; u < len; u += blockDim.x) {
      V x = U::ld_co(my_slot + u * UB);
      fan_out<U>(bufs, nT, a.fanout, a.dst_index, b0 + (unsigned long long)(c0 + u) * UB, x);
    }
  }
  ep.store();
}

// ------------------------------------------------------------------------------------------
// Intra-GPU fast path (world == 1): static thread count, 256-bit global accesses (LDG.E.256 /
// STG.E.256, new on sm_100) with L2 evict-first (every byte is touched once), and an
// over-subscribed grid so the hardware block scheduler balances SMs of unequal L2 distance.
// tools/tune_thread_reduce.cu measured this shape at 6.5 TB/s on B200 (128-bit + persistent
// grid: 5.9 TB/s; cudaMemcpy D2D: 6.3 TB/s).
constexpr int kLThreads = 128;

template <typename T, int OP, int NT>
__global__ void __launch_bounds__(kLThreads, 4) thread_local_kernel(const __grid_constant__ ThreadArgs a) {
  const unsigned long long b0 = (unsigned long long)a.from * sizeof(T);
  const long long nunits = (a.to - a.from) * (long long)sizeof(T) / 32;
  const long long gstride = (long long)gridDim.x * kLThreads;
  for (long long u = (long long)blockIdx.x * kLThreads + threadIdx.x; u < nunits; u += gstride) {
    const unsigned long long off = b0 + (unsigned long long)u * 32;
    Vec32 x[NT];
#pragma unroll
    for (int t = 0; t < NT; ++t) x[t] = ld_stream32((const char*)a.bufs[0][t] + off);
#pragma unroll
    for (int t = 1; t < NT; ++t) {  // thread-id order, as everywhere
      x[0].lo = vec_reduce<T, OP>(x[0].lo, x[t].lo);
      x[0].hi = vec_reduce<T, OP>(x[0].hi, x[t].hi);
    }
    if (a.fanout) {
#pragma unroll
      for (int t = 0; t < NT; ++t) st_stream32((char*)a.bufs[0][t] + off, x[0]);
    } else {
      st_stream32((char*)a.bufs[0][a.dst_index] + off, x[0]);
    }
  }
}

template <typename T, int OP, int NT>
void launch_local_fast(const ThreadArgs& args, cudaStream_t stream) {
  int dev = 0, sms = 0;
  MP4J_CUDA(cudaGetDevice(&dev));
  MP4J_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const long long nunits = (args.to - args.from) * (long long)sizeof(T) / 32;
  const long long want = (nunits + kLThreads - 1) / kLThreads;
  const int grid = (int)std::max<long long>(1, std::min<long long>(want, (long long)sms * 16));
  thread_local_kernel<T, OP, NT><<<grid, kLThreads, 0, stream>>>(args);
  MP4J_CUDA(cudaGetLastError());
}

// true if (T, OP) has the static-T 256-bit specialisations
template <typename T, int OP>
constexpr bool has_local_fast() {
  return sizeof(T) >= 4 && (OP == MP4J_SUM || OP == MP4J_MAX || OP == MP4J_MIN);
}

// ------------------------------------------------------------------------------------------
// Cluster / DSMEM variant (world == 1).
constexpr int kCThreads = 256;
constexpr int kTileBytes = 16 * 1024;  // per thread buffer per step
constexpr int kStages = 3;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE;\n"
      "bra WAIT_LOOP;\n"
      "DONE:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// TMA 1-D bulk copy global -> shared, completion on an mbarrier
__device__ __forceinline__ void bulk_load(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(smem_dst)),
               "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ Vec16 ld_dsmem(uint32_t local_addr, uint32_t cta) {
  uint32_t remote;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local_addr), "r"(cta));
  Vec16 v;
  asm volatile("ld.shared::cluster.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "r"(remote)
               : "memory");
  return v;
}

template <typename T, int OP, int CS>
__global__ void __launch_bounds__(kCThreads) thread_cluster_kernel(const __grid_constant__ ThreadArgs a) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t full[kStages];
  cg::cluster_group cluster = cg::this_cluster();
  const uint32_t cta = cluster.block_rank();
  const int nT = a.T;
  const int tpc = nT / CS;  // thread buffers per CTA
  void* const* bufs = a.bufs[0];
  const unsigned long long b0 = (unsigned long long)a.from * sizeof(T);
  const long long nbytes = (a.to - a.from) * (long long)sizeof(T);
  const long long ntiles = (nbytes + kTileBytes - 1) / kTileBytes;
  const long long ncluster = gridDim.x / CS, cid = blockIdx.x / CS;
  // smem: [stage][tpc][tile]
  auto tile_ptr = [&](int s, int j) { return smem_raw + ((size_t)s * tpc + j) * kTileBytes; };

  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) mbar_init(&full[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  auto issue = [&](long long k) {  // tile index within this cluster's sequence
    const long long tile = cid + k * ncluster;
    if (tile >= ntiles) return;
    const int s = (int)(k % kStages);
    const long long off = tile * kTileBytes;
    const uint32_t bytes = (uint32_t)min((long long)kTileBytes, nbytes - off);
    mbar_expect_tx(&full[s], bytes * tpc);
    for (int j = 0; j < tpc; ++j)
      bulk_load(tile_ptr(s, j), (const char*)bufs[cta + j * CS] + b0 + off, bytes, &full[s]);
  };

  if (threadIdx.x == 0) {
    issue(0);
    issue(1);
  }
  for (long long k = 0;; ++k) {
    const long long tile = cid + k * ncluster;
    if (tile >= ntiles) break;  // uniform across the cluster
    const int s = (int)(k % kStages);
    const long long off = tile * kTileBytes;
    const int nvec = (int)(min((long long)kTileBytes, nbytes - off) / 16);
    mbar_wait(&full[s], (uint32_t)((k / kStages) & 1));
    // fold this CTA's tpc tiles into tile 0 (thread-id order within the CTA's subset)
    if (tpc > 1) {
      for (int v = threadIdx.x; v < nvec; v += blockDim.x) {
        Vec16 acc = *(const Vec16*)(tile_ptr(s, 0) + v * 16);
        for (int j = 1; j < tpc; ++j) acc = vec_reduce<T, OP>(acc, *(const Vec16*)(tile_ptr(s, j) + v * 16));
        *(Vec16*)(tile_ptr(s, 0) + v * 16) = acc;
      }
    }
    // all CTAs: tile k is in smem everywhere, and everyone finished reading tile k-1
    cluster.sync();
    if (threadIdx.x == 0) issue(k + 2);  // stage (k+2)%3 held tile k-1: free now
    // reduce-scatter through DSMEM: this CTA owns vectors [cta*nvec/CS, (cta+1)*nvec/CS)
    const int lo = (int)((long long)nvec * cta / CS), hi = (int)((long long)nvec * (cta + 1) / CS);
    const uint32_t base = smem_u32(tile_ptr(s, 0));
    for (int v = lo + threadIdx.x; v < hi; v += blockDim.x) {
      Vec16 x[CS];
#pragma unroll
      for (int c = 0; c < CS; ++c) x[c] = ld_dsmem(base + v * 16, c);
      Vec16 acc = x[0];
#pragma unroll
      for (int c = 1; c < CS; ++c) acc = vec_reduce<T, OP>(acc, x[c]);
      const unsigned long long goff = b0 + off + (unsigned long long)v * 16;
      if (a.fanout) {
        for (int t = 0; t < nT; ++t) st_stream((char*)bufs[t] + goff, acc);
      } else {
        st_stream((char*)bufs[a.dst_index] + goff, acc);
      }
    }
  }
  cluster.sync();  // nobody exits while a peer may still read its shared memory
}

// ------------------------------------------------------------------------------------------
template <typename T, int OP, bool NVLS, bool VEC>
void launch_reg(const ThreadArgs& args, int grid, cudaStream_t stream) {
  auto kfn = thread_allreduce_kernel<T, OP, NVLS, VEC>;
  if (args.tab.emulate && args.tab.world > 1) {
    void* params[] = {(void*)&args};
    MP4J_CUDA(cudaLaunchCooperativeKernel((const void*)kfn, dim3(grid, args.tab.world), dim3(kTThreads), params, 0,
                                          stream));
  } else {
    kfn<<<grid, kTThreads, 0, stream>>>(args);
    MP4J_CUDA(cudaGetLastError());
  }
}

template <typename T, int OP, int CS>
void launch_cluster_cs(const ThreadArgs& args, cudaStream_t stream) {
  auto kfn = thread_cluster_kernel<T, OP, CS>;
  const int tpc = args.T / CS;
  const size_t smem = (size_t)kStages * tpc * kTileBytes;
  static bool configured = false;
  if (!configured) {
    MP4J_CUDA(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(200 * 1024)));
    configured = true;
  }
  int dev = 0, sms = 0;
  MP4J_CUDA(cudaGetDevice(&dev));
  MP4J_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int per_sm = std::max(1, std::min(2, (int)((200 * 1024) / smem)));
  const long long ntiles = ((args.to - args.from) * (long long)sizeof(T) + kTileBytes - 1) / kTileBytes;
  long long clusters = std::min<long long>((long long)(sms / CS) * per_sm, std::max<long long>(ntiles, 1));
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((unsigned)(clusters * CS));
  cfg.blockDim = dim3(kCThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CS;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  MP4J_CUDA(cudaLaunchKernelEx(&cfg, kfn, args));
}

template <typename T, int OP>
void launch_cluster(const ThreadArgs& args, cudaStream_t stream) {
  const int T_ = args.T;
  if (T_ % 8 == 0) return launch_cluster_cs<T, OP, 8>(args, stream);
  if (T_ % 4 == 0) return launch_cluster_cs<T, OP, 4>(args, stream);
  if (T_ % 2 == 0) return launch_cluster_cs<T, OP, 2>(args, stream);
  throw Error("cluster thread-reduce needs an even thread count");
}

template <typename T, int OP>
void launch_any(const ThreadArgs& args, bool nvls, bool vec, int grid, int algo, cudaStream_t stream) {
  if (algo == 1) {
    if constexpr (sizeof(T) >= 4 && (OP == MP4J_SUM || OP == MP4J_MAX || OP == MP4J_MIN)) {
      return launch_cluster<T, OP>(args, stream);
    } else {
      throw Error("cluster thread-reduce is built for SUM/MAX/MIN on 4/8-byte types");
    }
  }
  if (algo == 2) {  // host verified: world == 1, 32-byte aligned, T in {2, 4, 8}
    if constexpr (has_local_fast<T, OP>()) {
      switch (args.T) {
        case 2: return launch_local_fast<T, OP, 2>(args, stream);
        case 4: return launch_local_fast<T, OP, 4>(args, stream);
        case 8: return launch_local_fast<T, OP, 8>(args, stream);
      }
    }
    throw Error("256-bit thread-reduce fast path not built for this (dtype, op, T)");
  }
  if (!vec) return launch_reg<T, OP, false, false>(args, grid, stream);
  if constexpr (Nvls<T, OP>::supported) {
    if (nvls) return launch_reg<T, OP, true, true>(args, grid, stream);
  }
  launch_reg<T, OP, false, true>(args, grid, stream);
}

template <typename T, bool INTEGRAL, bool LOC>
void t_by_op(int op, const ThreadArgs& args, bool nvls, bool vec, int grid, int algo, cudaStream_t stream) {
  switch (op) {
    case MP4J_SUM: return launch_any<T, MP4J_SUM>(args, nvls, vec, grid, algo, stream);
    case MP4J_MAX: return launch_any<T, MP4J_MAX>(args, nvls, vec, grid, algo, stream);
    case MP4J_MIN: return launch_any<T, MP4J_MIN>(args, nvls, vec, grid, algo, stream);
    case MP4J_PROD: return launch_any<T, MP4J_PROD>(args, false, vec, grid, algo, stream);
  }
  if constexpr (INTEGRAL) {
    switch (op) {
      case MP4J_BAND: return launch_any<T, MP4J_BAND>(args, nvls, vec, grid, algo, stream);
      case MP4J_BOR: return launch_any<T, MP4J_BOR>(args, nvls, vec, grid, algo, stream);
      case MP4J_BXOR: return launch_any<T, MP4J_BXOR>(args, nvls, vec, grid, algo, stream);
    }
  }
  if constexpr (LOC) {
    switch (op) {
      case MP4J_MAXLOC: return launch_any<T, MP4J_MAXLOC>(args, false, vec, grid, algo, stream);
      case MP4J_MINLOC: return launch_any<T, MP4J_MINLOC>(args, false, vec, grid, algo, stream);
    }
  }
  throw Error("operator " + std::to_string(op) + " is not defined for this dtype on the GPU path");
}

}  // namespace

int thread_max_blocks_per_sm() {
  static int cached = -1;
  if (cached < 0) {
    int nb = 0;
    MP4J_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(
        &nb, thread_allreduce_kernel<float, MP4J_SUM, false, true>, kTThreads, 0));
    cached = std::max(nb, 1);
  }
  return cached;
}

int thread_threads() { return kTThreads; }

bool thread_local_fast_supported(int dtype, int op, int T) {
  if (!(T == 2 || T == 4 || T == 8)) return false;
  if (!(op == MP4J_SUM || op == MP4J_MAX || op == MP4J_MIN)) return false;
  return dtype == MP4J_F64 || dtype == MP4J_F32 || dtype == MP4J_I64 || dtype == MP4J_I32;
}

bool thread_cluster_supported(int dtype, int op, int T, int world) {
  if (world != 1 || T % 2 != 0 || T > 64) return false;
  if (!(op == MP4J_SUM || op == MP4J_MAX || op == MP4J_MIN)) return false;
  // smem: stages * (T/CS) * tile must fit
  const int cs = T % 8 == 0 ? 8 : (T % 4 == 0 ? 4 : 2);
  if ((size_t)kStages * (T / cs) * kTileBytes > 200 * 1024) return false;
  return dtype == MP4J_F64 || dtype == MP4J_F32 || dtype == MP4J_I64 || dtype == MP4J_I32;
}

// `vec` is decided by the caller: every buffer and the byte range are 16-byte aligned.
void launch_thread_allreduce(int dtype, int op, bool nvls, const ThreadArgs& args, int grid, int algo,
                             cudaStream_t stream) {
  const bool vec = (algo & 0x100) == 0;
  algo &= 0xff;
  switch (dtype) {
    case MP4J_F64: return t_by_op<double, false, true>(op, args, nvls, vec, grid, algo, stream);
    case MP4J_F32: return t_by_op<float, false, false>(op, args, nvls, vec, grid, algo, stream);
    case MP4J_I64: return t_by_op<int64_t, true, true>(op, args, nvls, vec, grid, algo, stream);
    case MP4J_I32: return t_by_op<int32_t, true, false>(op, args, nvls, vec, grid, algo, stream);
    case MP4J_I16: return t_by_op<int16_t, true, false>(op, args, false, vec, grid, algo, stream);
    case MP4J_I8: return t_by_op<int8_t, true, false>(op, args, false, vec, grid, algo, stream);
  }
  throw Error("dtype " + std::to_string(dtype) + " is not supported on the GPU path");
}

}  // namespace mp4j
