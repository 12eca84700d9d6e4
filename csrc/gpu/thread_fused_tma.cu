// Fused ThreadCommSlave allreduce (world > 1), second generation: the HBM streams of the kernel run
// through the TMA unit instead of through registers (TCS:1915-1963 on CUDA tensors).
//
// Same dataflow as thread_fused_ws.cu (T-way local reduce -> slot ring -> inter-GPU reduce-scatter +
// allgather of the slot over NVLink -> fan-out to the T buffers), same flags, but
//   * phase A is a TMA pipeline: one thread issues `cp.async.bulk` loads of one tile per thread
//     buffer into a 5-stage shared-memory ring (160 KiB of reads in flight per SM instead of the 48 KiB
//     that fit in the registers of the load warps), four reducer warps fold the T tiles from shared
//     memory in thread-id order and write the slot;
//   * phase C is a TMA pipeline too: one warp bulk-loads finished slot tiles into shared memory and
//     bulk-stores each tile T times (once per thread buffer) - no registers, no LSU traffic;
//   * T is a run-time value (any 1..64) and only 16-byte alignment is needed;
//   * phase B runs as teams of two warps per chunk with 16 independent 16-byte requests per lane, so a
//     chunk's share is one NVLink round trip at 8 ranks.
// The read stream therefore has head-room to catch up after a stall, which the register formulation
// (load warps at exactly the average rate) did not have.
#include <stdlib.h>

#include <algorithm>

#include "cuda_check.h"
#include "dense.cuh"
#include "launch.h"

namespace mp4j {

namespace {

constexpr int kRW = 4;                   // reducer warps (phase A consumers)
constexpr int kWB = 8;                   // NVLink warps (phase B), teams of kWPC
constexpr int kWPC = 2;                  // warps per chunk in phase B
constexpr int kTeams = kWB / kWPC;
constexpr int kStageA = 32 * 1024;       // bytes per phase-A stage: T tiles of kStageA / T bytes
constexpr int kStagesA = 5;
constexpr int kTileC = 16 * 1024;        // phase-C tile
constexpr int kStagesC = 4;
constexpr int kTmaThreads = 32 * (1 + kRW + 1 + kWB + 1);  // producer | reducers | signal | B | C
constexpr size_t kTmaSmem = (size_t)kStagesA * kStageA + (size_t)kStagesC * kTileC;

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_addr(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n.reg .pred p;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "selp.u32 %0, 1, 0, p;\n}\n"
      : "=r"(ok)
      : "r"(smem_addr(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  while (!mbar_try(bar, parity)) {
  }
}
__device__ __forceinline__ uint64_t policy_evict_first() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint64_t policy_evict_last() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  return p;
}
// TMA 1-D bulk copies
__device__ __forceinline__ void bulk_g2s(void* sdst, const void* gsrc, uint32_t bytes, uint64_t* bar, uint64_t pol) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
          smem_addr(sdst)),
      "l"(gsrc), "r"(bytes), "r"(smem_addr(bar)), "l"(pol)
      : "memory");
}
__device__ __forceinline__ void bulk_s2g(void* gdst, const void* ssrc, uint32_t bytes, uint64_t pol) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;" ::"l"(gdst),
               "r"(smem_addr(ssrc)), "r"(bytes), "l"(pol)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async;" ::: "memory"); }

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
__device__ __forceinline__ Vec16 lds16(const void* p) {
  Vec16 v;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "r"(smem_addr(p)));
  return v;
}
__device__ __forceinline__ void st_keep16(void* p, const Vec16& v, uint64_t pol) {  // staging: re-read from L2 soon
  asm volatile("st.global.L2::cache_hint.v4.u32 [%0], {%1,%2,%3,%4}, %5;" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z),
               "r"(v.w), "l"(pol)
               : "memory");
}


// diagnostics: stage timestamps of block 0 (dbg bit 128), read back by tools/ws_trace.py
#define WS_TRACE(ev, k)                                                                                  \
  do {                                                                                                   \
    if ((a.dbg & 128) && blockIdx.x == 0 && (k) < kWsTraceChunks)                                        \
      ((unsigned long long*)((char*)tab.sig[rank] + kSigWsTraceOff))[(k) * 8 + (ev)] = globaltimer_ns(); \
  } while (0)

// Walks the chunks of one block: chunk k covers bytes [off, off + len) of the range.
struct ChunkIter {
  long long nbytes, ch;
  long long first, stride;
  __device__ __forceinline__ long long off(long long k) const { return (first + k * stride) * ch; }
  __device__ __forceinline__ long long len(long long k) const { return min(ch, nbytes - off(k)); }
};

template <typename T, int OP, bool NVLS>
__global__ void __launch_bounds__(kTmaThreads, 1) thread_fused_tma_kernel(const __grid_constant__ ThreadArgs a) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ __align__(8) uint64_t fullA[kStagesA], emptyA[kStagesA], fullC[kStagesC];
  __shared__ volatile unsigned c_done;  // chunks whose slot phase C has fully read (slot recycling)
  // arrivals of the phase-A warps per ring slot. Per SLOT, not one running total: a warp that is a chunk ahead
  // of a slow one must not complete the slow warp's chunk (a warp cannot lap the ring, so per-slot counts are exact)
  __shared__ unsigned a_cnt[kWsMaxSlots];
  const PeerTable& tab = a.tab;
  const int rank = my_rank(tab);
  const int world = tab.world;
  const int erow = tab.emulate ? rank : 0;
  const int nT = a.T;
  const unsigned long long b0 = (unsigned long long)a.from * sizeof(T);
  ChunkIter it;
  it.nbytes = (a.to - a.from) * (long long)sizeof(T);
  it.ch = (long long)a.slot_bytes;
  it.first = blockIdx.x;
  it.stride = gridDim.x;
  const long long nchunks = (it.nbytes + it.ch - 1) / it.ch;
  const long long mine = nchunks > (long long)blockIdx.x ? (nchunks - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  uint32_t* ws_epoch = (uint32_t*)((char*)tab.sig[rank] + kSigWsEpochOff) + blockIdx.x;
  const uint32_t base = *(volatile uint32_t*)ws_epoch;
  const int kSlots = a.slots;
  const unsigned long long ring_off = a.stg_off + (unsigned long long)blockIdx.x * kSlots * a.slot_bytes;
  unsigned char* smA = smem;
  unsigned char* smC = smem + (size_t)kStagesA * kStageA;
  // bytes of one thread buffer per phase-A stage (multiple of 128)
  const int TB = (kStageA / nT) & ~127;
  if (threadIdx.x == 0) {
    c_done = 0;
    for (int i = 0; i < kWsMaxSlots; ++i) a_cnt[i] = 0;
    for (int s = 0; s < kStagesA; ++s) {
      mbar_init(&fullA[s], 1);
      mbar_init(&emptyA[s], kRW);
    }
    for (int s = 0; s < kStagesC; ++s) mbar_init(&fullC[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) {
    // ================= producer: TMA loads of the T input tiles, stage by stage
    if (lane == 0 && !(a.dbg & 4)) {
      const uint64_t pol = policy_evict_first();
      long long g = 0;
      for (long long k = 0; k < mine; ++k) {
        const long long off = it.off(k), len = it.len(k);
        for (long long t0 = 0; t0 < len; t0 += TB, ++g) {
          const int s = (int)(g % kStagesA);
          mbar_wait(&emptyA[s], (uint32_t)(((g / kStagesA) & 1) ^ 1));
          const uint32_t bytes = (uint32_t)min((long long)TB, len - t0);
          mbar_expect_tx(&fullA[s], bytes * (uint32_t)nT);
          for (int t = 0; t < nT; ++t)
            bulk_g2s(smA + (size_t)s * kStageA + (size_t)t * TB, (const char*)a.bufs[erow][t] + b0 + off + t0, bytes,
                     &fullA[s], pol);
        }
      }
    }
  } else if (warp <= kRW) {
    // ================= reducers: fold the T tiles of a stage (thread-id order) into the slot
    const int rtid = threadIdx.x - 32;
    const uint64_t pol_keep = policy_evict_last();
    long long g = 0;
    for (long long k = 0; k < mine; ++k) {
      const long long len = it.len(k);
      if (k >= kSlots) {
        if (lane == 0)
          while (c_done < (unsigned)(k - kSlots + 1)) {
          }
        __syncwarp();
      }
      char* slot = tab.heap[rank] + ring_off + (unsigned long long)(k % kSlots) * a.slot_bytes;
      if (a.dbg & 4) {
        // diagnostics: no HBM stream
      } else {
        for (long long t0 = 0; t0 < len; t0 += TB, ++g) {
          const int s = (int)(g % kStagesA);
          mbar_wait(&fullA[s], (uint32_t)((g / kStagesA) & 1));
          const int nvec = (int)(min((long long)TB, len - t0) / 16);
          const unsigned char* st = smA + (size_t)s * kStageA;
          for (int v = rtid; v < nvec; v += kRW * 32) {
            Vec16 acc = lds16(st + (size_t)v * 16);
#pragma unroll 8
            for (int t = 1; t < nT; ++t) acc = vec_reduce<T, OP>(acc, lds16(st + (size_t)t * TB + (size_t)v * 16));
            if (!(a.dbg & 32)) st_keep16(slot + t0 + (long long)v * 16, acc, pol_keep);
          }
          __syncwarp();
          if (lane == 0) mbar_arrive(&emptyA[s]);
        }
      }
      __syncwarp();
      if (lane == 0) {
        __threadfence_block();  // the warp's slot stores are ordered before the count
        atomicAdd(&a_cnt[k % kSlots], 1u);
        if (warp == 1) WS_TRACE(0, k);
      }
    }
  } else if (warp == kRW + 1) {
    // ================= signal warp: FA(k) to every peer once all reducer warps finished chunk k
    for (long long k = 0; k < mine; ++k) {
      if (lane < world) {
        while (*(volatile unsigned*)&a_cnt[k % kSlots] < (unsigned)kRW * (unsigned)(k / kSlots + 1)) {
        }
        __threadfence_block();
        if (lane == 0) WS_TRACE(1, k);
        st_release_sys((uint32_t*)((char*)tab.sig[lane] + kSigWsFaOff) + blockIdx.x * kMaxRanks + rank,
                       base + (uint32_t)k + 1u);
        if (lane == 0) WS_TRACE(2, k);
      }
    }
  } else if (warp <= kRW + 1 + kWB) {
    // ================= phase B: inter-GPU reduce-scatter + allgather of the slot over NVLink.
    // Team j (kWPC warps) owns the chunks k = j, j + kTeams, ...: kTeams chunks are in their NVLink
    // round trips at once, and the warps of a team split the rank's sub-slice.
    const int bw = warp - (kRW + 2);
    const int team = bw / kWPC, tw = bw % kWPC;
    for (long long k = team; k < mine; k += kTeams) {
      const int nv16 = (int)(it.len(k) / 16);
      if (lane < world && (!(a.dbg & 2) || lane == rank))
        wait_flag(tab, (uint32_t*)((char*)tab.sig[rank] + kSigWsFaOff) + blockIdx.x * kMaxRanks + lane,
                  base + (uint32_t)k + 1u);
      __syncwarp();
      if (tw == 0 && lane == 0) WS_TRACE(3, k);
      const int sl = (int)(k % kSlots);
      const unsigned long long so = ring_off + (unsigned long long)sl * a.slot_bytes;
      const int lo = (int)((long long)nv16 * rank / world), hi = (int)((long long)nv16 * (rank + 1) / world);
      constexpr int kBatch = NVLS ? 16 : 8;
      const bool loc = (a.dbg & 8) != 0;
      for (int v0 = lo + tw * 32 + lane; v0 < ((a.dbg & 1) ? lo : hi); v0 += 32 * kWPC * kBatch) {
        Vec16 acc[kBatch];
        if constexpr (NVLS) {
#pragma unroll
          for (int j = 0; j < kBatch; ++j) {
            const int v = v0 + j * 32 * kWPC;
            if (v < hi) acc[j] = Nvls<T, OP>::ld_reduce(tab.mc_heap + so + (unsigned long long)v * 16);
          }
#pragma unroll
          for (int j = 0; j < kBatch; ++j) {
            const int v = v0 + j * 32 * kWPC;
            if (v < hi) multimem_st(tab.mc_heap + so + (unsigned long long)v * 16, acc[j]);
          }
        } else {
          for (int q = 0; q < world; ++q) {  // ascending rank order => identical bits on every rank
            if (loc && q != rank) continue;
            Vec16 x[kBatch];
#pragma unroll
            for (int j = 0; j < kBatch; ++j) {
              const int v = v0 + j * 32 * kWPC;
              if (v < hi) x[j] = ld_coherent(tab.heap[q] + so + (unsigned long long)v * 16);
            }
#pragma unroll
            for (int j = 0; j < kBatch; ++j) acc[j] = (q == 0 || loc) ? x[j] : vec_reduce<T, OP>(acc[j], x[j]);
          }
          for (int q = 0; q < world; ++q) {
            if (loc && q != rank) continue;
#pragma unroll
            for (int j = 0; j < kBatch; ++j) {
              const int v = v0 + j * 32 * kWPC;
              if (v < hi) st_stream(tab.heap[q] + so + (unsigned long long)v * 16, acc[j]);
            }
          }
        }
      }
      named_bar_sync(4 + team, 32 * kWPC);  // the team's stores precede the release below
      if (tw == 0 && lane == 0) WS_TRACE(4, k);
      if (tw == 0 && lane < world)
        st_release_sys((uint32_t*)((char*)tab.sig[lane] + kSigWsFbOff) +
                           ((size_t)blockIdx.x * kWsMaxSlots + sl) * kMaxRanks + rank,
                       base + (uint32_t)k + 1u);
      if (tw == 0 && lane == 0) WS_TRACE(5, k);
    }
  } else {
    // ================= phase C: slot tile -> shared memory -> T bulk stores (one warp, lane 0 drives the TMA)
    const uint64_t pol = policy_evict_first();
    // the slot lines are rewritten by phase A a few chunks later: keep them in L2 (a dirty line that is
    // evicted costs an HBM write of dead data)
    const uint64_t pol_slot = (a.dbg & 64) ? pol : policy_evict_last();
    // load cursor (lk, lj) runs up to kStagesC - 2 tiles ahead of the store cursor (sk, sj)
    long long lk = 0, sk = 0, li = 0, si = 0;
    int lj = 0, sj = 0;
    bool lready = false;  // FB(lk) observed
    while (sk < mine) {
      // ---- prefetch
      while (lk < mine && li < si + (kStagesC - 1)) {
        if (!lready) {
          // FB(lk) from every rank; block only if there is nothing loaded to store meanwhile
          const uint32_t* fl = (uint32_t*)((char*)tab.sig[rank] + kSigWsFbOff) +
                               ((size_t)blockIdx.x * kWsMaxSlots + (int)(lk % kSlots)) * kMaxRanks + lane;
          const uint32_t want = base + (uint32_t)lk + 1u;
          const bool need = lane < world && (!(a.dbg & 2) || lane == rank);
          bool ok = true;
          if (li == si) {
            if (need) wait_flag(tab, fl, want);
          } else {
            if (need) ok = (int32_t)(ld_relaxed_sys(fl) - want) >= 0;
            ok = __all_sync(0xffffffffu, ok);
            if (ok && need) (void)ld_acquire_sys(fl);
          }
          __syncwarp();
          if (!ok) break;
          if (lane == 0) fence_proxy_async();  // peers' (generic-proxy) slot writes before the TMA reads
          if (lane == 0) WS_TRACE(6, lk);
          lready = true;
        }
        const long long len = it.len(lk);
        const long long t0 = (long long)lj * kTileC;
        const uint32_t bytes = (uint32_t)min((long long)kTileC, len - t0);
        const int s = (int)(li % kStagesC);
        if (lane == 0 && !(a.dbg & 4)) {
          if (li >= kStagesC) bulk_wait_read<1>();  // stores of tile li - kStagesC have left shared memory
          mbar_expect_tx(&fullC[s], bytes);
          bulk_g2s(smC + (size_t)s * kTileC,
                   tab.heap[rank] + ring_off + (unsigned long long)(lk % kSlots) * a.slot_bytes + t0, bytes, &fullC[s],
                   pol_slot);
        }
        ++li;
        if (t0 + kTileC >= len) {
          ++lk;
          lj = 0;
          lready = false;
        } else {
          ++lj;
        }
      }
      // ---- store tile si
      {
        const long long off = it.off(sk), len = it.len(sk);
        const long long t0 = (long long)sj * kTileC;
        const uint32_t bytes = (uint32_t)min((long long)kTileC, len - t0);
        const int s = (int)(si % kStagesC);
        const bool last = t0 + kTileC >= len;
        if (lane == 0) {
          if (!(a.dbg & 4)) {
            mbar_wait(&fullC[s], (uint32_t)((si / kStagesC) & 1));
            if (last) c_done = (unsigned)(sk + 1);  // the slot is in shared memory: phase A may reuse it
            if (last) WS_TRACE(7, sk);
            const unsigned long long o = b0 + off + t0;
            if (a.fanout) {
              for (int t = 0; t < nT; ++t) bulk_s2g((char*)a.bufs[erow][t] + o, smC + (size_t)s * kTileC, bytes, pol);
            } else {
              bulk_s2g((char*)a.bufs[erow][a.dst_index] + o, smC + (size_t)s * kTileC, bytes, pol);
            }
            bulk_commit();
          } else if (last) {
            c_done = (unsigned)(sk + 1);
            WS_TRACE(7, sk);
          }
        }
        ++si;
        if (last) {
          ++sk;
          sj = 0;
        } else {
          ++sj;
        }
      }
      __syncwarp();
    }
    if (lane == 0) bulk_wait_all();
  }
  __syncthreads();
  if (threadIdx.x == 0) *(volatile uint32_t*)ws_epoch = base + (uint32_t)mine;
}

template <typename T, int OP, bool NVLS>
void launch_tma_k(const ThreadArgs& args, int grid, cudaStream_t stream) {
  auto kfn = thread_fused_tma_kernel<T, OP, NVLS>;
  static bool configured[64] = {};  // per instantiation and device
  int dev = 0;
  MP4J_CUDA(cudaGetDevice(&dev));
  if (!configured[dev & 63]) {
    MP4J_CUDA(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kTmaSmem));
    configured[dev & 63] = true;
  }
  if (args.tab.emulate) {
    void* params[] = {(void*)&args};
    MP4J_CUDA(cudaLaunchCooperativeKernel((const void*)kfn, dim3(grid, args.tab.world), dim3(kTmaThreads), params,
                                          kTmaSmem, stream));
  } else {
    kfn<<<grid, kTmaThreads, kTmaSmem, stream>>>(args);
    MP4J_CUDA(cudaGetLastError());
  }
}

template <typename T, int OP>
void launch_tma_op(const ThreadArgs& args, bool nvls, int grid, cudaStream_t stream) {
  if constexpr (Nvls<T, OP>::supported) {
    if (nvls) return launch_tma_k<T, OP, true>(args, grid, stream);
  }
  launch_tma_k<T, OP, false>(args, grid, stream);
}

template <typename T>
void launch_tma_type(int op, const ThreadArgs& args, bool nvls, int grid, cudaStream_t stream) {
  switch (op) {
    case MP4J_SUM: return launch_tma_op<T, MP4J_SUM>(args, nvls, grid, stream);
    case MP4J_MAX: return launch_tma_op<T, MP4J_MAX>(args, nvls, grid, stream);
    case MP4J_MIN: return launch_tma_op<T, MP4J_MIN>(args, nvls, grid, stream);
  }
  throw Error("TMA fused thread kernel is built for SUM/MAX/MIN");
}

}  // namespace

bool thread_fused_tma_supported(int dtype, int op, int T) {
  if (T < 1 || T > 64) return false;
  if (!(op == MP4J_SUM || op == MP4J_MAX || op == MP4J_MIN)) return false;
  return dtype == MP4J_F64 || dtype == MP4J_F32 || dtype == MP4J_I64 || dtype == MP4J_I32;
}

// Needs: every buffer + range 16-byte aligned, slot_bytes a multiple of kTileC... (any multiple of 16 works;
// tiles are cut per chunk), grid * slots * slot_bytes of staging, one CTA per SM.
void launch_thread_fused_tma(int dtype, int op, bool nvls, const ThreadArgs& args, int grid, cudaStream_t stream) {
  switch (dtype) {
    case MP4J_F64: return launch_tma_type<double>(op, args, nvls, grid, stream);
    case MP4J_F32: return launch_tma_type<float>(op, args, nvls, grid, stream);
    case MP4J_I64: return launch_tma_type<int64_t>(op, args, nvls, grid, stream);
    case MP4J_I32: return launch_tma_type<int32_t>(op, args, nvls, grid, stream);
  }
  throw Error("TMA fused thread kernel is built for 4/8-byte types");
}

}  // namespace mp4j
