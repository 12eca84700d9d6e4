// Device-side building blocks of the GPU data plane: the peer table every kernel receives,
// system-scope flag primitives, the per-block cross-GPU barrier, 16-byte vector helpers and
// multimem (NVLS) wrappers.
//
// Replaces, on the device, what the reference does with a TCP connect per message plus a master
// barrier RPC per phase (PCS:391-425, 503, 731, 1153, 1367): kernels load/store peer HBM directly
// over NVLink and synchronise through signal pads in symmetric memory.
#pragma once
#ifndef __CUDACC_RTC__
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/mp4j_types.h"
#endif
#include "../core/ops.h"

namespace mp4j {

constexpr int kMaxRanks = MP4J_MAX_RANKS;
constexpr int kMaxBlocks = 1024;                      // flag rows per signal pad
constexpr size_t kSigFlagsBytes = kMaxBlocks * kMaxRanks * sizeof(uint32_t);
constexpr size_t kSigEpochOff = kSigFlagsBytes;       // per-block epoch counters (local use only)
// warp-specialised fused thread kernel: its own chunk counters and two flag arrays. FA (phase A of
// chunk k done): [block][source rank], monotonic in k. FB (phase B done): [block][ring slot][source
// rank] - phase B of different chunks runs concurrently in different warps and may finish out of order.
constexpr int kWsMaxBlocks = 256;
constexpr int kWsMaxSlots = 8;
constexpr size_t kSigWsEpochOff = kSigEpochOff + 4096;
constexpr size_t kSigWsFaOff = kSigWsEpochOff + 4096;
constexpr size_t kSigWsFbOff = kSigWsFaOff + kWsMaxBlocks * kMaxRanks * sizeof(uint32_t);
constexpr size_t kSigWsEnd = kSigWsFbOff + (size_t)kWsMaxBlocks * kWsMaxSlots * kMaxRanks * sizeof(uint32_t);
// FC (pull formulation, thread_fused_pull.cu): "rank r has pulled what it needs of chunk k from this slot"
constexpr size_t kSigWsFcOff = kSigWsEnd;
constexpr size_t kSigWsFcEnd = kSigWsFcOff + (size_t)kWsMaxBlocks * kWsMaxSlots * kMaxRanks * sizeof(uint32_t);
constexpr size_t kSigPadBytes = 256 * 1024;           // everything above, zero-initialised once
// diagnostics (MP4J_WS_DBG bit 128): per-chunk stage timestamps of block 0, kWsTraceChunks x 8 x u64
constexpr size_t kSigWsTraceOff = kSigWsFcEnd;
constexpr int kWsTraceChunks = 64;
// thread_seg.cu: "reader r has pulled this block's k-th window" [block][reader rank]
constexpr size_t kSigSegFcOff = kSigWsTraceOff + kWsTraceChunks * 8 * 8;
// dense collectives: every rank's view of where the array lives (symmetric offset, or "staged window j"), one
// u64 per source rank, compared after the entry barrier - ranks that disagree must not touch each other's memory
constexpr size_t kSigLayoutKeyOff = kSigSegFcOff + kSigFlagsBytes;
static_assert(kSigLayoutKeyOff + kMaxRanks * 8 <= kSigPadBytes, "signal pad layout overflows");

struct PeerTable {
  int world;
  int rank;                       // real mode; emulation derives the rank from blockIdx.y
  int emulate;                    // 1: all "ranks" live on this GPU, one grid row per rank
  int poll_ns;                    // sleep between flag polls; 0 = tight spin (MP4J_POLL_NS)
  char* heap[kMaxRanks];          // every rank's symmetric heap as mapped HERE
  uint32_t* sig[kMaxRanks];       // every rank's signal pad as mapped HERE
  char* mc_heap;                  // multicast (NVLS) mapping of the heaps, or nullptr
  uint32_t* err;                  // host-visible error word: 1 = peer wait timed out
  unsigned long long timeout_ns;  // device-side watchdog for flag waits
};

__device__ __forceinline__ int my_rank(const PeerTable& t) { return t.emulate ? (int)blockIdx.y : t.rank; }

// ---- system-scope flag access
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void st_relaxed_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// Waits until *flag reaches `want` (wrap-safe). Returns false on watchdog timeout, after raising
// the host-visible error word: a dead peer must not hang every other rank's kernel.
// Polling uses RELAXED loads with a short sleep and ONE acquire fence once the flag has arrived:
// an acquire at system scope invalidates the SM's L1 (CCTL.IVALL) and drains the LSU, so spinning
// on ld.acquire would tax the streaming warps that share the SM with the poller.
__device__ __forceinline__ uint32_t ld_relaxed_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

__device__ __forceinline__ bool wait_flag(const PeerTable& t, const uint32_t* flag, uint32_t want) {
  if ((int32_t)(ld_relaxed_sys(flag) - want) >= 0) {
    (void)ld_acquire_sys(flag);  // acquire once, on the already-arrived flag (cheaper than a full fence)
    return true;
  }
  unsigned long long t0 = globaltimer_ns();
  for (unsigned spins = 1;; ++spins) {
    if (t.poll_ns > 0) __nanosleep((unsigned)t.poll_ns);
    if ((int32_t)(ld_relaxed_sys(flag) - want) >= 0) {
      (void)ld_acquire_sys(flag);  // acquire once, on the already-arrived flag (cheaper than a full fence)
      return true;
    }
    if ((spins & 1023u) == 0) {
      // somebody else already gave up: the launch is lost, leave quickly instead of timing out again
      if (*(volatile uint32_t*)t.err != 0u) return false;
      if (globaltimer_ns() - t0 > t.timeout_ns) {
        atomicExch((unsigned*)t.err, 1u);
        return false;
      }
    }
  }
}

// Per-block epoch counter: persists across launches in the local pad so flags never need a reset.
struct BlockEpoch {
  uint32_t* slot;
  uint32_t value;
  __device__ __forceinline__ void load(const PeerTable& t, int rank) {
    slot = (uint32_t*)((char*)t.sig[rank] + kSigEpochOff) + blockIdx.x;
    value = *(volatile uint32_t*)slot;
  }
  __device__ __forceinline__ void store() {
    if (threadIdx.x == 0) *(volatile uint32_t*)slot = value;
  }
};

// Cross-GPU barrier between the same-index blocks of all ranks. Data written by any thread of the
// block before the call is visible to the peers' block after their call returns (bar.sync orders
// the CTA's writes before the releasing thread; st.release.sys is cumulative).
// Returns false if a peer did not arrive within the watchdog time: the caller must then leave without
// storing anything (the host finds the error word and declares the plane unusable).
__device__ __forceinline__ bool block_barrier_peers(const PeerTable& t, int rank, BlockEpoch& ep) {
  ep.value += 1;
  __syncthreads();
  bool ok = true;
  if (threadIdx.x < (unsigned)t.world) {
    const int q = threadIdx.x;
    st_release_sys(t.sig[q] + blockIdx.x * kMaxRanks + rank, ep.value);
    ok = wait_flag(t, t.sig[rank] + blockIdx.x * kMaxRanks + q, ep.value);
  }
  return __syncthreads_and(ok ? 1 : 0) != 0;
}

// ---- 16-byte vector moves (streaming: touched once, keep them out of L1)
struct __align__(16) Vec16 {
  uint32_t x, y, z, w;
};
__device__ __forceinline__ Vec16 ld_stream(const void* p) {
  Vec16 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p));
  return v;
}
// plain (coherent) load: for data another GPU or an earlier phase of this kernel produced
__device__ __forceinline__ Vec16 ld_coherent(const void* p) {
  Vec16 v;
  asm volatile("ld.global.relaxed.sys.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ void st_stream(void* p, const Vec16& v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y),
               "r"(v.z), "r"(v.w)
               : "memory");
}

// ---- 32-byte vector moves: LDG.E.256 / STG.E.256 are new on sm_100 (PTX .v8.b32, needs an L2
// eviction-priority qualifier); they are what lifts a many-stream reduce to the copy roofline.
struct __align__(32) Vec32 {
  Vec16 lo, hi;
};
__device__ __forceinline__ Vec32 ld_stream32(const void* p) {  // read-once input
  Vec32 v;
  asm volatile("ld.global.nc.L1::no_allocate.L2::evict_first.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(v.lo.x), "=r"(v.lo.y), "=r"(v.lo.z), "=r"(v.lo.w), "=r"(v.hi.x), "=r"(v.hi.y), "=r"(v.hi.z),
                 "=r"(v.hi.w)
               : "l"(p));
  return v;
}
__device__ __forceinline__ Vec32 ld_coherent32(const void* p) {  // data produced during this kernel
  Vec32 v;
  asm volatile("ld.global.relaxed.sys.L2::evict_first.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(v.lo.x), "=r"(v.lo.y), "=r"(v.lo.z), "=r"(v.lo.w), "=r"(v.hi.x), "=r"(v.hi.y), "=r"(v.hi.z),
                 "=r"(v.hi.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ void st_stream32(void* p, const Vec32& v) {  // write-once output
  asm volatile("st.global.L1::no_allocate.L2::evict_first.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p),
               "r"(v.lo.x), "r"(v.lo.y), "r"(v.lo.z), "r"(v.lo.w), "r"(v.hi.x), "r"(v.hi.y), "r"(v.hi.z), "r"(v.hi.w)
               : "memory");
}
__device__ __forceinline__ void st_keep32(void* p, const Vec32& v) {  // staging: will be re-read from L2
  asm volatile("st.global.L2::evict_last.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "r"(v.lo.x),
               "r"(v.lo.y), "r"(v.lo.z), "r"(v.lo.w), "r"(v.hi.x), "r"(v.hi.y), "r"(v.hi.z), "r"(v.hi.w)
               : "memory");
}

template <typename T, int OP>
__device__ __forceinline__ Vec16 vec_reduce(const Vec16& a, const Vec16& b) {
  constexpr int N = 16 / sizeof(T);
  union U {
    Vec16 v;
    T e[N];
    __device__ U() {}
  } ua, ub;
  ua.v = a;
  ub.v = b;
#pragma unroll
  for (int i = 0; i < N; ++i) ua.e[i] = Reduce<T, OP>::apply(ua.e[i], ub.e[i]);
  return ua.v;
}

// ---- NVLS: in-switch reduction / broadcast through the multicast mapping
template <typename T, int OP>
struct Nvls {
  static constexpr bool supported = false;
  __device__ static Vec16 ld_reduce(const void*) { return Vec16{}; }
};
template <>
struct Nvls<float, MP4J_SUM> {
  static constexpr bool supported = true;
  __device__ __forceinline__ static Vec16 ld_reduce(const void* mc) {
    Vec16 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(mc)
                 : "memory");
    return v;
  }
};
template <>
struct Nvls<double, MP4J_SUM> {
  static constexpr bool supported = true;
  __device__ __forceinline__ static Vec16 ld_reduce(const void* mc) {
    union {
      Vec16 v;
      double d[2];
    } u;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.f64 %0, [%1];" : "=d"(u.d[0]) : "l"(mc) : "memory");
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.f64 %0, [%1];"
                 : "=d"(u.d[1])
                 : "l"((const char*)mc + 8)
                 : "memory");
    return u.v;
  }
};
#define MP4J_NVLS_INT(T, OPC, PTXOP, PTXTY, W)                                                            \
  template <>                                                                                            \
  struct Nvls<T, OPC> {                                                                                  \
    static constexpr bool supported = true;                                                              \
    __device__ __forceinline__ static Vec16 ld_reduce(const void* mc) {                                  \
      union {                                                                                            \
        Vec16 v;                                                                                         \
        T e[16 / sizeof(T)];                                                                             \
      } u;                                                                                               \
      _Pragma("unroll") for (int i = 0; i < (int)(16 / sizeof(T)); ++i) {                                \
        asm volatile("multimem.ld_reduce.relaxed.sys.global." PTXOP "." PTXTY " %0, [%1];"               \
                     : "=" W(u.e[i])                                                                     \
                     : "l"((const char*)mc + i * sizeof(T))                                              \
                     : "memory");                                                                        \
      }                                                                                                  \
      return u.v;                                                                                        \
    }                                                                                                    \
  };
MP4J_NVLS_INT(int32_t, MP4J_SUM, "add", "u32", "r")
MP4J_NVLS_INT(int32_t, MP4J_MAX, "max", "s32", "r")
MP4J_NVLS_INT(int32_t, MP4J_MIN, "min", "s32", "r")
MP4J_NVLS_INT(int32_t, MP4J_BAND, "and", "b32", "r")
MP4J_NVLS_INT(int32_t, MP4J_BOR, "or", "b32", "r")
MP4J_NVLS_INT(int32_t, MP4J_BXOR, "xor", "b32", "r")
MP4J_NVLS_INT(int64_t, MP4J_SUM, "add", "u64", "l")
MP4J_NVLS_INT(int64_t, MP4J_MAX, "max", "s64", "l")
MP4J_NVLS_INT(int64_t, MP4J_MIN, "min", "s64", "l")
MP4J_NVLS_INT(int64_t, MP4J_BAND, "and", "b64", "l")
MP4J_NVLS_INT(int64_t, MP4J_BOR, "or", "b64", "l")
MP4J_NVLS_INT(int64_t, MP4J_BXOR, "xor", "b64", "l")
#undef MP4J_NVLS_INT

__device__ __forceinline__ void multimem_st(void* mc, const Vec16& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "r"(v.x), "r"(v.y),
               "r"(v.z), "r"(v.w)
               : "memory");
}

inline bool nvls_supported(int dtype, int op) {
  if (op == MP4J_COPY) return true;  // multimem.st broadcast
  switch (dtype) {
    case MP4J_F32: case MP4J_F64: return op == MP4J_SUM;
    // bf16 / fp16 SUM have multimem forms too (half_types.cuh) but stay on the P2P kernels until the
    // in-switch path has been verified on a multi-GPU box (it rounds differently: fp32 accumulation)
    case MP4J_I32: case MP4J_I64:
      return op == MP4J_SUM || op == MP4J_MAX || op == MP4J_MIN || op == MP4J_BAND || op == MP4J_BOR ||
             op == MP4J_BXOR;
  }
  return false;
}

}  // namespace mp4j
