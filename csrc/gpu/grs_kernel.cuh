// The gather-reduce-scatter kernel template (see dense.cu for the collective instances). It lives in a
// header because it is compiled twice: by nvcc for the predefined (dtype, operator) pairs, and at run
// time by NVRTC for custom device operators (custom_op.cu), where Reduce<T, MP4J_CUSTOM_OP> is the
// user's expression.
#pragma once
#include "dense.cuh"

namespace mp4j {
namespace grs {

constexpr int kThreads = 512;

template <typename T, int OP>
__device__ __forceinline__ T reduce_or_first(T a, T b) {
  if constexpr (OP == MP4J_COPY) return a;
  else return Reduce<T, OP>::apply(a, b);
}

template <typename T, int OP>
__device__ __forceinline__ Vec16 vreduce(const Vec16& a, const Vec16& b) {
  if constexpr (OP == MP4J_COPY) return a;
  else return vec_reduce<T, OP>(a, b);
}

// 16 bytes through element-sized accesses (local pointer not 16-byte congruent with the heap)
template <typename T>
__device__ __forceinline__ Vec16 ld_elementwise(const void* p) {
  union {
    Vec16 v;
    T e[16 / sizeof(T)];
  } u;
#pragma unroll
  for (int i = 0; i < (int)(16 / sizeof(T)); ++i) u.e[i] = ((const T*)p)[i];
  return u.v;
}
template <typename T>
__device__ __forceinline__ void st_elementwise(void* p, const Vec16& v) {
  union {
    Vec16 v;
    T e[16 / sizeof(T)];
  } u;
  u.v = v;
#pragma unroll
  for (int i = 0; i < (int)(16 / sizeof(T)); ++i) ((T*)p)[i] = u.e[i];
}

template <typename T, int OP, bool NVLS, int W>
__global__ void __launch_bounds__(kThreads) grs_kernel(const __grid_constant__ GrsArgs a) {
  const PeerTable& tab = a.tab;
  const int rank = my_rank(tab);
  const GrsItem& it = a.item[tab.emulate ? rank : 0];
  constexpr int ES = sizeof(T);
  constexpr int EPV = 16 / ES;
  const int world = tab.world;
  const unsigned full = (world >= 32) ? 0xffffffffu : ((1u << world) - 1u);

  const unsigned long long src_off = a.src_off + (unsigned long long)it.adj, dst_off = a.dst_off + (unsigned long long)it.adj;
  BlockEpoch ep;
  ep.load(tab, rank);
  if (a.dead_rank_p1 != 0 && rank == a.dead_rank_p1 - 1) return;  // test hook: this rank never shows up

  // [lo, hi) = scalar head | 16-byte vector body | scalar tail, decided by the HEAP address only so
  // that every rank splits identically; a local pointer of different alignment is read/written
  // element-wise inside the same vector slots (vec_ok == 0).
  const long long n = it.hi > it.lo ? it.hi - it.lo : 0;
  long long head = 0, nvec = 0;
  if (n > 0) {
    const unsigned mis = (unsigned)((src_off + (unsigned long long)it.lo * ES) & 15ull);
    head = mis ? (16 - mis) / ES : 0;
    if (head > n) head = n;
    nvec = (n - head) / EPV;
  }
  const long long body_end = head + nvec * EPV;  // scalar set: [0,head) U [body_end, n)
  const long long nscalar = head + (n - body_end);
  const unsigned long long src_b0 = src_off + (unsigned long long)(it.lo + head) * ES;  // body start
  const unsigned long long dst_b0 = dst_off + (unsigned long long)(it.lo + head) * ES;
  const long long gstride = (long long)gridDim.x * blockDim.x;
  const long long gtid = (long long)blockIdx.x * blockDim.x + threadIdx.x;

  // ---- optional fused copy-in (one-shot): same thread->element mapping as the reduce loop below,
  // and every rank has the same item, so block b's barrier covers exactly what block b reads.
  if (it.src_local != nullptr) {
    const char* sl = (const char*)it.src_local;
    char* mine = tab.heap[rank];
    for (long long v = gtid; v < nvec; v += gstride) {
      const char* p = sl + (unsigned long long)(it.lo + head) * ES + v * 16;
      st_stream(mine + src_b0 + v * 16, a.vec_ok ? ld_stream(p) : ld_elementwise<T>(p));
    }
    if (blockIdx.x == 0) {
      for (long long s = threadIdx.x; s < nscalar; s += blockDim.x) {
        const long long i = it.lo + (s < head ? s : body_end + (s - head));
        *(T*)(mine + src_off + (unsigned long long)i * ES) = ((const T*)sl)[i];
      }
    }
  }

  // every rank tells every rank where it believes the array lives (published before the barrier, compared after)
  if (a.layout_key != 0 && a.entry_barrier && blockIdx.x == 0 && threadIdx.x < (unsigned)world)
    ((volatile unsigned long long*)((char*)tab.sig[threadIdx.x] + kSigLayoutKeyOff))[rank] = a.layout_key;
  if (a.entry_barrier && !block_barrier_peers(tab, rank, ep)) {
    ep.store();
    return;  // a peer is missing: store nothing; the host raises on the error word
  }
  if (a.layout_key != 0 && a.entry_barrier && blockIdx.x == 0 && threadIdx.x < (unsigned)world) {
    const unsigned long long theirs = ((volatile unsigned long long*)((char*)tab.sig[rank] + kSigLayoutKeyOff))[threadIdx.x];
    if (theirs != a.layout_key) atomicMax((unsigned*)tab.err, 2u);  // host: "ranks disagree on the array's layout"
  }

  const unsigned smask = it.src_mask & full, dmask = it.dst_mask & full;
  const bool multi_src = (smask & (smask - 1)) != 0;

  // ---- vector body. An NVLink round trip costs ~2-3 us and the asm-volatile accessors are never
  // reordered by the compiler, so memory-level parallelism is explicit: every thread keeps ~8
  // independent 16-byte requests in flight (kU vectors x sources). W > 0 is the compile-time world
  // size that lets the (vector, source) tile live in registers; W == 0 is the generic fallback.
  const bool single_src = smask != 0 && !multi_src;
  const int src0 = single_src ? (__ffs(smask) - 1) : 0;
  if (NVLS || single_src || (W > 0 && smask == full)) {
    constexpr int kU = (NVLS || W == 0) ? 8 : (W >= 8 ? 1 : 8 / W);
    constexpr int kW = W > 0 ? W : 1;
    for (long long v0 = gtid; v0 < nvec; v0 += gstride * 8) {
#pragma unroll
      for (int b = 0; b < 8 / kU; ++b) {
        const long long vb = v0 + (long long)b * kU * gstride;
        Vec16 acc[kU];
        if (single_src) {
#pragma unroll
          for (int j = 0; j < kU; ++j) {
            const long long v = vb + j * gstride;
            if (v < nvec) acc[j] = ld_coherent(tab.heap[src0] + src_b0 + v * 16);
          }
        } else if constexpr (NVLS) {
#pragma unroll
          for (int j = 0; j < kU; ++j) {
            const long long v = vb + j * gstride;
            if (v < nvec) acc[j] = Nvls<T, OP>::ld_reduce(tab.mc_heap + src_b0 + v * 16);
          }
        } else {
          Vec16 x[kU][kW];
#pragma unroll
          for (int j = 0; j < kU; ++j) {
            const long long v = vb + j * gstride;
#pragma unroll
            for (int q = 0; q < kW; ++q)
              if (v < nvec) x[j][q] = ld_coherent(tab.heap[q] + src_b0 + v * 16);
          }
#pragma unroll
          for (int j = 0; j < kU; ++j) {
            acc[j] = x[j][0];
#pragma unroll
            for (int q = 1; q < kW; ++q) acc[j] = vreduce<T, OP>(acc[j], x[j][q]);  // ascending rank order
          }
        }
        bool stored = false;
        if constexpr (NVLS) {
          if (dmask == full) {
#pragma unroll
            for (int j = 0; j < kU; ++j) {
              const long long v = vb + j * gstride;
              if (v < nvec) multimem_st(tab.mc_heap + dst_b0 + v * 16, acc[j]);
            }
            stored = true;
          }
        }
        if (!stored) {
          for (int q = 0; q < world; ++q) {
            if (!((dmask >> q) & 1u)) continue;
#pragma unroll
            for (int j = 0; j < kU; ++j) {
              const long long v = vb + j * gstride;
              if (v < nvec) st_stream(tab.heap[q] + dst_b0 + v * 16, acc[j]);
            }
          }
        }
        if (it.dst_local != nullptr) {
#pragma unroll
          for (int j = 0; j < kU; ++j) {
            const long long v = vb + j * gstride;
            if (v < nvec) {
              char* p = (char*)it.dst_local + (unsigned long long)(it.lo + head) * ES + v * 16;
              if (a.vec_ok) st_stream(p, acc[j]);
              else st_elementwise<T>(p, acc[j]);
            }
          }
        }
      }
    }
  } else {
    // generic: any source mask, runtime world, one vector at a time
    for (long long v = gtid; v < nvec; v += gstride) {
      Vec16 x[kMaxRanks];
#pragma unroll
      for (int q = 0; q < kMaxRanks; ++q)
        if (q < world && ((smask >> q) & 1u)) x[q] = ld_coherent(tab.heap[q] + src_b0 + v * 16);
      Vec16 acc;
      bool first = true;
#pragma unroll
      for (int q = 0; q < kMaxRanks; ++q) {
        if (q < world && ((smask >> q) & 1u)) {
          acc = first ? x[q] : vreduce<T, OP>(acc, x[q]);
          first = false;
        }
      }
#pragma unroll
      for (int q = 0; q < kMaxRanks; ++q)
        if (q < world && ((dmask >> q) & 1u)) st_stream(tab.heap[q] + dst_b0 + v * 16, acc);
      if (it.dst_local != nullptr) {
        char* p = (char*)it.dst_local + (unsigned long long)(it.lo + head) * ES + v * 16;
        if (a.vec_ok) st_stream(p, acc);
        else st_elementwise<T>(p, acc);
      }
    }
  }

  // ---- scalar head/tail (block 0)
  if (blockIdx.x == 0) {
    for (long long s = threadIdx.x; s < nscalar; s += blockDim.x) {
      const long long i = it.lo + (s < head ? s : body_end + (s - head));
      T acc = T();
      bool first = true;
      for (int q = 0; q < world; ++q) {
        if ((smask >> q) & 1u) {
          const T x = *(const volatile T*)(tab.heap[q] + src_off + (unsigned long long)i * ES);
          acc = first ? x : reduce_or_first<T, OP>(acc, x);
          first = false;
        }
      }
      for (int q = 0; q < world; ++q)
        if ((dmask >> q) & 1u) *(volatile T*)(tab.heap[q] + dst_off + (unsigned long long)i * ES) = acc;
      if (it.dst_local != nullptr) ((T*)it.dst_local)[i] = acc;
    }
  }

  if (a.exit_barrier) (void)block_barrier_peers(tab, rank, ep);
  ep.store();
}

}  // namespace grs
}  // namespace mp4j
