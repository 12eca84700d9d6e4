// Dense array collectives on the GPU: ONE kernel shape, "gather-reduce-scatter" (GRS).
//
//   for every element i of this rank's work item:
//       acc  = reduce over ranks q in src_mask (ascending)  of  heap[q][src_off + i]
//       heap[q][dst_off + i] = acc   for ranks q in dst_mask      (+ optional local output)
//
// with a per-block cross-GPU barrier on entry (peers' inputs are ready) and on exit (all pushes
// landed, nobody still reads my buffer). Instances:
//   allreduce two-shot : item = balanced slice r, src = all,    dst = all         (PCS:1733-1763)
//   allreduce one-shot : item = whole range,     src = all,    dst = local only  (PCS:1776, small)
//   reduce-scatter     : item = slice r,         src = all,    dst = {r}         (PCS:1174)
//   allgather          : item = slice r,         src = {r},    dst = all         (PCS:620)
//   broadcast          : item = balanced slice r, src = {root}, dst = all        (PCS:750)
//   gather             : item = slice r,         src = {r},    dst = {root}      (PCS:535)
//   scatter            : item = slice r,         src = {root}, dst = {r}         (PCS:900)
//   reduce             : item = balanced slice r, src = all,    dst = {root}      (PCS:1390)
// The reference runs these as p-1 serial ring steps or lg p tree rounds over TCP, each followed by a
// master barrier RPC; on NVSwitch every GPU reaches every peer at full bandwidth, so all peers are
// pulled/pushed concurrently inside one launch. With NVLS the reduce is a multimem.ld_reduce (the
// switch adds) and the scatter a multimem.st (the switch replicates).
//
// The operator reduction and the dtype are template parameters: the collective, the Operator and
// the element type are fused in the same loop (S1-S9 of SURVEY §2.6).
#include <algorithm>

#include "cuda_check.h"
#include "dense.cuh"
#include "launch.h"

namespace mp4j {

namespace {

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

  BlockEpoch ep;
  ep.load(tab, rank);

  // [lo, hi) = scalar head | 16-byte vector body | scalar tail, decided by the HEAP address only so
  // that every rank splits identically; a local pointer of different alignment is read/written
  // element-wise inside the same vector slots (vec_ok == 0).
  const long long n = it.hi > it.lo ? it.hi - it.lo : 0;
  long long head = 0, nvec = 0;
  if (n > 0) {
    const unsigned mis = (unsigned)((a.src_off + (unsigned long long)it.lo * ES) & 15ull);
    head = mis ? (16 - mis) / ES : 0;
    if (head > n) head = n;
    nvec = (n - head) / EPV;
  }
  const long long body_end = head + nvec * EPV;  // scalar set: [0,head) U [body_end, n)
  const long long nscalar = head + (n - body_end);
  const unsigned long long src_b0 = a.src_off + (unsigned long long)(it.lo + head) * ES;  // body start
  const unsigned long long dst_b0 = a.dst_off + (unsigned long long)(it.lo + head) * ES;
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
        *(T*)(mine + a.src_off + (unsigned long long)i * ES) = ((const T*)sl)[i];
      }
    }
  }

  if (a.entry_barrier) block_barrier_peers(tab, rank, ep);

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
          const T x = *(const volatile T*)(tab.heap[q] + a.src_off + (unsigned long long)i * ES);
          acc = first ? x : reduce_or_first<T, OP>(acc, x);
          first = false;
        }
      }
      for (int q = 0; q < world; ++q)
        if ((dmask >> q) & 1u) *(volatile T*)(tab.heap[q] + a.dst_off + (unsigned long long)i * ES) = acc;
      if (it.dst_local != nullptr) ((T*)it.dst_local)[i] = acc;
    }
  }

  if (a.exit_barrier) block_barrier_peers(tab, rank, ep);
  ep.store();
}

template <typename T, int OP, bool NVLS, int W = 0>
void launch_one(const GrsArgs& args, int grid, cudaStream_t stream) {
  auto kfn = grs_kernel<T, OP, NVLS, W>;
  if (args.tab.emulate) {
    void* params[] = {(void*)&args};
    MP4J_CUDA(cudaLaunchCooperativeKernel((const void*)kfn, dim3(grid, args.tab.world), dim3(kThreads), params, 0,
                                          stream));
  } else {
    kfn<<<grid, kThreads, 0, stream>>>(args);
    MP4J_CUDA(cudaGetLastError());
  }
}

template <typename T, int OP>
void launch_nvls_or_not(const GrsArgs& args, bool nvls, int grid, cudaStream_t stream) {
  if constexpr (Nvls<T, OP>::supported || OP == MP4J_COPY) {
    if (nvls) return launch_one<T, OP, true>(args, grid, stream);
  }
  // compile-time world size for the operators and types that carry the bulk of the traffic
  if constexpr (sizeof(T) >= 4 && (OP == MP4J_SUM || OP == MP4J_MAX || OP == MP4J_MIN)) {
    switch (args.tab.world) {
      case 2: return launch_one<T, OP, false, 2>(args, grid, stream);
      case 4: return launch_one<T, OP, false, 4>(args, grid, stream);
      case 8: return launch_one<T, OP, false, 8>(args, grid, stream);
    }
  }
  launch_one<T, OP, false>(args, grid, stream);
}

template <typename T, bool INTEGRAL, bool LOC>
void by_op(int op, const GrsArgs& args, bool nvls, int grid, cudaStream_t stream) {
  switch (op) {
    case MP4J_SUM: return launch_nvls_or_not<T, MP4J_SUM>(args, nvls, grid, stream);
    case MP4J_MAX: return launch_nvls_or_not<T, MP4J_MAX>(args, nvls, grid, stream);
    case MP4J_MIN: return launch_nvls_or_not<T, MP4J_MIN>(args, nvls, grid, stream);
    case MP4J_PROD: return launch_nvls_or_not<T, MP4J_PROD>(args, false, grid, stream);
  }
  if constexpr (INTEGRAL) {
    switch (op) {
      case MP4J_BAND: return launch_nvls_or_not<T, MP4J_BAND>(args, nvls, grid, stream);
      case MP4J_BOR: return launch_nvls_or_not<T, MP4J_BOR>(args, nvls, grid, stream);
      case MP4J_BXOR: return launch_nvls_or_not<T, MP4J_BXOR>(args, nvls, grid, stream);
    }
  }
  if constexpr (LOC) {
    switch (op) {
      case MP4J_MAXLOC: return launch_one<T, MP4J_MAXLOC, false>(args, grid, stream);
      case MP4J_MINLOC: return launch_one<T, MP4J_MINLOC, false>(args, grid, stream);
    }
  }
  throw Error("operator " + std::to_string(op) + " is not defined for this dtype on the GPU path");
}

}  // namespace

int grs_max_blocks_per_sm() {
  static int cached = -1;
  if (cached < 0) {
    int nb = 0;
    MP4J_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, grs_kernel<float, MP4J_SUM, false, 0>, kThreads, 0));
    cached = std::max(nb, 1);
  }
  return cached;
}

int grs_threads() { return kThreads; }

void launch_grs(int dtype, int op, bool nvls, const GrsArgs& args, int grid, cudaStream_t stream) {
  if (op == MP4J_COPY) {  // ranges are in bytes
    if (nvls) return launch_one<int8_t, MP4J_COPY, true>(args, grid, stream);
    return launch_one<int8_t, MP4J_COPY, false>(args, grid, stream);
  }
  switch (dtype) {
    case MP4J_F64: return by_op<double, false, true>(op, args, nvls, grid, stream);
    case MP4J_F32: return by_op<float, false, false>(op, args, nvls, grid, stream);
    case MP4J_I64: return by_op<int64_t, true, true>(op, args, nvls, grid, stream);
    case MP4J_I32: return by_op<int32_t, true, false>(op, args, nvls, grid, stream);
    case MP4J_I16: return by_op<int16_t, true, false>(op, args, false, grid, stream);
    case MP4J_I8: return by_op<int8_t, true, false>(op, args, false, grid, stream);
  }
  throw Error("dtype " + std::to_string(dtype) + " is not supported on the GPU path");
}

}  // namespace mp4j
