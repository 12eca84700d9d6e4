// ThreadCommSlave gather / allgather / broadcast / scatter / reduce-scatter / reduce on CUDA tensors:
// ONE launch per collective (TCS:448, 621, 802, 1010, 1174, 1396).
//
// The reference runs every one of them as "thread merge -> process collective -> thread copy" (three
// passes over the data plus lg T Exchanger rounds). Here a collective is a table of SEGMENTS of the
// common index space; each segment says where its value comes from and where it goes:
//
//   source       REDUCE : op over all T thread buffers of all ranks (thread-id order, then rank order)
//                COPY   : thread ts of rank qs
//   destination  thread td (or every thread) of rank qd (or of every rank)
//
//   reduceScatterArray  one REDUCE segment per (rank q, thread t) -> (q, t)      counts[q][t]
//   reduceArray         one REDUCE segment                        -> (root rank, root thread)
//   allgatherArray      one COPY segment per (q, t) from (q, t)   -> everybody
//   gatherArray         one COPY segment per (q, t) from (q, t)   -> (root rank, root thread)
//   scatterArray        one COPY segment per (q, t) from the root -> (q, t)
//   broadcastArray      one COPY segment from the root            -> everybody
//
// The kernel cuts the segments into windows of one staging slot (64 KiB); block b walks windows b, b + G,
// ... on every rank:
//   stage 1 (source ranks)       fold / copy the local thread buffers of the window; write the local
//                                destinations directly when the value is already final (COPY, or one
//                                rank), else put it into this block's slot (double buffered)
//   cross-GPU barrier            device-scope fence + relaxed flags: all cross-GPU traffic of this kernel
//                                is LOADS from the peers' slots, nothing remote is ever stored
//   stage 2 (destination ranks)  load the window from the source ranks' slots over NVLink (rank order for
//                                REDUCE), write the local destination thread buffers
// Only what a destination needs crosses NVLink, once. Unaligned segment ends (counts are arbitrary) are
// handled element-wise; the 16-byte aligned interior of every window moves as 16-byte vectors.
#include <stdlib.h>

#include <algorithm>
#include <vector>

#include "../core/partition.h"
#include "cuda_check.h"
#include "custom_op.h"
#include "dense.cuh"
#include "gpu_plane.h"
#include "launch.h"

namespace mp4j {

namespace {

constexpr int kSegThreads = 256;
constexpr int kMaxSegs = 128;
constexpr int kAll = -1;

struct SegItem {
  long long lo, hi;  // element range
  int src_rank, src_thread;  // kAll / kAll = REDUCE over everybody; else COPY from that thread
  int dst_rank, dst_thread;  // kAll = every rank / every thread
  long long first_window;    // index of the segment's first window in the launch's window list
};

struct SegArgs {
  PeerTable tab;
  int T;
  int nseg;
  int vec_ok;                    // all thread buffers share one alignment modulo 16
  int base_mis;                  // that alignment (bytes)
  unsigned long long stg_off;    // heap offset of the slots: [block][kSegSlots][slot_bytes]
  unsigned long long slot_bytes;
  long long nwindows;
  int lag, pad_;                 // stage 2 of a window runs `lag` iterations after its stage 1 (1 .. kSegSlots - 1)
  long long wstride;             // window visited at position i of the walk: (i * wstride) % nwindows (coprime)
  void* bufs[kMaxRanks][MP4J_MAX_THREADS];  // [rank][thread] in emulation, [0][thread] otherwise
  SegItem seg[kMaxSegs];
};

constexpr int kSegSlots = 4;  // staging slots per block: a source may run this many windows ahead of its readers

template <typename T, int OP>
__device__ __forceinline__ T red1(T a, T b) {
  if constexpr (OP == MP4J_COPY) return a;
  else return Reduce<T, OP>::apply(a, b);
}
template <typename T, int OP>
__device__ __forceinline__ Vec16 red16(const Vec16& a, const Vec16& b) {
  if constexpr (OP == MP4J_COPY) return a;
  else return vec_reduce<T, OP>(a, b);
}

// stage 2 of a REDUCE window for a compile-time world size: kB vectors x W source ranks = 8 loads in flight
// per thread (an NVLink round trip is ~3 us; one vector at a time would make a window 16 round trips long)
template <typename T, int OP, int W>
__device__ __forceinline__ void pull_reduce(const SegArgs& a, int erow, unsigned long long slot_rel, int td0, int td1,
                                            long long dst_rel, long long yva, long long nvec) {
  const PeerTable& tab = a.tab;
  constexpr int kB = 8 / W;
  #pragma unroll 1
  for (long long v0 = threadIdx.x; v0 < nvec; v0 += (long long)kSegThreads * kB) {
    Vec16 x[kB][W];
#pragma unroll
    for (int j = 0; j < kB; ++j) {
      const long long v = v0 + (long long)j * kSegThreads;
#pragma unroll
      for (int q = 0; q < W; ++q)
        if (v < nvec) x[j][q] = ld_coherent(tab.heap[q] + slot_rel + yva + v * 16);
    }
#pragma unroll
    for (int j = 0; j < kB; ++j) {
      const long long v = v0 + (long long)j * kSegThreads;
      if (v < nvec) {
        Vec16 acc = x[j][0];
#pragma unroll
        for (int q = 1; q < W; ++q) acc = red16<T, OP>(acc, x[j][q]);  // ascending rank order
        for (int t = td0; t < td1; ++t) st_stream((char*)a.bufs[erow][t] + dst_rel + yva + v * 16, acc);
      }
    }
  }
}

// stage 1 of a REDUCE window for a compile-time thread count: 2 vectors x NT buffers = 16 independent 16-byte
// loads in flight per thread, folded in thread-id order
template <typename T, int OP, int NT>
__device__ __forceinline__ void fold_reduce(const SegArgs& a, int erow, long long c, long long yva, long long nvec,
                                            char* slot, bool write_local, int td0, int td1) {
#pragma unroll 1
  for (long long v0 = threadIdx.x; v0 < nvec; v0 += (long long)kSegThreads * 2) {
    Vec16 x[2][NT];
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const long long v = v0 + (long long)j * kSegThreads;
#pragma unroll
      for (int t = 0; t < NT; ++t)
        if (v < nvec) x[j][t] = ld_stream((const char*)a.bufs[erow][t] + (yva - c) + v * 16);
    }
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const long long v = v0 + (long long)j * kSegThreads;
      if (v < nvec) {
        Vec16 acc = x[j][0];
#pragma unroll
        for (int t = 1; t < NT; ++t) acc = red16<T, OP>(acc, x[j][t]);
        const long long y = yva + v * 16;
        if (slot) st_stream(slot + y, acc);
        if (write_local)
          for (int t = td0; t < td1; ++t) st_stream((char*)a.bufs[erow][t] + (y - c), acc);
      }
    }
  }
}

// One window of the launch: which segment it belongs to and its byte range.
struct Win {
  int seg;
  bool reduce;
  int src_rank, dst_rank;
  long long win, ya, yb;
};

template <typename T, int OP>
__global__ void __launch_bounds__(kSegThreads, 2) thread_seg_kernel(const __grid_constant__ SegArgs a) {
  constexpr long long ES = sizeof(T);
  const PeerTable& tab = a.tab;
  const int rank = my_rank(tab);
  const int world = tab.world;
  const int erow = tab.emulate ? rank : 0;
  const int nT = a.T;
  const long long CH = (long long)a.slot_bytes;
  const long long c = a.base_mis;
  BlockEpoch ep;
  ep.load(tab, rank);
  const uint32_t base = ep.value;
  // flags. FA: "source rank src finished stage 1 of this block's k-th window" (in the reader's pad);
  //        FC: "reader rank finished pulling this block's k-th window" (in the source's pad). Values base + k + 1.
  auto fa = [&](int owner, int src) { return tab.sig[owner] + blockIdx.x * kMaxRanks + src; };
  auto fc = [&](int owner, int reader) {
    return (uint32_t*)((char*)tab.sig[owner] + kSigSegFcOff) + blockIdx.x * kMaxRanks + reader;
  };
  auto decode = [&](long long w) {
    Win o;
    int lo_s = 0, hi_s = a.nseg - 1;
    while (lo_s < hi_s) {
      const int mid = (lo_s + hi_s + 1) >> 1;
      if (a.seg[mid].first_window <= w) lo_s = mid;
      else hi_s = mid - 1;
    }
    const SegItem& sg = a.seg[lo_s];
    o.seg = lo_s;
    o.reduce = sg.src_rank == kAll;
    o.src_rank = sg.src_rank;
    o.dst_rank = sg.dst_rank;
    // virtual byte offsets y = element offset * ES + c: y % 16 == address % 16 in every thread buffer
    const long long ylo = sg.lo * ES + c, yhi = sg.hi * ES + c;
    o.win = ylo / CH + (w - sg.first_window);
    o.ya = max(ylo, o.win * CH);
    o.yb = min(yhi, (o.win + 1) * CH);
    return o;
  };
  // does rank d pull window `x` out of rank q's slot?
  auto pulls = [&](const Win& x, int d, int q) {
    const bool q_src = x.reduce || x.src_rank == q;
    const bool d_dst = x.dst_rank == kAll || x.dst_rank == d;
    return q_src && d_dst && d != q;
  };
  // waits (threads 0..world-1) until every reader of my slot for window index kk has pulled it
  auto wait_consumed = [&](long long kk) {
    bool ok = true;
    if (threadIdx.x < (unsigned)world) {
      const Win x = decode((((long long)blockIdx.x + kk * gridDim.x) * a.wstride) % a.nwindows);
      if (pulls(x, (int)threadIdx.x, rank)) ok = wait_flag(tab, fc(rank, (int)threadIdx.x), base + (uint32_t)kk + 1u);
    }
    return __syncthreads_and(ok ? 1 : 0) != 0;
  };

  // Software pipeline inside the block: iteration i runs stage 1 of this block's window i and THEN stage 2 of its
  // window i - D. Every window needs all its source ranks, so a destination that did "stage 1, stage 2" back to
  // back kept everybody waiting for its next stage 1 while it was pulling (4 GPUs: 1.42 ms for the 8 x 64 Mi
  // reduce-scatter); with the lag its stage 1 runs D windows ahead and the pull is off the other ranks' path. D must
  // leave slack in the slot ring (a source reuses a slot kSegSlots windows later and then needs the reader's "pulled"
  // flag): D = kSegSlots - 1 couples the ranks iteration by iteration again; D = 1 leaves three windows of slack.
  const long long K = a.nwindows > (long long)blockIdx.x ? (a.nwindows - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  const long long D = world > 1 ? a.lag : 0;
  bool alive = true;
  for (long long it = 0; it < K + D && alive; ++it)
  for (int ph = 0; ph < 2 && alive; ++ph) {
    const long long k = ph == 0 ? it : it - D;
    if (k < 0 || k >= K || (ph == 1 && world == 1)) continue;  // one rank: stage 1 wrote the results already
    const long long w = (long long)blockIdx.x + k * gridDim.x;
    // positions are spread over the whole window list: at any moment the blocks work on windows of MANY segments,
    // so every rank is pulling (and being pulled from) at the same time. Walking the list in order made the ranks
    // take turns - consecutive windows belong to one segment, i.e. to one destination rank.
    const Win x = decode((w * a.wstride) % a.nwindows);
    const SegItem& sg = a.seg[x.seg];
    const bool reduce = x.reduce;
    const bool is_src = reduce || sg.src_rank == rank;
    const bool is_dst = sg.dst_rank == kAll || sg.dst_rank == rank;
    const long long win = x.win, ya = x.ya, yb = x.yb;
    long long yva = (ya + 15) & ~15ll, yvb = yb & ~15ll;  // 16-byte aligned interior
    if (!a.vec_ok || yva >= yvb) yva = yvb = ya;          // all scalar
    const long long nvec = (yvb - yva) / 16;
    const long long nhead = (yva - ya) / ES, ntail = (yb - max(yvb, ya)) / ES;
    const unsigned long long slot_off =
        a.stg_off + ((unsigned long long)blockIdx.x * kSegSlots + (unsigned long long)(k % kSegSlots)) * a.slot_bytes;
    const unsigned long long slot_rel = slot_off - (unsigned long long)(win * CH);  // heap + slot_rel + y
    char* my_slot = tab.heap[rank] + slot_rel;
    const bool local_final = !reduce || world == 1;  // COPY, or one rank: the local value is the result
    const bool to_slot = world > 1 && is_src && (reduce || sg.dst_rank == kAll || sg.dst_rank != rank);
    const int td0 = sg.dst_thread == kAll ? 0 : sg.dst_thread, td1 = sg.dst_thread == kAll ? nT : sg.dst_thread + 1;
    auto buf = [&](int t, long long y) { return (char*)a.bufs[erow][t] + (y - c); };

    if (ph == 0) {
    // the slot of this window was last used kSegSlots windows ago: its readers must be done with it
    if (world > 1 && k >= kSegSlots && !wait_consumed(k - kSegSlots)) {
      alive = false;
      break;
    }

    // ================= stage 1: local fold / copy
    if (is_src) {
      const bool write_local = local_final && is_dst;
      if (reduce) {
        // the loads of up to 8 thread buffers are in flight together (the accessors are asm volatile, so a
        // load placed after a dependent reduce would wait for it), thread-id order
        #pragma unroll 1
        for (long long v = threadIdx.x; v < nvec; v += kSegThreads) {
          const long long y = yva + v * 16;
          Vec16 acc;
#pragma unroll 1
          for (int t0 = 0; t0 < nT; t0 += 8) {
            Vec16 xx[8];
#pragma unroll
            for (int j = 0; j < 8; ++j)
              if (t0 + j < nT) xx[j] = ld_stream(buf(t0 + j, y));
#pragma unroll
            for (int j = 0; j < 8; ++j)
              if (t0 + j < nT) acc = (t0 + j == 0) ? xx[j] : red16<T, OP>(acc, xx[j]);
          }
          if (to_slot) st_stream(my_slot + y, acc);
          if (write_local)
            for (int t = td0; t < td1; ++t) st_stream(buf(t, y), acc);
        }
      } else {
        #pragma unroll 1
        for (long long v0 = threadIdx.x; v0 < nvec; v0 += (long long)kSegThreads * 4) {
          Vec16 acc[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const long long v = v0 + (long long)j * kSegThreads;
            if (v < nvec) acc[j] = ld_stream(buf(sg.src_thread, yva + v * 16));
          }
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const long long v = v0 + (long long)j * kSegThreads;
            if (v < nvec) {
              const long long y = yva + v * 16;
              if (to_slot) st_stream(my_slot + y, acc[j]);
              if (write_local)
                for (int t = td0; t < td1; ++t)
                  if (t != sg.src_thread) st_stream(buf(t, y), acc[j]);
            }
          }
        }
      }
      #pragma unroll 1
      for (long long i = threadIdx.x; i < nhead + ntail; i += kSegThreads) {
        const long long y = i < nhead ? ya + i * ES : max(yvb, ya) + (i - nhead) * ES;
        T acc;
        if (reduce) {
          acc = *(const T*)buf(0, y);
          for (int t = 1; t < nT; ++t) acc = red1<T, OP>(acc, *(const T*)buf(t, y));
        } else {
          acc = *(const T*)buf(sg.src_thread, y);
        }
        if (to_slot) *(T*)(my_slot + y) = acc;
        if (write_local)
          for (int t = td0; t < td1; ++t)
            if (reduce || t != sg.src_thread) *(T*)buf(t, y) = acc;
      }
    }
    if (world == 1) continue;

    // ================= publish: the slot is in THIS GPU's memory, readers load it through this GPU's L2, so
    // a device-scope fence by every writer orders it before the (relaxed) flag to each reader
    if (to_slot) {
      __threadfence();
      __syncthreads();
      if (threadIdx.x < (unsigned)world && pulls(x, (int)threadIdx.x, rank))
        st_relaxed_sys(fa((int)threadIdx.x, rank), base + (uint32_t)k + 1u);
    }
    continue;
    }  // ph == 0

    // ================= stage 2 (ph == 1): pull from the source ranks' slots, write the local destinations
    if (is_dst && (reduce || sg.src_rank != rank)) {
      bool ok = true;
      if (threadIdx.x < (unsigned)world && pulls(x, rank, (int)threadIdx.x))
        ok = wait_flag(tab, fa(rank, (int)threadIdx.x), base + (uint32_t)k + 1u);
      if (!__syncthreads_and(ok ? 1 : 0)) {
        alive = false;
        break;
      }
      const int q0 = reduce ? 0 : sg.src_rank, q1 = reduce ? world : sg.src_rank + 1;
      if (reduce) {
        if (world == 2 || world == 4 || world == 8) {
          switch (world) {
            case 2: pull_reduce<T, OP, 2>(a, erow, slot_rel, td0, td1, -c, yva, nvec); break;
            case 4: pull_reduce<T, OP, 4>(a, erow, slot_rel, td0, td1, -c, yva, nvec); break;
            default: pull_reduce<T, OP, 8>(a, erow, slot_rel, td0, td1, -c, yva, nvec); break;
          }
        } else {
          #pragma unroll 1
          for (long long v = threadIdx.x; v < nvec; v += kSegThreads) {
            const long long y = yva + v * 16;
            Vec16 acc = ld_coherent(tab.heap[0] + slot_rel + y);
            for (int q = 1; q < world; ++q) acc = red16<T, OP>(acc, ld_coherent(tab.heap[q] + slot_rel + y));
            for (int t = td0; t < td1; ++t) st_stream(buf(t, y), acc);
          }
        }
      } else {
        #pragma unroll 1
        for (long long v0 = threadIdx.x; v0 < nvec; v0 += (long long)kSegThreads * 8) {
          Vec16 acc[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const long long v = v0 + (long long)j * kSegThreads;
            if (v < nvec) acc[j] = ld_coherent(tab.heap[q0] + slot_rel + yva + v * 16);
          }
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const long long v = v0 + (long long)j * kSegThreads;
            if (v < nvec)
              for (int t = td0; t < td1; ++t) st_stream(buf(t, yva + v * 16), acc[j]);
          }
        }
      }
      #pragma unroll 1
      for (long long i = threadIdx.x; i < nhead + ntail; i += kSegThreads) {
        const long long y = i < nhead ? ya + i * ES : max(yvb, ya) + (i - nhead) * ES;
        T acc = *(const volatile T*)(tab.heap[q0] + slot_rel + y);
        for (int q = q0 + 1; q < q1; ++q) acc = red1<T, OP>(acc, *(const volatile T*)(tab.heap[q] + slot_rel + y));
        for (int t = td0; t < td1; ++t) *(T*)buf(t, y) = acc;
      }
      // every load of the window has returned (its value was stored above): tell the sources
      __syncthreads();
      if (threadIdx.x < (unsigned)world && pulls(x, rank, (int)threadIdx.x))
        st_relaxed_sys(fc((int)threadIdx.x, rank), base + (uint32_t)k + 1u);
    }
  }
  // nobody leaves (and lets a later launch reuse the slots) while a reader may still pull from them
  if (world > 1 && alive)
    for (long long kk = K > kSegSlots ? K - kSegSlots : 0; kk < K; ++kk)
      if (!wait_consumed(kk)) break;
  ep.value = base + (uint32_t)K;  // K windows on every rank: the per-block counters stay in step
  ep.store();
}

template <typename T, int OP>
void launch_seg(const SegArgs& args, int grid, cudaStream_t stream) {
  auto kfn = thread_seg_kernel<T, OP>;
  if (args.tab.emulate && args.tab.world > 1) {
    void* params[] = {(void*)&args};
    MP4J_CUDA(cudaLaunchCooperativeKernel((const void*)kfn, dim3(grid, args.tab.world), dim3(kSegThreads), params, 0,
                                          stream));
  } else {
    kfn<<<grid, kSegThreads, 0, stream>>>(args);
    MP4J_CUDA(cudaGetLastError());
  }
}

template <typename T, bool INTEGRAL, bool LOC>
void seg_by_op(int op, const SegArgs& a, int grid, cudaStream_t s) {
  switch (op) {
    case MP4J_SUM: return launch_seg<T, MP4J_SUM>(a, grid, s);
    case MP4J_MAX: return launch_seg<T, MP4J_MAX>(a, grid, s);
    case MP4J_MIN: return launch_seg<T, MP4J_MIN>(a, grid, s);
    case MP4J_PROD: return launch_seg<T, MP4J_PROD>(a, grid, s);
  }
  if constexpr (INTEGRAL) {
    switch (op) {
      case MP4J_BAND: return launch_seg<T, MP4J_BAND>(a, grid, s);
      case MP4J_BOR: return launch_seg<T, MP4J_BOR>(a, grid, s);
      case MP4J_BXOR: return launch_seg<T, MP4J_BXOR>(a, grid, s);
    }
  }
  if constexpr (LOC) {
    switch (op) {
      case MP4J_MAXLOC: return launch_seg<T, MP4J_MAXLOC>(a, grid, s);
      case MP4J_MINLOC: return launch_seg<T, MP4J_MINLOC>(a, grid, s);
    }
  }
  throw Error("operator " + std::to_string(op) + " is not defined for this dtype on the GPU path");
}

int seg_blocks_per_sm() {
  static int cached = -1;
  if (cached < 0) {
    int nb = 0;
    MP4J_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, thread_seg_kernel<float, MP4J_SUM>, kSegThreads, 0));
    cached = std::max(1, std::min(nb, 4));
  }
  return cached;
}

}  // namespace

// segs: nseg x {lo, hi, src_rank, src_thread, dst_rank, dst_thread} (int64 each; -1 = all).
void GpuPlane::thread_segments(void* const* bufs, int T, int dtype, int op, const int64_t* segs, int nseg,
                               const char* name, cudaStream_t s) {
  entry_check();
  if (T < 1 || T > MP4J_MAX_THREADS) throw Error("thread count out of range");
  if (nseg < 0 || nseg > kMaxSegs) throw Error("too many segments for one launch (max " + std::to_string(kMaxSegs) + ")");
  const size_t es = dtype_size(dtype);
  if (es == 0) throw Error("unknown dtype");
  const int p = world();
  const int L = local_ranks();
  // Relay for the two collectives with ONE root (broadcastArray, reduceArray). As a single launch everybody pulls
  // from the root's slots (broadcast) or the root pulls from everybody's (reduce): (p - 1) x the array through one
  // NVLink port. Two launches move (p - 1) / p of the array through every port instead: scatter + allgather for a
  // broadcast, reduce-scatter + gather for a reduce (the decomposition the reference uses at process level,
  // PCS:750-770 and 1390-1416). Piece q rests in thread 0 of rank q between the launches (in the root thread on the
  // root rank); a destination rank writes its own buffers, so stream order is all the two launches need.
  if (opt_.seg_relay && nseg == 1 && p > 2 && !in_relay_) {
    const int64_t* g = segs;
    const bool bcast = g[2] != kAll && g[3] != kAll && g[4] == kAll && g[5] == kAll;
    const bool red = g[2] == kAll && g[3] == kAll && g[4] != kAll && g[5] != kAll;
    const int64_t len = g[1] - g[0];
    if ((bcast || red) && g[0] >= 0 && len > 0 && (size_t)len * es >= opt_.seg_relay_min_bytes) {
      const int64_t we = (int64_t)(std::max<size_t>(opt_.thread_slot_bytes, 64 << 10) / 16 * 16 / es);
      const int64_t per = ((len + p - 1) / p + we - 1) / we * we;  // whole windows per piece
      const int r0 = (int)(bcast ? g[2] : g[4]), t0 = (int)(bcast ? g[3] : g[5]);
      if (r0 < 0 || r0 >= p || t0 < 0 || t0 >= T) throw Error("segment endpoint out of range");
      std::vector<int64_t> s1, s2;
      for (int q = 0; q < p; ++q) {
        const int64_t lo = std::min<int64_t>(g[0] + q * per, g[1]), hi = std::min<int64_t>(lo + per, g[1]);
        if (hi <= lo) continue;
        const int64_t hold = q == r0 ? t0 : 0;
        if (bcast) {
          if (q != r0) s1.insert(s1.end(), {lo, hi, r0, t0, q, 0});
          s2.insert(s2.end(), {lo, hi, q, hold, kAll, kAll});
        } else {
          s1.insert(s1.end(), {lo, hi, kAll, kAll, q, hold});
          if (q != r0) s2.insert(s2.end(), {lo, hi, q, 0, r0, t0});
        }
      }
      in_relay_ = true;
      try {
        thread_segments(bufs, T, dtype, op, s1.data(), (int)(s1.size() / 6), name, s);
        thread_segments(bufs, T, dtype, op, s2.data(), (int)(s2.size() / 6), name, s);
      } catch (...) {
        in_relay_ = false;
        throw;
      }
      in_relay_ = false;
      last_.algo = "fused-segments-relay";
      last_.bytes = (size_t)len * es;
      return;
    }
  }
  bool any_reduce = false;
  SegArgs a{};
  a.tab = table();
  a.T = T;
  for (int i = 0; i < L; ++i)
    for (int t = 0; t < T; ++t) a.bufs[i][t] = bufs[(size_t)i * T + t];
  const long long scale = 1;
  a.vec_ok = 1;
  a.base_mis = (int)((uintptr_t)bufs[0] & 15);
  for (int i = 0; i < L * T; ++i)
    if ((int)((uintptr_t)bufs[i] & 15) != a.base_mis) a.vec_ok = 0;
  if (!a.vec_ok) a.base_mis = 0;
  const size_t CH = std::max<size_t>(opt_.thread_slot_bytes, 64 << 10) / 16 * 16;
  a.slot_bytes = CH;
  a.stg_off = thread_off_;
  long long nwin = 0;
  int n = 0;
  for (int i = 0; i < nseg; ++i) {
    const int64_t* g = segs + (size_t)i * 6;
    if (g[0] < 0 || g[1] < g[0]) throw Error("segment range is not legal");
    if (g[1] == g[0]) continue;
    SegItem& it = a.seg[n++];
    it.lo = g[0] * scale;
    it.hi = g[1] * scale;
    it.src_rank = (int)g[2];
    it.src_thread = (int)g[3];
    it.dst_rank = (int)g[4];
    it.dst_thread = (int)g[5];
    if ((it.src_rank == kAll) != (it.src_thread == kAll)) throw Error("segment source must be (all, all) or one thread");
    if (it.src_rank >= p || it.dst_rank >= p || it.src_thread >= T || it.dst_thread >= T) throw Error("segment endpoint out of range");
    if (it.src_rank == kAll) any_reduce = true;
    it.first_window = nwin;
    const long long ylo = it.lo * (long long)es + a.base_mis, yhi = it.hi * (long long)es + a.base_mis;
    nwin += (yhi - 1) / (long long)CH - ylo / (long long)CH + 1;
  }
  a.nseg = n;
  a.nwindows = nwin;
  if (n == 0) return;
  static const int seg_lag = getenv("MP4J_SEG_LAG") ? atoi(getenv("MP4J_SEG_LAG")) : 1;
  a.lag = std::max(1, std::min(seg_lag, kSegSlots - 1));
  // stride of the walk: about one segment's worth of windows, coprime with the number of windows
  a.wstride = 1;
  if (nwin > 2 * n) {
    long long st = nwin / n + 1;
    auto gcd = [](long long x, long long y) { while (y) { long long t = x % y; x = y; y = t; } return x; };
    while (gcd(st, nwin) != 1) ++st;
    a.wstride = st % nwin;
  }
  if (any_reduce && (!plane_op_ok(dtype, op) || op == MP4J_COPY || is_custom_op(op)))
    throw Error("operator is not defined for this dtype on the fused thread collectives");
  long long cap = (long long)sms_ * seg_blocks_per_sm();
  if (emulated()) cap = std::max(1, sms_ / p);
  cap = std::min<long long>(cap, (long long)(thread_bytes_ / (kSegSlots * CH)));
  cap = std::min<long long>(cap, kMaxBlocks);
  if (opt_.max_blocks > 0) cap = std::min<long long>(cap, opt_.max_blocks);
  const int grid = (int)std::max<long long>(1, std::min(cap, nwin));
  last_ = CollStat{name, "fused-segments", 0, grid, false};
  for (int i = 0; i < n; ++i) last_.bytes += (size_t)(a.seg[i].hi - a.seg[i].lo) * es;
  const int kop = any_reduce ? op : MP4J_COPY;
  if (kop == MP4J_COPY) {
    // COPY segments never combine values: move them as raw elements of the same width
    switch (es) {
      case 8: return launch_seg<int64_t, MP4J_COPY>(a, grid, s);
      case 4: return launch_seg<int32_t, MP4J_COPY>(a, grid, s);
      case 2: return launch_seg<int16_t, MP4J_COPY>(a, grid, s);
      default: return launch_seg<int8_t, MP4J_COPY>(a, grid, s);
    }
  }
  switch (dtype) {
    case MP4J_F64: return seg_by_op<double, false, true>(kop, a, grid, s);
    case MP4J_F32: return seg_by_op<float, false, false>(kop, a, grid, s);
    case MP4J_I64: return seg_by_op<int64_t, true, true>(kop, a, grid, s);
    case MP4J_I32: return seg_by_op<int32_t, true, false>(kop, a, grid, s);
    case MP4J_I16: return seg_by_op<int16_t, true, false>(kop, a, grid, s);
    case MP4J_I8: return seg_by_op<int8_t, true, false>(kop, a, grid, s);
  }
  throw Error("dtype " + std::to_string(dtype) + " is not supported by the fused thread collectives");
}

}  // namespace mp4j
