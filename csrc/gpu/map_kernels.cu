// Sparse (map container) collectives on the GPU: allreduceMap / reduceScatterMap / allgatherMap
// (PCS:2053-2081, 1299-1327, 666-692) as hash-partition -> peer exchange -> hash-merge.
//
// The reference hash-partitions a HashMap<String,T> into p HashMaps on the CPU (PCS:2059-2072),
// ring-reduce-scatters them with a get/put + operator call per (String, value) pair (DOP:225-257),
// ring-allgathers the p reduced maps and merges them (PCS:2074-2081). Here:
//   K10 map_partition_push : every (key, value) pair is written STRAIGHT into the owner GPU's
//        receive region over NVLink (no local bucket copy, no serialisation); per-destination
//        cursors are warp-aggregated atomics; the last block publishes the counts and meets the
//        peers at a flag barrier.
//   K11 map_table          : one cooperative kernel: the owner clears as much of its open-addressing
//        table as the received counts need, inserts what it received (atomicCAS on the key, atomic
//        operator on the value), compacts the table into dense (key, value) arrays, publishes the
//        owned count to the peers and meets them.
//   map_allgather_push     : every rank pushes its owned pairs to the right offset of every peer's
//        output (offsets = prefix sums of the published counts, read on the device: no host sync)
//        and writes the 4-word info block the caller reads after synchronising.
// A collective is these three launches and nothing else on the stream (no memsets). The separate
// clear / merge / compact kernels remain for the local (one GPU, no peers) reduce-by-key.
// Float sums use atomicAdd: the order of additions, hence the last bits, is not deterministic
// (the reference's order is arrival order, PCS:1329-1373 - also unspecified).
#include <cooperative_groups.h>

#include <algorithm>
#include <mutex>
#include <string>

#include "cuda_check.h"
#include "launch.h"
#include "map_kernels.cuh"

namespace mp4j {

namespace cg = cooperative_groups;

namespace {

constexpr int kMapThreads = 256;
constexpr int kCtrlRow = kMaxBlocks - 1;  // flag row used by the single "last block" of a map kernel

__device__ __forceinline__ unsigned long long splitmix64(unsigned long long x) {
  x += 0x9e3779b97f4a7c15ull;
  x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ull;
  x = (x ^ (x >> 27)) * 0x94d049bb133111ebull;
  return x ^ (x >> 31);
}

__device__ __forceinline__ int owner_of(long long key, int p, int partition) {
  if (partition == MP4J_MAP_PART_JAVA32) {
    int h = (int)(unsigned)(unsigned long long)key;  // low 32 bits = String.hashCode()
    int m = h % p;                                   // C remainder has Java's sign rule
    return m < 0 ? m + p : m;
  }
  if (partition == MP4J_MAP_PART_TAGGED) {
    // the caller chose the destination: it sits in the top 4 bits of the key (reduceScatterMap / scatter-like
    // uses where the reference addresses rank i with list element i, PCS:1299-1327)
    const int d = (int)((unsigned long long)key >> 60);
    return d < p ? d : p - 1;
  }
  return (int)(splitmix64((unsigned long long)key) % (unsigned)p);
}

// cross-rank barrier executed by ONE block per rank (whichever finished last) on a fixed flag row
__device__ __forceinline__ void ctrl_barrier_peers(const PeerTable& t, int rank) {
  uint32_t* eslot = (uint32_t*)((char*)t.sig[rank] + kSigEpochOff) + kCtrlRow;
  __shared__ uint32_t epoch_s;
  if (threadIdx.x == 0) epoch_s = *(volatile uint32_t*)eslot + 1;
  __syncthreads();
  const uint32_t e = epoch_s;
  if (threadIdx.x < (unsigned)t.world) {
    const int q = threadIdx.x;
    st_release_sys(t.sig[q] + kCtrlRow * kMaxRanks + rank, e);
    wait_flag(t, t.sig[rank] + kCtrlRow * kMaxRanks + q, e);
  }
  __syncthreads();
  if (threadIdx.x == 0) *(volatile uint32_t*)eslot = e;
}

// returns true for exactly one block per rank row: the one that finished last
__device__ __forceinline__ bool last_block_of_rank(unsigned long long* ticket) {
  __shared__ bool last;
  __threadfence_system();  // this block's peer stores are performed before it takes a ticket
  __syncthreads();
  if (threadIdx.x == 0) last = (atomicAdd(ticket, 1ull) == (unsigned long long)gridDim.x - 1);
  __syncthreads();
  if (last) __threadfence_system();
  return last;
}

// ---------------------------------------------------------------------------------------- K10
template <typename V>
__global__ void __launch_bounds__(kMapThreads) map_partition_push_kernel(const __grid_constant__ MapExchangeArgs a) {
  const PeerTable& tab = a.tab;
  const int rank = my_rank(tab);
  const int p = tab.world;
  const int row = tab.emulate ? rank : 0;
  const long long n = a.n[row];
  const long long* keys = a.keys[row];
  const V* vals = (const V*)a.vals[row];
  unsigned long long* cursor = (unsigned long long*)(tab.heap[rank] + a.cursor_off);  // [p] + overflow + ticket
  const long long stride = (long long)gridDim.x * blockDim.x;
  const unsigned lane = threadIdx.x & 31;
  for (long long i0 = (long long)blockIdx.x * blockDim.x; i0 < n; i0 += stride) {
    const long long i = i0 + threadIdx.x;
    const bool active = i < n;
    long long key = active ? keys[i] : 0;
    V val = active ? vals[i] : V();
    const int d = active ? owner_of(key, p, a.partition) : -1;
    // warp-aggregated slot claim: one atomic per distinct destination per warp
    const unsigned peers = __match_any_sync(0xffffffffu, d);
    const int leader = __ffs(peers) - 1;
    unsigned long long base = 0;
    if (active && (int)lane == leader) base = atomicAdd(&cursor[d], (unsigned long long)__popc(peers));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (active) {
      const unsigned long long slot = base + __popc(peers & ((1u << lane) - 1u));
      if (slot < (unsigned long long)a.cap) {
        // region `rank` of destination d
        long long* rk = (long long*)(tab.heap[d] + a.rkeys_off) + (unsigned long long)rank * a.cap + slot;
        V* rv = (V*)(tab.heap[d] + a.rvals_off) + (unsigned long long)rank * a.cap + slot;
        *rk = key;
        *rv = val;
      } else {
        cursor[p] = 1;  // overflow: reported by the host after the call
      }
    }
  }
  if (last_block_of_rank(&cursor[p + 1])) {
    if (threadIdx.x < (unsigned)p) {
      const int q = threadIdx.x;
      const unsigned long long c = min(*(volatile unsigned long long*)&cursor[q], (unsigned long long)a.cap);
      *((volatile unsigned long long*)(tab.heap[q] + a.rcnt_off) + rank) = c;
    }
    __syncthreads();
    ctrl_barrier_peers(tab, rank);
    if (threadIdx.x <= (unsigned)p + 1 && threadIdx.x != (unsigned)p) cursor[threadIdx.x] = 0;  // keep overflow flag
  }
}

// ---------------------------------------------------------------------------------------- K11
template <typename V, int OP>
__device__ __forceinline__ void atomic_reduce(V* addr, V v) {
  if constexpr (OP == MP4J_SUM) {
    if constexpr (sizeof(V) == 8 && !is_fp<V>::value) atomicAdd((unsigned long long*)addr, (unsigned long long)v);
    else atomicAdd(addr, v);
  } else if constexpr (!is_fp<V>::value) {
    if constexpr (OP == MP4J_MAX) atomicMax(addr, v);
    else atomicMin(addr, v);
  } else {
    // float/double max/min: CAS loop with the reference's Math.max/min semantics (ops.h)
    using U = typename std::conditional<sizeof(V) == 8, unsigned long long, unsigned int>::type;
    U* ua = (U*)addr;
    U old = *ua, assumed;
    do {
      assumed = old;
      const V cur = bit_cast<V>(assumed);
      const V nv = Reduce<V, OP>::apply(cur, v);
      if (bit_cast<U>(nv) == assumed) break;
      old = atomicCAS(ua, assumed, bit_cast<U>(nv));
    } while (old != assumed);
  }
}

template <typename V, int OP>
__device__ __forceinline__ V identity() {
  if constexpr (OP == MP4J_SUM) return V(0);
  else if constexpr (is_fp<V>::value) {
    // -inf / +inf
    if constexpr (sizeof(V) == 8) return bit_cast<V>(OP == MP4J_MAX ? 0xfff0000000000000ull : 0x7ff0000000000000ull);
    else return bit_cast<V>(OP == MP4J_MAX ? 0xff800000u : 0x7f800000u);
  } else if constexpr (sizeof(V) == 8) {
    return OP == MP4J_MAX ? (V)0x8000000000000000ull : (V)0x7fffffffffffffffull;
  } else {
    return OP == MP4J_MAX ? (V)0x80000000u : (V)0x7fffffffu;
  }
}

template <typename V, int OP>
__global__ void __launch_bounds__(kMapThreads) map_clear_kernel(MapTable t) {
  const unsigned long long size = t.mask + 1;
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < size;
       i += (unsigned long long)gridDim.x * blockDim.x) {
    t.keys[i] = kMapEmpty;
    ((V*)t.vals)[i] = identity<V, OP>();
  }
}

// inserts `cnt` pairs; table must have room (size >= 2 * pairs)
template <typename V, int OP>
__device__ __forceinline__ void table_insert(const MapTable& t, long long key, V val) {
  // INT64_MIN marks a free slot: a pair with that key cannot be stored (it would be folded into whatever
  // free slot it probed first and leak into the next key placed there). Such pairs are ignored - documented
  // in mp4j_gpu.h; the String-key bridge never produces the value.
  if (key == kMapEmpty) return;
  unsigned long long h = splitmix64((unsigned long long)key ^ 0x5851f42d4c957f2dull) & t.mask;
  for (;;) {
    const long long prev = (long long)atomicCAS((unsigned long long*)&t.keys[h], (unsigned long long)kMapEmpty,
                                                (unsigned long long)key);
    if (prev == kMapEmpty || prev == key) {
      atomic_reduce<V, OP>((V*)t.vals + h, val);
      return;
    }
    h = (h + 1) & t.mask;
  }
}

// merges a plain local array of pairs (thread-level map reduce / local pre-aggregation)
template <typename V, int OP>
__global__ void __launch_bounds__(kMapThreads) map_merge_local_kernel(MapTable t, const long long* keys,
                                                                      const V* vals, long long n) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    table_insert<V, OP>(t, keys[i], vals[i]);
}

// table -> dense arrays; count accumulates in *count (zeroed by the caller)
template <typename V>
__global__ void __launch_bounds__(kMapThreads) map_compact_kernel(MapTable t, long long* out_keys, V* out_vals,
                                                                  unsigned long long* count,
                                                                  unsigned long long out_cap) {
  const unsigned long long size = t.mask + 1;
  const unsigned lane = threadIdx.x & 31;
  for (unsigned long long i0 = (unsigned long long)blockIdx.x * blockDim.x; i0 < size;
       i0 += (unsigned long long)gridDim.x * blockDim.x) {
    const unsigned long long i = i0 + threadIdx.x;
    const bool live = i < size && t.keys[i] != kMapEmpty;
    const unsigned m = __ballot_sync(0xffffffffu, live);
    if (m == 0) continue;
    unsigned long long base = 0;
    const int leader = __ffs(m) - 1;
    if ((int)lane == leader) base = atomicAdd(count, (unsigned long long)__popc(m));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (live) {
      const unsigned long long o = base + __popc(m & ((1u << lane) - 1u));
      if (o < out_cap) {
        out_keys[o] = t.keys[i];
        out_vals[o] = ((const V*)t.vals)[i];
      }
    }
  }
}

// Reduce-by-key of a plain local array in ONE cooperative launch (thread-level map merge, local pre-aggregation
// of a sparse gradient): clear the table, insert, compact. The three kernels above remain for the allocation-free
// entry point, which must not depend on cooperative launch limits.
template <typename V, int OP>
__global__ void __launch_bounds__(kMapThreads) map_local_kernel(MapTable t, const long long* keys, const V* vals,
                                                                long long n, long long* out_keys, V* out_vals,
                                                                unsigned long long* count,
                                                                unsigned long long out_cap) {
  cg::grid_group grid = cg::this_grid();
  const unsigned long long size = t.mask + 1;
  const unsigned long long tid = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
  const unsigned long long nthr = (unsigned long long)gridDim.x * blockDim.x;
  for (unsigned long long i = tid; i < size; i += nthr) {
    t.keys[i] = kMapEmpty;
    ((V*)t.vals)[i] = identity<V, OP>();
  }
  if (tid == 0) *count = 0;
  grid.sync();
  for (unsigned long long i = tid; i < (unsigned long long)n; i += nthr) table_insert<V, OP>(t, keys[i], vals[i]);
  grid.sync();
  const unsigned lane = threadIdx.x & 31;
  for (unsigned long long i0 = (unsigned long long)blockIdx.x * blockDim.x; i0 < size; i0 += nthr) {
    const unsigned long long i = i0 + threadIdx.x;
    const bool live = i < size && t.keys[i] != kMapEmpty;
    const unsigned m = __ballot_sync(0xffffffffu, live);
    if (m == 0) continue;
    unsigned long long base = 0;
    const int leader = __ffs(m) - 1;
    if ((int)lane == leader) base = atomicAdd(count, (unsigned long long)__popc(m));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (live) {
      const unsigned long long o = base + __popc(m & ((1u << lane) - 1u));
      if (o < out_cap) {
        out_keys[o] = t.keys[i];
        out_vals[o] = ((const V*)t.vals)[i];
      }
    }
  }
}

// ---------------------------------------------------------------------------------------- K11f
// One cooperative kernel for everything between the two exchanges of a map collective: clear the part
// of the table this call needs, merge the p receive regions, compact the table into dense arrays,
// publish the owned count to every peer and meet the peers. The table is allocated for the worst
// case (every pair of every rank lands here: 2 * p * cap slots, 512 MiB at 8 GPUs x 2M pairs), but
// the part that is cleared, probed and scanned is sized on the device from the counts that actually
// arrived (2x the received pairs, rounded up to a power of two): with balanced keys that is 1/p of it.
struct MapTableArgs {
  PeerTable tab;
  unsigned long long rkeys_off, rvals_off, rcnt_off, ocnt_off;
  long long cap;
  unsigned long long own_cap;
  MapTable t[kMaxRanks];  // indexed by rank in emulation, [0] otherwise (like every per-rank array here)
  long long* own_keys[kMaxRanks];
  void* own_vals[kMaxRanks];
  unsigned long long* own_count[kMaxRanks];
};

template <typename V, int OP>
__global__ void __launch_bounds__(kMapThreads) map_table_kernel(const __grid_constant__ MapTableArgs a) {
  cg::grid_group grid = cg::this_grid();
  const PeerTable& tab = a.tab;
  const int rank = my_rank(tab);
  const int p = tab.world;
  const int row = tab.emulate ? rank : 0;
  const unsigned long long* cnt = (const unsigned long long*)(tab.heap[rank] + a.rcnt_off);
  const long long* rk = (const long long*)(tab.heap[rank] + a.rkeys_off);
  const V* rv = (const V*)(tab.heap[rank] + a.rvals_off);
  unsigned long long total = 0;
  for (int s = 0; s < p; ++s) total += cnt[s];
  unsigned long long size = 1024;
  while (size < 2 * total) size <<= 1;
  if (size > a.t[row].mask + 1) size = a.t[row].mask + 1;
  const MapTable t{a.t[row].keys, a.t[row].vals, size - 1};
  const unsigned long long tid = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
  const unsigned long long nthr = (unsigned long long)gridDim.x * blockDim.x;
  unsigned long long* count = a.own_count[row];
  for (unsigned long long i = tid; i < size; i += nthr) {
    t.keys[i] = kMapEmpty;
    ((V*)t.vals)[i] = identity<V, OP>();
  }
  if (tid == 0) *count = 0;
  grid.sync();
  for (int s = 0; s < p; ++s) {
    const unsigned long long c = cnt[s];
    for (unsigned long long i = tid; i < c; i += nthr)
      table_insert<V, OP>(t, rk[(unsigned long long)s * a.cap + i], rv[(unsigned long long)s * a.cap + i]);
  }
  grid.sync();
  {
    const unsigned lane = threadIdx.x & 31;
    long long* out_keys = a.own_keys[row];
    V* out_vals = (V*)a.own_vals[row];
    for (unsigned long long i0 = (unsigned long long)blockIdx.x * blockDim.x; i0 < size; i0 += nthr) {
      const unsigned long long i = i0 + threadIdx.x;
      const bool live = i < size && t.keys[i] != kMapEmpty;
      const unsigned m = __ballot_sync(0xffffffffu, live);
      if (m == 0) continue;
      unsigned long long base = 0;
      const int leader = __ffs(m) - 1;
      if ((int)lane == leader) base = atomicAdd(count, (unsigned long long)__popc(m));
      base = __shfl_sync(0xffffffffu, base, leader);
      if (live) {
        const unsigned long long o = base + __popc(m & ((1u << lane) - 1u));
        if (o < a.own_cap) {
          out_keys[o] = t.keys[i];
          out_vals[o] = ((const V*)t.vals)[i];
        }
      }
    }
  }
  grid.sync();
  if (blockIdx.x == 0) {
    if (threadIdx.x < (unsigned)p)
      *((volatile unsigned long long*)(tab.heap[threadIdx.x] + a.ocnt_off) + rank) = *(volatile unsigned long long*)count;
    __syncthreads();
    // after this barrier: every peer's owned count is here, and no peer still reads its receive regions
    ctrl_barrier_peers(tab, rank);
  }
}

// publishes this rank's owned count to every peer, meets the peers, then pushes the owned pairs to
// offset sum(counts[q < rank]) of every peer's output arrays; total goes to *total_out.
struct MapGatherArgs {
  PeerTable tab;
  unsigned long long ocnt_off;   // heap: uint64 [p] owned counts (slot r written by rank r)
  unsigned long long okeys_off;  // heap: int64 [out_cap]
  unsigned long long ovals_off;  // heap: V [out_cap]
  unsigned long long out_cap;
  const long long* own_keys[kMaxRanks];
  const void* own_vals[kMaxRanks];
  const unsigned long long* own_count[kMaxRanks];
  // what the caller reads after synchronising the stream (pinned, mapped): pairs in the result, receive overflow,
  // output overflow, pairs this rank owned. Written by block 0 of the push kernel when info[row] is set.
  unsigned long long* info[kMaxRanks];
  unsigned long long overflow_off;  // heap: receive-overflow flag of the partition step (0 = none in this mode)
  int use_mc;                       // push through the multicast mapping when the plane has one
};

__global__ void __launch_bounds__(kMapThreads) map_publish_count_kernel(const __grid_constant__ MapGatherArgs a) {
  const PeerTable& tab = a.tab;
  const int rank = my_rank(tab);
  const int row = tab.emulate ? rank : 0;
  if (threadIdx.x < (unsigned)tab.world)
    *((volatile unsigned long long*)(tab.heap[threadIdx.x] + a.ocnt_off) + rank) = *a.own_count[row];
  __syncthreads();
  ctrl_barrier_peers(tab, rank);
}

// src[0, n) -> element offset `off` of the array at heap offset `base_off` of EVERY rank
template <typename E>
__device__ __forceinline__ void mc_broadcast(const PeerTable& tab, unsigned long long base_off, unsigned long long off,
                                             const E* src, unsigned long long n, unsigned long long tid,
                                             unsigned long long nthr) {
  constexpr unsigned long long kPer = 16 / sizeof(E);
  const unsigned long long head = min(n, (kPer - (off % kPer)) % kPer);  // elements before the first aligned vector
  const unsigned long long nvec = (n - head) / kPer;
  const unsigned long long tail0 = head + nvec * kPer;
  for (unsigned long long v = tid; v < nvec; v += nthr) {
    const E* s = src + head + v * kPer;
    Vec16 x;
    if constexpr (sizeof(E) == 8) {
      const unsigned long long a0 = ((const unsigned long long*)s)[0], a1 = ((const unsigned long long*)s)[1];
      x.x = (uint32_t)a0, x.y = (uint32_t)(a0 >> 32), x.z = (uint32_t)a1, x.w = (uint32_t)(a1 >> 32);
    } else {
      const uint32_t* u = (const uint32_t*)s;
      x.x = u[0], x.y = u[1], x.z = u[2], x.w = u[3];
    }
    multimem_st(tab.mc_heap + base_off + (off + head + v * kPer) * sizeof(E), x);
  }
  // ragged ends: at most 2 * (kPer - 1) elements
  const unsigned long long nrag = head + (n - tail0);
  for (unsigned long long j = tid; j < nrag; j += nthr) {
    const unsigned long long i = j < head ? j : tail0 + (j - head);
    const E e = src[i];
    for (int q = 0; q < tab.world; ++q) ((E*)(tab.heap[q] + base_off))[off + i] = e;
  }
}

template <typename V>
__global__ void __launch_bounds__(kMapThreads) map_allgather_push_kernel(const __grid_constant__ MapGatherArgs a) {
  const PeerTable& tab = a.tab;
  const int rank = my_rank(tab);
  const int p = tab.world;
  const int row = tab.emulate ? rank : 0;
  const volatile unsigned long long* oc = (const volatile unsigned long long*)(tab.heap[rank] + a.ocnt_off);
  unsigned long long off = 0;
  for (int q = 0; q < rank; ++q) off += oc[q];
  const unsigned long long mine = oc[rank];
  const long long* sk = a.own_keys[row];
  const V* sv = (const V*)a.own_vals[row];
  BlockEpoch ep;
  ep.load(tab, rank);
  if (blockIdx.x == 0 && threadIdx.x == 0 && a.info[row]) {
    unsigned long long total = 0;
    for (int q = 0; q < p; ++q) total += oc[q];
    unsigned long long* info = a.info[row];
    info[0] = total < a.out_cap ? total : a.out_cap;
    info[1] = 0;
    if (a.overflow_off) {  // read and re-arm: the flag is not sticky across calls
      volatile unsigned long long* f = (volatile unsigned long long*)(tab.heap[rank] + a.overflow_off);
      info[1] = *f;
      *f = 0;
    }
    info[2] = total > a.out_cap ? 1 : 0;
    info[3] = *a.own_count[row];
  }
  const unsigned long long tid = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
  const unsigned long long nthr = (unsigned long long)gridDim.x * blockDim.x;
  const unsigned long long end = off + mine < a.out_cap ? mine : (off < a.out_cap ? a.out_cap - off : 0);  // pairs to push
  if (tab.mc_heap && a.use_mc) {
    // NVLS: one multimem.st of 16 bytes lands in every rank's output (the switch replicates it), so this rank
    // sends its owned pairs once instead of p times. 16-byte stores need a 16-byte aligned destination: the
    // few elements before the first aligned one and after the last full vector go out as per-peer stores.
    mc_broadcast<long long>(tab, a.okeys_off, off, sk, end, tid, nthr);
    mc_broadcast<V>(tab, a.ovals_off, off, sv, end, tid, nthr);
  } else {
    for (unsigned long long i = tid; i < end; i += nthr) {
      const long long k = sk[i];
      const V v = sv[i];
      for (int q = 0; q < p; ++q) {
        ((long long*)(tab.heap[q] + a.okeys_off))[off + i] = k;
        ((V*)(tab.heap[q] + a.ovals_off))[off + i] = v;
      }
    }
  }
  if (!block_barrier_peers(tab, rank, ep)) return;  // all pushes landed everywhere before anyone reads its output
  ep.store();
}

// ---------------------------------------------------------------------------------------- dispatch
template <typename V>
void launch_partition(const MapExchangeArgs& a, int grid, cudaStream_t s) {
  auto kfn = map_partition_push_kernel<V>;
  if (a.tab.emulate) {
    void* params[] = {(void*)&a};
    MP4J_CUDA(cudaLaunchCooperativeKernel((const void*)kfn, dim3(grid, a.tab.world), dim3(kMapThreads), params, 0, s));
  } else {
    kfn<<<grid, kMapThreads, 0, s>>>(a);
    MP4J_CUDA(cudaGetLastError());
  }
}

template <typename V>
void launch_gather(const MapGatherArgs& a, bool publish, bool push, int grid, cudaStream_t s) {
  if (a.tab.emulate) {
    void* params[] = {(void*)&a};
    if (publish)
      MP4J_CUDA(cudaLaunchCooperativeKernel((const void*)map_publish_count_kernel, dim3(1, a.tab.world),
                                            dim3(kMapThreads), params, 0, s));
    if (push)
      MP4J_CUDA(cudaLaunchCooperativeKernel((const void*)map_allgather_push_kernel<V>, dim3(grid, a.tab.world),
                                            dim3(kMapThreads), params, 0, s));
  } else {
    if (publish) map_publish_count_kernel<<<1, kMapThreads, 0, s>>>(a);
    if (push) map_allgather_push_kernel<V><<<grid, kMapThreads, 0, s>>>(a);
    MP4J_CUDA(cudaGetLastError());
  }
}

template <typename F>
void dispatch_v(int dtype, F&& f) {
  switch (dtype) {
    case MP4J_F32: return f(float());
    case MP4J_F64: return f(double());
    case MP4J_I32: return f(int32_t());
    case MP4J_I64: return f((long long)0);
  }
  throw Error("GPU maps support float/double/int/long values");
}
template <typename F>
void dispatch_op(int op, F&& f) {
  switch (op) {
    case MP4J_SUM: return f(std::integral_constant<int, MP4J_SUM>());
    case MP4J_MAX: return f(std::integral_constant<int, MP4J_MAX>());
    case MP4J_MIN: return f(std::integral_constant<int, MP4J_MIN>());
  }
  throw Error("GPU maps support SUM/MAX/MIN");
}

}  // namespace

// ---- type/op erased entry points used by the host side (map_plane.cu)
int map_threads() { return kMapThreads; }

void map_launch_partition(int dtype, const MapExchangeArgs& a, int grid, cudaStream_t s) {
  dispatch_v(dtype, [&](auto v) { launch_partition<decltype(v)>(a, grid, s); });
}

void map_launch_clear(int dtype, int op, const MapTable& t, int grid, cudaStream_t s) {
  dispatch_v(dtype, [&](auto v) {
    dispatch_op(op, [&](auto o) {
      map_clear_kernel<decltype(v), decltype(o)::value><<<grid, kMapThreads, 0, s>>>(t);
    });
  });
  MP4J_CUDA(cudaGetLastError());
}

void map_launch_local(int dtype, int op, const MapTable& t, const long long* keys, const void* vals, long long n,
                      long long* out_keys, void* out_vals, unsigned long long* count, unsigned long long out_cap,
                      int sms, cudaStream_t s) {
  dispatch_v(dtype, [&](auto v) {
    using V = decltype(v);
    dispatch_op(op, [&](auto o) {
      auto kfn = map_local_kernel<V, decltype(o)::value>;
      int per_sm = 0;
      MP4J_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kfn, kMapThreads, 0));
      const unsigned long long work = std::max<unsigned long long>(t.mask + 1, (unsigned long long)n);
      const int grid = (int)std::max<unsigned long long>(
          1, std::min<unsigned long long>((unsigned long long)std::min(per_sm, 4) * sms, (work + kMapThreads - 1) / kMapThreads));
      MapTable tt = t;
      const V* vv = (const V*)vals;
      V* ov = (V*)out_vals;
      void* params[] = {(void*)&tt, (void*)&keys, (void*)&vv, (void*)&n, (void*)&out_keys, (void*)&ov, (void*)&count,
                        (void*)&out_cap};
      MP4J_CUDA(cudaLaunchCooperativeKernel((const void*)kfn, dim3(grid), dim3(kMapThreads), params, 0, s));
    });
  });
}

void map_launch_merge_local(int dtype, int op, const MapTable& t, const long long* keys, const void* vals,
                            long long n, int grid, cudaStream_t s) {
  dispatch_v(dtype, [&](auto v) {
    using V = decltype(v);
    dispatch_op(op, [&](auto o) {
      map_merge_local_kernel<V, decltype(o)::value><<<grid, kMapThreads, 0, s>>>(t, keys, (const V*)vals, n);
    });
  });
  MP4J_CUDA(cudaGetLastError());
}

void map_launch_compact(int dtype, const MapTable& t, long long* out_keys, void* out_vals,
                        unsigned long long* count, unsigned long long out_cap, int grid, cudaStream_t s) {
  dispatch_v(dtype, [&](auto v) {
    using V = decltype(v);
    map_compact_kernel<V><<<grid, kMapThreads, 0, s>>>(t, out_keys, (V*)out_vals, count, out_cap);
  });
  MP4J_CUDA(cudaGetLastError());
}

void map_launch_table(int dtype, int op, const PeerTable& tab, const MapTable* t, long long* const* own_keys,
                      void* const* own_vals, unsigned long long* const* own_count, int nlocal,
                      unsigned long long own_cap, unsigned long long rkeys_off, unsigned long long rvals_off,
                      unsigned long long rcnt_off, unsigned long long ocnt_off, long long cap, int sms, cudaStream_t s) {
  MapTableArgs a{};
  a.tab = tab;
  a.rkeys_off = rkeys_off;
  a.rvals_off = rvals_off;
  a.rcnt_off = rcnt_off;
  a.ocnt_off = ocnt_off;
  a.cap = cap;
  a.own_cap = own_cap;
  for (int i = 0; i < nlocal; ++i) {
    a.t[i] = t[i];
    a.own_keys[i] = own_keys[i];
    a.own_vals[i] = own_vals[i];
    a.own_count[i] = own_count[i];
  }
  dispatch_v(dtype, [&](auto v) {
    dispatch_op(op, [&](auto o) {
      auto kfn = map_table_kernel<decltype(v), decltype(o)::value>;
      int per_sm = 0;
      MP4J_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kfn, kMapThreads, 0));
      const int rows = tab.emulate ? tab.world : 1;
      const int grid = std::max(1, std::min(per_sm, 4) * sms / rows);  // co-resident: the kernel has grid-wide syncs
      void* params[] = {(void*)&a};
      MP4J_CUDA(cudaLaunchCooperativeKernel((const void*)kfn, dim3(grid, rows), dim3(kMapThreads), params, 0, s));
    });
  });
}

void map_launch_gather(int dtype, const PeerTable& tab, unsigned long long ocnt_off, unsigned long long okeys_off,
                       unsigned long long ovals_off, unsigned long long out_cap, const long long* const* own_keys,
                       const void* const* own_vals, const unsigned long long* const* own_count,
                       unsigned long long* const* info, unsigned long long overflow_off, int nlocal, bool publish,
                       bool push, bool use_mc, int grid, cudaStream_t s) {
  MapGatherArgs a{};
  a.tab = tab;
  a.ocnt_off = ocnt_off;
  a.okeys_off = okeys_off;
  a.ovals_off = ovals_off;
  a.out_cap = out_cap;
  for (int i = 0; i < nlocal; ++i) {
    a.own_keys[i] = own_keys[i];
    a.own_vals[i] = own_vals[i];
    a.own_count[i] = own_count[i];
    a.info[i] = info ? info[i] : nullptr;
  }
  a.overflow_off = overflow_off;
  a.use_mc = use_mc ? 1 : 0;
  dispatch_v(dtype, [&](auto v) { launch_gather<decltype(v)>(a, publish, push, grid, s); });
}

// =========================================================================================
// Allocation-free reduce-by-key.
//
// mp4j_map_reduce_local_noalloc is the one entry point of this library that never asks CUDA
// for memory. It exists for callers that are not allowed, or not able, to allocate on the device
// at the moment they need a reduce-by-key:
//   * code that must not synchronise the device (cudaMalloc / cudaFree are synchronising calls, so
//     an allocation in the middle of a pipelined step stalls every stream of the GPU);
//   * processes whose device memory has been handed over in full to another owner (a framework's
//     caching pool, a serving runtime with a fixed budget): a late cudaMalloc there fails or
//     fragments the pool;
//   * embedders that link only this C ABI and have no plane object (mp4j_gpu) whose cached
//     scratch they could reuse.
// The hash table lives in the file-scope device array below instead: kNoAllocTableSlots slots of
// (int64 key, 8-byte value) = 8 MiB that every GPU that loads the library keeps until the process
// exits. That bounds one call to kNoAllocMaxPairs = 262,144 input pairs (load factor <= 0.5).
//
// All calls on a device share that one array and it keeps whatever the previous call left in it,
// so each call first resets the part of the table it is going to use (keys to the empty marker,
// values to the operator's identity) and zeroes *out_count. Two calls that overlap on the GPU
// (different streams, or different host threads) would scribble over each other's table: issue
// them on one stream, or serialise them yourself.
// =========================================================================================
namespace {

constexpr unsigned long long kNoAllocMaxPairs = 1ull << 18;
constexpr unsigned long long kNoAllocTableSlots = 2 * kNoAllocMaxPairs;
constexpr size_t kNoAllocScratchBytes = kNoAllocTableSlots * 16;

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

constexpr int kNoAllocMaxDevices = 64;
std::mutex g_noalloc_mu;
void* g_noalloc_ptr[kNoAllocMaxDevices] = {};
int g_noalloc_sms[kNoAllocMaxDevices] = {};

// In host code `g_noalloc_scratch` is only a handle for the runtime, not a device address: the
// address a kernel can use has to be asked for, once per device, with that device current. The
// first call of the entry point on a device does it, so make that call before any stream capture
// is started on that device.
void* noalloc_scratch_for_current_device(int* sms) {
  int dev = -1;
  MP4J_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kNoAllocMaxDevices) throw Error("no-alloc reduce-by-key: device ordinal out of range");
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (!g_noalloc_ptr[dev]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess || p == nullptr)
      throw Error(std::string("no-alloc reduce-by-key: cudaGetSymbolAddress failed: ") + cudaGetErrorString(e));
    int n = 0;
    MP4J_CUDA(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
    g_noalloc_sms[dev] = n;
    g_noalloc_ptr[dev] = p;
  }
  *sms = g_noalloc_sms[dev];
  return g_noalloc_ptr[dev];
}

}  // namespace
}  // namespace mp4j

#include "../capi/capi_util.h"

extern "C" int mp4j_map_reduce_local_noalloc(const int64_t* keys, const void* vals, int64_t n, int dtype, int op,
                                             int64_t* out_keys, void* out_vals, int64_t out_cap, uint64_t* out_count,
                                             void* stream) {
  using namespace mp4j;
  MP4J_API_BEGIN
  if (n < 0 || out_cap < 0) throw Error("no-alloc reduce-by-key: negative size");
  if ((unsigned long long)n > kNoAllocMaxPairs)
    throw Error("no-alloc reduce-by-key: n = " + std::to_string(n) + " pairs exceeds the built-in scratch limit of " +
                std::to_string(kNoAllocMaxPairs) + " pairs; use mp4j_gpu_map_reduce_local (allocating) for larger inputs");
  if (!(dtype == MP4J_F32 || dtype == MP4J_F64 || dtype == MP4J_I32 || dtype == MP4J_I64))
    throw Error("no-alloc reduce-by-key: values must be float/double/int/long");
  if (!(op == MP4J_SUM || op == MP4J_MAX || op == MP4J_MIN)) throw Error("no-alloc reduce-by-key: op must be SUM/MAX/MIN");
  if (!out_count || (n > 0 && (!keys || !vals)) || (out_cap > 0 && (!out_keys || !out_vals)))
    throw Error("no-alloc reduce-by-key: null argument");
  int sms = 0;
  char* scratch = (char*)noalloc_scratch_for_current_device(&sms);
  cudaStream_t s = (cudaStream_t)stream;
  unsigned long long tsize = 1024;
  while (tsize < 2ull * (unsigned long long)n) tsize <<= 1;  // <= kNoAllocTableSlots by the check above
  MapTable t{(long long*)scratch, scratch + kNoAllocTableSlots * 8, tsize - 1};
  const int blocks = std::max(1, std::min(sms * 4, kMaxBlocks - 2));
  map_launch_clear(dtype, op, t, blocks, s);         // stale contents of the previous call go here
  MP4J_CUDA(cudaMemsetAsync(out_count, 0, 8, s));    // the counter accumulates, so it starts at zero
  if (n > 0) map_launch_merge_local(dtype, op, t, (const long long*)keys, vals, n, blocks, s);
  map_launch_compact(dtype, t, (long long*)out_keys, out_vals, (unsigned long long*)out_count,
                     (unsigned long long)out_cap, blocks, s);
  MP4J_API_END
}
