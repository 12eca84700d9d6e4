// Host-side orchestration of the sparse (map container) collectives: buffers, launch order, and the
// small info block the caller reads after synchronising the stream. Kernels: map_kernels.cu.
#include <algorithm>

#include "cuda_check.h"
#include "gpu_plane.h"
#include "launch.h"

namespace mp4j {

struct GpuPlane::MapWork {
  // symmetric (same offsets on every rank)
  long long cap = 0, out_cap = 0;
  size_t rkeys_off = 0, rvals_off = 0, rcnt_off = 0, cursor_off = 0, ocnt_off = 0, okeys_off = 0, ovals_off = 0;
  bool have_sym = false;
  unsigned long long calls = 0;  // owned-count slots alternate by call parity (see map_collective)
  // local scratch, one set per local rank
  struct Local {
    long long* tkeys = nullptr;
    void* tvals = nullptr;
    unsigned long long tsize = 0;
    long long* own_keys = nullptr;
    void* own_vals = nullptr;
    unsigned long long own_cap = 0;
    unsigned long long* own_count = nullptr;  // device
    unsigned long long* info_host = nullptr;  // pinned + mapped, 4 x u64
    unsigned long long* info_dev = nullptr;
  } local[kMaxRanks];
};

namespace {

__global__ void map_info_kernel(const unsigned long long* ocnt, int p, unsigned long long* cursor_overflow,
                                const unsigned long long* own_count, unsigned long long out_cap, int mode,
                                unsigned long long* info) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    unsigned long long total = 0;
    if (mode == 1) total = *own_count;
    else
      for (int q = 0; q < p; ++q) total += ocnt[q];
    info[0] = total < out_cap ? total : out_cap;
    info[1] = cursor_overflow ? *cursor_overflow : 0;
    if (cursor_overflow) *cursor_overflow = 0;  // not sticky across calls
    info[2] = total > out_cap ? 1 : 0;
    info[3] = *own_count;
  }
}

__global__ void map_set_count_kernel(unsigned long long* dst, unsigned long long v) {
  if (threadIdx.x == 0 && blockIdx.x == 0) *dst = v;
}

unsigned long long next_pow2(unsigned long long x) {
  unsigned long long p = 1;
  while (p < x) p <<= 1;
  return p;
}

}  // namespace

void GpuPlane::map_release() {
  if (!map_) return;
  for (auto& l : map_->local) {
    if (l.tkeys) cudaFree(l.tkeys);
    if (l.tvals) cudaFree(l.tvals);
    if (l.own_keys) cudaFree(l.own_keys);
    if (l.own_vals) cudaFree(l.own_vals);
    if (l.own_count) cudaFree(l.own_count);
    if (l.info_host) cudaFreeHost(l.info_host);
  }
  delete map_;
  map_ = nullptr;
}

void GpuPlane::map_reduce_local(int local_rank, const long long* keys, const void* vals, long long n, int dtype,
                                int op, long long* out_keys, void* out_vals, long long out_cap,
                                unsigned long long* out_count, cudaStream_t s) {
  if (local_rank < 0 || local_rank >= local_ranks()) throw Error("map_reduce_local: bad local rank");
  if (!(dtype == MP4J_F32 || dtype == MP4J_F64 || dtype == MP4J_I32 || dtype == MP4J_I64))
    throw Error("GPU maps support float/double/int/long values");
  if (!(op == MP4J_SUM || op == MP4J_MAX || op == MP4J_MIN)) throw Error("GPU maps support SUM/MAX/MIN");
  if (n < 0 || out_cap < 0) throw Error("map_reduce_local: bad sizes");
  if (!map_) map_ = new MapWork();
  auto& l = map_->local[local_rank];
  const unsigned long long tsize = next_pow2(std::max<unsigned long long>(2ull * (unsigned long long)n, 1024));
  if (l.tsize < tsize) {
    MP4J_CUDA(cudaStreamSynchronize(s));
    if (l.tkeys) cudaFree(l.tkeys);
    if (l.tvals) cudaFree(l.tvals);
    MP4J_CUDA(cudaMalloc(&l.tkeys, tsize * 8));
    MP4J_CUDA(cudaMalloc(&l.tvals, tsize * 8));
    l.tsize = tsize;
  }
  // use only as much of the table as this call needs: clearing dominates small calls otherwise
  MapTable t{l.tkeys, l.tvals, tsize - 1};
  const int blocks = std::max(1, std::min(sms_ * 4, kMaxBlocks - 2));
  map_launch_local(dtype, op, t, keys, vals, n, out_keys, out_vals, out_count, (unsigned long long)out_cap, sms_, s);
  last_ = CollStat{"map_reduce_local", "hash-merge", (size_t)n * (8 + dtype_size(dtype)), blocks, false};
}

void GpuPlane::map_collective(int mode, const long long* const* keys, const void* const* vals, const long long* n,
                              long long max_n, long long out_cap, int dtype, int op, int partition,
                              long long** out_keys, void** out_vals, unsigned long long** info_host,
                              cudaStream_t s) {
  entry_check();
  const int p = world();
  const int L = local_ranks();
  if (mode < 0 || mode > 2) throw Error("map_collective: bad mode");
  if (!(dtype == MP4J_F32 || dtype == MP4J_F64 || dtype == MP4J_I32 || dtype == MP4J_I64))
    throw Error("GPU maps support float/double/int/long values");
  if (mode != 2 && !(op == MP4J_SUM || op == MP4J_MAX || op == MP4J_MIN)) throw Error("GPU maps support SUM/MAX/MIN");
  if (max_n < 0 || out_cap <= 0) throw Error("map_collective: bad sizes");
  for (int i = 0; i < L; ++i)
    if (n[i] < 0 || n[i] > max_n) throw Error("map_collective: n exceeds max_n");
  if (!map_) map_ = new MapWork();
  MapWork& w = *map_;
  const size_t vs = dtype_size(dtype);

  // ---- symmetric buffers: receive regions sized for the worst case when that is small, else for a
  // 2x-balanced key distribution (overflow is detected and reported, never silent)
  long long cap = max_n;
  if ((size_t)p * (size_t)max_n * 16 > (256u << 20)) cap = std::min<long long>(max_n, 2 * ((max_n + p - 1) / p) + 4096);
  cap = std::max<long long>(cap, 1);
  if (!w.have_sym || cap > w.cap || out_cap > w.out_cap) {
    MP4J_CUDA(cudaStreamSynchronize(s));
    if (w.have_sym)
      for (size_t off : {w.rkeys_off, w.rvals_off, w.rcnt_off, w.cursor_off, w.ocnt_off, w.okeys_off, w.ovals_off})
        free(off);
    w.cap = std::max(cap, w.cap);
    w.out_cap = std::max(out_cap, w.out_cap);
    w.rkeys_off = alloc((size_t)p * w.cap * 8);
    w.rvals_off = alloc((size_t)p * w.cap * 8);
    w.rcnt_off = alloc(256);
    w.cursor_off = alloc(256);
    w.ocnt_off = alloc(256);
    w.okeys_off = alloc((size_t)w.out_cap * 8);
    w.ovals_off = alloc((size_t)w.out_cap * 8);
    w.have_sym = true;
    for (int i = 0; i < L; ++i) {
      const int r = emulated() ? i : rank();
      MP4J_CUDA(cudaMemsetAsync(heap_.base(r) + w.rcnt_off, 0, 256, s));
      MP4J_CUDA(cudaMemsetAsync(heap_.base(r) + w.cursor_off, 0, 256, s));
      MP4J_CUDA(cudaMemsetAsync(heap_.base(r) + w.ocnt_off, 0, 256, s));
    }
    MP4J_CUDA(cudaStreamSynchronize(s));
    // Every rank re-created its buffers in the same order, so offsets agree. A peer must not push
    // pairs or counts into them before this rank's zeroing has completed (it would be wiped out:
    // found at 8 GPUs x 2M keys, where a fast rank's counts arrived before a slow rank's memset),
    // so the ranks meet on the host once per (re)allocation.
    host_barrier();
  }
  // ---- local scratch
  const unsigned long long recv_upper = (unsigned long long)p * (unsigned long long)w.cap;
  const unsigned long long tsize = next_pow2(std::max<unsigned long long>(2 * recv_upper, 1024));
  for (int i = 0; i < L; ++i) {
    auto& l = w.local[i];
    if (!l.own_count) {
      MP4J_CUDA(cudaMalloc(&l.own_count, 64));
      void* h = nullptr;
      MP4J_CUDA(cudaHostAlloc(&h, 64, cudaHostAllocMapped));
      l.info_host = (unsigned long long*)h;
      void* d = nullptr;
      MP4J_CUDA(cudaHostGetDevicePointer(&d, h, 0));
      l.info_dev = (unsigned long long*)d;
    }
    if (mode != 2 && l.tsize < tsize) {
      MP4J_CUDA(cudaStreamSynchronize(s));
      if (l.tkeys) cudaFree(l.tkeys);
      if (l.tvals) cudaFree(l.tvals);
      MP4J_CUDA(cudaMalloc(&l.tkeys, tsize * 8));
      MP4J_CUDA(cudaMalloc(&l.tvals, tsize * 8));
      l.tsize = tsize;
    }
    if (mode != 2 && l.own_cap < recv_upper) {
      MP4J_CUDA(cudaStreamSynchronize(s));
      if (l.own_keys) cudaFree(l.own_keys);
      if (l.own_vals) cudaFree(l.own_vals);
      MP4J_CUDA(cudaMalloc(&l.own_keys, recv_upper * 8));
      MP4J_CUDA(cudaMalloc(&l.own_vals, recv_upper * 8));
      l.own_cap = recv_upper;
    }
  }
  // Owned counts are published into the peers' memory BEFORE the barrier of this call, i.e. possibly
  // while a slower peer still runs the info kernel of the previous call: alternate between two slots.
  const size_t ocnt_off = w.ocnt_off + (size_t)(w.calls++ & 1) * 128;
  const PeerTable tab = table();
  const long long* own_k[kMaxRanks];
  const void* own_v[kMaxRanks];
  const unsigned long long* own_c[kMaxRanks];

  unsigned long long* own_cw[kMaxRanks];
  unsigned long long* info_dev[kMaxRanks];
  for (int i = 0; i < L; ++i) {
    own_cw[i] = w.local[i].own_count;
    own_c[i] = w.local[i].own_count;
    info_dev[i] = w.local[i].info_dev;
    info_host[i] = w.local[i].info_host;
  }
  int ggrid = std::max(1, std::min(sms_ * 2, kMaxBlocks - 2));
  if (emulated()) ggrid = std::max(1, sms_ / p);
  if (mode != 2) {
    // launch 1: partition + push to owners (meets the peers); launch 2: clear + merge + compact + publish (meets
    // the peers again: nobody starts the next call's pushes into a rank that is still merging)
    MapExchangeArgs ea{};
    ea.tab = tab;
    ea.partition = partition;
    ea.cap = w.cap;
    ea.rkeys_off = w.rkeys_off;
    ea.rvals_off = w.rvals_off;
    ea.rcnt_off = w.rcnt_off;
    ea.cursor_off = w.cursor_off;
    for (int i = 0; i < L; ++i) {
      ea.keys[i] = keys[i];
      ea.vals[i] = vals[i];
      ea.n[i] = n[i];
    }
    int pgrid = (int)std::max<long long>(1, std::min<long long>((max_n + map_threads() - 1) / map_threads(), sms_ * 2));
    if (emulated()) pgrid = std::max(1, std::min(pgrid, sms_ * 2 / p));
    map_launch_partition(dtype, ea, pgrid, s);
    MapTable tt[kMaxRanks];
    long long* own_kw[kMaxRanks];
    void* own_vw[kMaxRanks];
    for (int i = 0; i < L; ++i) {
      auto& l = w.local[i];
      tt[i] = MapTable{l.tkeys, l.tvals, l.tsize - 1};
      own_kw[i] = l.own_keys;
      own_vw[i] = l.own_vals;
      own_k[i] = l.own_keys;
      own_v[i] = l.own_vals;
    }
    map_launch_table(dtype, op, tab, tt, own_kw, own_vw, own_cw, L, w.local[0].own_cap, w.rkeys_off, w.rvals_off,
                     w.rcnt_off, ocnt_off, w.cap, sms_, s);
  } else {
    for (int i = 0; i < L; ++i) {
      map_set_count_kernel<<<1, 32, 0, s>>>(w.local[i].own_count, (unsigned long long)n[i]);
      own_k[i] = keys[i];
      own_v[i] = vals[i];
    }
  }
  const unsigned long long overflow_off = w.cursor_off + (size_t)p * 8;
  if (mode == 1) {
    for (int i = 0; i < L; ++i) {
      auto& l = w.local[i];
      const int r = emulated() ? i : rank();
      map_info_kernel<<<1, 32, 0, s>>>(nullptr, p, (unsigned long long*)(heap_.base(r) + overflow_off), l.own_count,
                                       l.own_cap, mode, l.info_dev);
      out_keys[i] = l.own_keys;
      out_vals[i] = l.own_vals;
    }
    MP4J_CUDA(cudaGetLastError());
    last_ = CollStat{"reduce_scatter_map", "partition+table", (size_t)max_n * (8 + vs), ggrid, false};
    return;
  }
  // map_nvls: 1 = multicast push when AUTO would use NVLS for a dense collective (world > nvls_min_world), 2 = always
  const bool use_mc = opt_.map_nvls >= 2 || (opt_.map_nvls == 1 && p > opt_.nvls_min_world);
  // launch 3: push owned pairs to every peer's output at the right offset (the counts were published by launch 2;
  // allgather-only calls publish them here) and write the info block
  map_launch_gather(dtype, tab, ocnt_off, w.okeys_off, w.ovals_off, (unsigned long long)out_cap, own_k, own_v, own_c,
                    info_dev, mode == 0 ? overflow_off : 0, L, mode == 2, true, use_mc, ggrid, s);
  for (int i = 0; i < L; ++i) {
    const int r = emulated() ? i : rank();
    out_keys[i] = (long long*)(heap_.base(r) + w.okeys_off);
    out_vals[i] = heap_.base(r) + w.ovals_off;
  }
  MP4J_CUDA(cudaGetLastError());
  last_ = CollStat{mode == 0 ? "allreduce_map" : "allgather_map", mode == 0 ? "partition+table+gather" : "gather",
                   (size_t)max_n * (8 + vs), ggrid, false};
}

}  // namespace mp4j
