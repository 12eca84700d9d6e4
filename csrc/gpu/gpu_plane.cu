#include "gpu_plane.h"

#include <stdlib.h>
#include <string.h>

#include <algorithm>

#include "../host/process_comm.h"
#include "cuda_check.h"
#include "custom_op.h"
#include "launch.h"

namespace mp4j {

namespace {
size_t env_size(const char* name, size_t dflt) {
  const char* v = getenv(name);
  return (v && *v) ? (size_t)strtoull(v, nullptr, 10) : dflt;
}
size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }
}  // namespace

// predefined operators are checked against the dtype table; a custom one against the type it was compiled for
bool plane_op_ok(int dtype, int op) {
  if (is_custom_op(op)) return custom_op_dtype(op) == dtype;
  return op_valid_for(dtype, op);
}

void GpuOptions::load_env() {
  heap_bytes = env_size("MP4J_HEAP_BYTES", heap_bytes);
  staging_bytes = env_size("MP4J_STAGING_BYTES", staging_bytes);
  thread_slot_bytes = env_size("MP4J_THREAD_SLOT_BYTES", thread_slot_bytes);
  if (getenv("MP4J_THREAD_SLOT_BYTES")) thread_slot_auto = false;
  thread_slots = (int)env_size("MP4J_THREAD_SLOTS", (size_t)thread_slots);
  oneshot_max_bytes = env_size("MP4J_ONE_SHOT_MAX_BYTES", oneshot_max_bytes);
  nvls_min_bytes = env_size("MP4J_NVLS_MIN_BYTES", nvls_min_bytes);
  rs_nvls_min_bytes = env_size("MP4J_RS_NVLS_MIN_BYTES", rs_nvls_min_bytes);
  ag_nvls_min_bytes = env_size("MP4J_AG_NVLS_MIN_BYTES", ag_nvls_min_bytes);
  rs_nvls_max_bytes = env_size("MP4J_RS_NVLS_MAX_BYTES", rs_nvls_max_bytes);
  ag_nvls_max_bytes = env_size("MP4J_AG_NVLS_MAX_BYTES", ag_nvls_max_bytes);
  nvls64_max_bytes = env_size("MP4J_NVLS64_MAX_BYTES", nvls64_max_bytes);
  map_nvls = (int)env_size("MP4J_MAP_NVLS", (size_t)map_nvls);
  seg_relay = (int)env_size("MP4J_SEG_RELAY", (size_t)seg_relay);
  seg_relay_min_bytes = env_size("MP4J_SEG_RELAY_MIN_BYTES", seg_relay_min_bytes);
  max_blocks = (int)env_size("MP4J_MAX_BLOCKS", (size_t)max_blocks);
  nvls_min_world = (int)env_size("MP4J_NVLS_MIN_WORLD", (size_t)nvls_min_world);
  poll_ns = (int)env_size("MP4J_POLL_NS", (size_t)poll_ns);
  if (getenv("MP4J_NVLS")) want_multicast = atoi(getenv("MP4J_NVLS")) != 0;
  if (getenv("MP4J_GPU_TIMEOUT_S")) timeout_s = atof(getenv("MP4J_GPU_TIMEOUT_S"));
  host_chunk_bytes = env_size("MP4J_HOST_CHUNK_BYTES", host_chunk_bytes);
  ws_ver = (int)env_size("MP4J_WS_VER", (size_t)ws_ver);
  ws_tail_split = (int)env_size("MP4J_WS_TAIL_SPLIT", (size_t)ws_tail_split);
  ws_fa_gpu = (int)env_size("MP4J_WS_FA_GPU", (size_t)ws_fa_gpu);
  ws_dbg = (int)env_size("MP4J_WS_DBG", (size_t)ws_dbg);
}

bool GpuPlane::set_option(const std::string& name, long long v) {
  if (name == "ws_ver") opt_.ws_ver = (int)v;
  else if (name == "ws_dbg") opt_.ws_dbg = (int)v;
  else if (name == "ws_tail_split") opt_.ws_tail_split = std::max(1, (int)v);
  else if (name == "ws_fa_gpu") opt_.ws_fa_gpu = (int)v;
  else if (name == "thread_slots") opt_.thread_slots = std::max(2, std::min((int)v, thread_ws_slots()));
  else if (name == "thread_slot_bytes") {
    opt_.thread_slot_bytes = align_up((size_t)v, 16 * 8 * 8);
    opt_.thread_slot_auto = false;
  }
  else if (name == "nvls_min_world") opt_.nvls_min_world = (int)v;
  else if (name == "nvls_min_bytes") opt_.nvls_min_bytes = (size_t)v;
  else if (name == "rs_nvls_min_bytes") opt_.rs_nvls_min_bytes = (size_t)v;
  else if (name == "ag_nvls_min_bytes") opt_.ag_nvls_min_bytes = (size_t)v;
  else if (name == "rs_nvls_max_bytes") opt_.rs_nvls_max_bytes = (size_t)v;
  else if (name == "ag_nvls_max_bytes") opt_.ag_nvls_max_bytes = (size_t)v;
  else if (name == "nvls64_max_bytes") opt_.nvls64_max_bytes = (size_t)v;
  else if (name == "map_nvls") opt_.map_nvls = (int)v;
  else if (name == "seg_relay") opt_.seg_relay = (int)v;
  else if (name == "seg_relay_min_bytes") opt_.seg_relay_min_bytes = (size_t)v;
  else if (name == "oneshot_max_bytes") opt_.oneshot_max_bytes = std::min((size_t)v, oneshot_slot_bytes_);
  else if (name == "max_blocks") opt_.max_blocks = (int)v;
  else if (name == "poll_ns") opt_.poll_ns = (int)v;
  else if (name == "timeout_ms") opt_.timeout_s = (double)v / 1e3;
  else if (name == "dead_rank") dead_rank_p1_ = (int)v + 1;  // test hook: -1 = none
  else if (name == "host_chunk_bytes") {
    hpipe_release();
    opt_.host_chunk_bytes = (size_t)v;
  }
  else return false;
  return true;
}

GpuPlane* GpuPlane::create_real(ProcessComm* pc, int device, GpuOptions opt) {
  std::unique_ptr<GpuPlane> g(new GpuPlane());
  g->device_ = device;
  MP4J_CUDA(cudaSetDevice(device));
  g->pc_ = pc;
  g->heap_.init_real(pc, device, opt.heap_bytes, opt.want_multicast);
  g->finish_init(opt);
  return g.release();
}

GpuPlane* GpuPlane::create_emulated(int world, int device, GpuOptions opt) {
  std::unique_ptr<GpuPlane> g(new GpuPlane());
  g->device_ = device;
  MP4J_CUDA(cudaSetDevice(device));
  g->heap_.init_emulated(world, device, opt.heap_bytes);
  g->finish_init(opt);
  return g.release();
}

void GpuPlane::finish_init(GpuOptions opt) {
  opt_ = opt;
  MP4J_CUDA(cudaDeviceGetAttribute(&sms_, cudaDevAttrMultiProcessorCount, device_));
  void* h = nullptr;
  MP4J_CUDA(cudaHostAlloc(&h, 64, cudaHostAllocMapped));
  err_host_ = (uint32_t*)h;
  *err_host_ = 0;
  void* d = nullptr;
  MP4J_CUDA(cudaHostGetDevicePointer(&d, h, 0));
  err_dev_ = (uint32_t*)d;
  // heap layout
  size_t off = kSigPadBytes;
  oneshot_slot_bytes_ = std::max<size_t>(opt_.oneshot_max_bytes, 256u << 10);  // capacity is not the AUTO threshold
  const size_t one = align_up(oneshot_slot_bytes_ + 256, 256);
  oneshot_off_[0] = off;
  oneshot_off_[1] = off + one;
  off += 2 * one;
  const int tgrid = std::min(kMaxBlocks, sms_ * thread_max_blocks_per_sm());
  opt_.thread_slot_bytes = align_up(opt_.thread_slot_bytes, 16 * 8 * 8);
  thread_off_ = align_up(off, 4096);
  opt_.thread_slots = std::max(2, std::min(opt_.thread_slots, thread_ws_slots()));
  // room for the deepest ring of the sweeps (8 slots of >= 64 KiB per SM), so set_option() can retune in place
  thread_bytes_ = (size_t)std::max(tgrid, sms_ * std::max(opt_.thread_slots, thread_ws_slots())) *
                  std::max<size_t>(opt_.thread_slot_bytes, 64 << 10);
  off = thread_off_ + thread_bytes_;
  staging_off_ = align_up(off, 4096);
  opt_.staging_bytes = align_up(opt_.staging_bytes, 4096);
  off = staging_off_ + opt_.staging_bytes;
  user_off_ = align_up(off, 4096);
  if (user_off_ > heap_.bytes())
    throw Error("symmetric heap too small: need at least " + std::to_string(user_off_) + " bytes for pads+staging, have " +
                std::to_string(heap_.bytes()));
}

// all ranks of the communicator meet on the host (master barrier); no-op for emulated planes
void GpuPlane::host_barrier() {
  if (pc_) pc_->barrier();
}

GpuPlane::~GpuPlane() {
  cudaDeviceSynchronize();
  map_release();
  hpipe_release();
  for (float* a : hist_amax_)
    if (a) cudaFree(a);
  for (float* a : lr_scratch_)
    if (a) cudaFree(a);
  for (int* a : lr_tile_cnt_)
    if (a) cudaFree(a);
  if (err_host_) cudaFreeHost(err_host_);
}

std::string GpuPlane::describe() const {
  return std::string("mode=") + heap_.mode() + " world=" + std::to_string(world()) + " rank=" + std::to_string(rank()) +
         " heap=" + std::to_string(heap_.bytes()) + " staging=" + std::to_string(opt_.staging_bytes) +
         " user_off=" + std::to_string(user_off_) + " sms=" + std::to_string(sms_) +
         (heap_.notes().empty() ? "" : " notes=[" + heap_.notes() + "]");
}

PeerTable GpuPlane::table() const {
  PeerTable t{};
  t.world = world();
  t.rank = rank();
  t.emulate = emulated() ? 1 : 0;
  for (int q = 0; q < world(); ++q) {
    t.heap[q] = heap_.base(q);
    t.sig[q] = (uint32_t*)heap_.base(q);
  }
  t.mc_heap = heap_.mc_base();
  t.err = err_dev_;
  t.timeout_ns = (unsigned long long)(opt_.timeout_s * 1e9);
  t.poll_ns = opt_.poll_ns;
  return t;
}

size_t GpuPlane::alloc(size_t bytes) {
  std::lock_guard<std::mutex> lk(alloc_mu_);
  bytes = align_up(std::max<size_t>(bytes, 1), 256);
  size_t cur = user_off_;
  for (auto& kv : used_) {  // ordered by offset: first fit
    if (kv.first - cur >= bytes) break;
    cur = kv.first + kv.second;
  }
  if (cur + bytes > heap_.bytes())
    throw Error("symmetric heap exhausted (raise MP4J_HEAP_BYTES): want " + std::to_string(bytes));
  used_[cur] = bytes;
  return cur;
}

void GpuPlane::free(size_t off) {
  std::lock_guard<std::mutex> lk(alloc_mu_);
  used_.erase(off);
}

void GpuPlane::check_error() {
  if (err_host_ && *err_host_ >= 2) {
    poisoned_ = true;
    *err_host_ = 0;
    throw Error("GPU collective: the ranks disagree on where the array lives - a symmetric tensor (comm.empty) must be "
                "the same allocation, at the same offset, on every rank, and symmetric and ordinary tensors must not "
                "be mixed in one call; the results of that call are undefined and this GPU plane is now unusable");
  }
  if (err_host_ && *err_host_) {
    poisoned_ = true;
    *err_host_ = 0;
    throw Error("GPU collective: a peer did not arrive within the device-side timeout (dead or diverged rank?); "
                "the kernel stored nothing after the failed wait and this GPU plane is now unusable");
  }
  if (poisoned_)
    throw Error("GPU plane is unusable: an earlier collective hit the peer-wait timeout (close the communicator)");
}

void GpuPlane::entry_check() { check_error(); }

int GpuPlane::pick_grid(long long max_units, int max_blocks_per_sm) const {
  long long cap = (long long)sms_ * max_blocks_per_sm;
  // emulation: all rows of the cooperative grid must be co-resident, and instantiations other than
  // the probed one may only fit one 512-thread block per SM
  if (emulated()) cap = sms_ / world();
  cap = std::min<long long>(cap, kMaxBlocks);
  if (opt_.max_blocks > 0) cap = std::min<long long>(cap, opt_.max_blocks);
  cap = std::max<long long>(cap, 1);
  long long want = (max_units + (long long)grs_threads() * 2 - 1) / ((long long)grs_threads() * 2);
  return (int)std::max<long long>(1, std::min(want, cap));
}

// true if every local array sits inside the symmetric user region at one common offset
bool GpuPlane::offsets_symmetric(void* const* arrs, size_t* off) const {
  const int L = local_ranks();
  for (int i = 0; i < L; ++i) {
    const int r = emulated() ? i : rank();
    const char* p = (const char*)arrs[i];
    const char* b = heap_.base(r);
    if (p < b + user_off_ || p >= b + heap_.bytes()) return false;
    const size_t o = (size_t)(p - b);
    if (i == 0) *off = o;
    else if (o != *off) return false;
  }
  return true;
}

void GpuPlane::run_grs(const char* name, void* const* arrs, int dtype, int op, const int64_t* lo, const int64_t* hi,
                       const unsigned* src_mask, const unsigned* dst_mask, bool reduce_like, int algo,
                       cudaStream_t s) {
  const int p = world();
  const int L = local_ranks();
  const size_t es = dtype_size(dtype);
  if (es == 0) throw Error("unknown dtype");
  if (op != MP4J_COPY && !plane_op_ok(dtype, op)) throw Error("operator is not defined for this dtype");
  // unit conversion: COPY kernels work on bytes
  const long long scale = (op == MP4J_COPY) ? (long long)es : 1;
  const size_t kes = (op == MP4J_COPY) ? 1 : es;
  long long span_lo = INT64_MAX, span_hi = INT64_MIN, max_n = 0;
  for (int q = 0; q < p; ++q) {
    if (hi[q] <= lo[q]) continue;
    span_lo = std::min<long long>(span_lo, lo[q]);
    span_hi = std::max<long long>(span_hi, hi[q]);
    max_n = std::max<long long>(max_n, hi[q] - lo[q]);
  }
  if (max_n == 0) return;
  const size_t span_bytes = (size_t)(span_hi - span_lo) * es;
  const int family = (!strcmp(name, "reduce_scatter") || !strcmp(name, "reduce")) ? 1
                     : (!strcmp(name, "allgather") || !strcmp(name, "broadcast")) ? 2 : 0;

  GrsArgs a{};
  a.tab = table();
  a.entry_barrier = 1;
  a.exit_barrier = 1;
  a.vec_ok = 1;
  a.dead_rank_p1 = dead_rank_p1_;
  const bool nvls_ok = has_multicast() && nvls_supported(dtype, op) &&
                       (algo == MP4J_ALGO_NVLS ||
                        (algo == MP4J_ALGO_AUTO && nvls_auto(span_bytes, family) &&
                         !(es == 8 && op != MP4J_COPY && span_bytes > opt_.nvls64_max_bytes)));
  last_.name = name;
  last_.bytes = span_bytes;

  size_t sym_off = 0;
  const bool zero_copy = offsets_symmetric(arrs, &sym_off);
  last_.staged = !zero_copy;
  last_.algo = nvls_ok ? "nvls" : "p2p";
  if (zero_copy) {
    a.src_off = a.dst_off = sym_off;
    a.layout_key = 0x5000000000000000ull | (unsigned long long)sym_off;  // "symmetric at this offset"
    for (int i = 0; i < L; ++i) {
      const int q = emulated() ? i : rank();
      GrsItem& it = a.item[i];
      it.lo = lo[q] * scale;
      it.hi = std::max(hi[q], lo[q]) * scale;
      it.src_mask = src_mask[q];
      it.dst_mask = dst_mask[q];
    }
    const int grid = pick_grid((max_n * (long long)es + 15) / 16, grs_max_blocks_per_sm());
    last_.grid = grid;
    launch_grs(dtype, op, nvls_ok, a, grid, s);
    return;
  }
  // ---- staged: the arrays go through the symmetric staging buffer. If they fit, in one piece. If not, in
  // nwin rounds, and round j carries sub-range j of EVERY rank's segment (packed side by side in the window),
  // not the j-th contiguous piece of the span: a contiguous piece of a reduce-scatter / allgather / partitioned
  // allreduce belongs to one rank, so windowing the span made the ranks take turns (1 GiB at 8 GPUs: 88 GB/s).
  const long long epv = (long long)(16 / es);
  const size_t slack = (size_t)p * 256 + 4096;  // alignment padding between the packed sub-ranges
  long long nwin = 1;
  if (span_bytes > opt_.staging_bytes) {
    long long total = 0;
    for (int q = 0; q < p; ++q) total += std::max<long long>(hi[q] - lo[q], 0);
    nwin = (long long)((total * es + (opt_.staging_bytes - slack) - 1) / (opt_.staging_bytes - slack));
  }
  for (long long j = 0; j < nwin; ++j) {
    long long l[kMaxRanks], h[kMaxRanks];
    size_t S[kMaxRanks];
    size_t used = 0;
    long long wmax = 0;
    if (nwin == 1) {
      // one window: keep the arrays' own layout (element e at staging + (e - span_lo) * es)
      for (int q = 0; q < p; ++q) {
        l[q] = lo[q];
        h[q] = std::max(hi[q], lo[q]);
        wmax = std::max(wmax, h[q] - l[q]);
      }
      for (int i = 0; i < L; ++i) {
        const int r = emulated() ? i : rank();
        MP4J_CUDA(cudaMemcpyAsync(heap_.base(r) + staging_off_, (const char*)arrs[i] + (size_t)span_lo * es, span_bytes,
                                  cudaMemcpyDeviceToDevice, s));
      }
    } else {
      for (int q = 0; q < p; ++q) {
        const long long len = std::max<long long>(hi[q] - lo[q], 0);
        const long long step = ((len + nwin - 1) / nwin + epv - 1) / epv * epv;
        l[q] = std::min(lo[q] + j * step, lo[q] + len);
        h[q] = std::min(l[q] + step, lo[q] + len);
        S[q] = used;
        used = align_up(used + (size_t)(h[q] - l[q]) * es, 256);
        wmax = std::max(wmax, h[q] - l[q]);
      }
      if (used > opt_.staging_bytes) throw Error("internal: staged window overflows the staging buffer");
      for (int i = 0; i < L; ++i) {
        const int r = emulated() ? i : rank();
        for (int q = 0; q < p; ++q)
          if (h[q] > l[q])
            MP4J_CUDA(cudaMemcpyAsync(heap_.base(r) + staging_off_ + S[q], (const char*)arrs[i] + (size_t)l[q] * es,
                                      (size_t)(h[q] - l[q]) * es, cudaMemcpyDeviceToDevice, s));
      }
    }
    if (wmax <= 0) continue;
    a.src_off = a.dst_off = (unsigned long long)staging_off_;
    a.layout_key = 0x6000000000000000ull | (unsigned long long)(nwin << 20) | (unsigned long long)j;  // "staged, round j of nwin"
    for (int i = 0; i < L; ++i) {
      const int q = emulated() ? i : rank();
      GrsItem& it = a.item[i];
      it.lo = l[q] * scale;
      it.hi = h[q] * scale;
      it.src_mask = src_mask[q];
      it.dst_mask = dst_mask[q];
      // element e of this item's sub-range lives at staging + adj + e * es
      it.adj = nwin == 1 ? -(long long)((size_t)span_lo * es) : (long long)S[q] - (long long)((size_t)l[q] * es);
    }
    const int grid = pick_grid((wmax * (long long)es + 15) / 16, grs_max_blocks_per_sm());
    last_.grid = grid;
    launch_grs(dtype, op, nvls_ok, a, grid, s);
    for (int i = 0; i < L; ++i) {
      const int r = emulated() ? i : rank();
      if (nwin == 1) {
        MP4J_CUDA(cudaMemcpyAsync((char*)arrs[i] + (size_t)span_lo * es, heap_.base(r) + staging_off_, span_bytes,
                                  cudaMemcpyDeviceToDevice, s));
      } else {
        for (int q = 0; q < p; ++q)
          if (h[q] > l[q])
            MP4J_CUDA(cudaMemcpyAsync((char*)arrs[i] + (size_t)l[q] * es, heap_.base(r) + staging_off_ + S[q],
                                      (size_t)(h[q] - l[q]) * es, cudaMemcpyDeviceToDevice, s));
      }
    }
  }
  (void)reduce_like;
  (void)kes;
}

void GpuPlane::allreduce(void* const* arrs, int dtype, int op, int64_t from, int64_t to, int algo, cudaStream_t s) {
  entry_check();
  check_from_to(from, to);
  const int p = world();
  if (p == 1 || from == to) return;
  const size_t es = dtype_size(dtype);
  const size_t bytes = (size_t)(to - from) * es;
  if (!plane_op_ok(dtype, op) || op == MP4J_COPY) throw Error("operator is not defined for this dtype");
  const unsigned full = (1u << p) - 1u;
  const bool oneshot = algo == MP4J_ALGO_ONESHOT || (algo == MP4J_ALGO_AUTO && bytes <= opt_.oneshot_max_bytes);
  if (oneshot && bytes <= oneshot_slot_bytes_) {
    // every rank copies its input into a double-buffered symmetric slot, then reads all p slots
    GrsArgs a{};
    a.tab = table();
    a.entry_barrier = 1;
    // no exit barrier: the slot alternates between calls (DESIGN.md "one-shot slot reuse"). A graph
    // replay repeats the SAME slot sequence, so under capture the barrier is kept.
    a.exit_barrier = capturing(s) ? 1 : 0;
    const size_t slot = oneshot_off_[oneshot_next_];
    oneshot_next_ ^= 1;
    a.src_off = a.dst_off = (unsigned long long)slot - (unsigned long long)from * es;
    a.vec_ok = 1;
    const int L = local_ranks();
    for (int i = 0; i < L; ++i) {
      GrsItem& it = a.item[i];
      it.lo = from;
      it.hi = to;
      it.src_mask = full;
      it.dst_mask = 0;
      it.src_local = arrs[i];
      it.dst_local = arrs[i];
      if ((((uintptr_t)arrs[i] + (size_t)from * es) & 15) != 0) a.vec_ok = 0;
    }
    const int grid = pick_grid((long long)(bytes + 15) / 16, grs_max_blocks_per_sm());
    last_ = CollStat{"allreduce", "oneshot", bytes, grid, false};
    launch_grs(dtype, op, false, a, grid, s);
    return;
  }
  int64_t lo[kMaxRanks], hi[kMaxRanks];
  unsigned sm[kMaxRanks], dm[kMaxRanks];
  // internal partition is free to be 16-byte friendly: results do not depend on who reduces what
  const long long n = to - from, epv = (long long)(16 / es);
  const long long per = ((n + p - 1) / p + epv - 1) / epv * epv;
  for (int q = 0; q < p; ++q) {
    lo[q] = std::min<long long>(to, from + q * per);
    hi[q] = std::min<long long>(to, from + (q + 1) * per);
    sm[q] = dm[q] = full;
  }
  run_grs("allreduce", arrs, dtype, op, lo, hi, sm, dm, true, algo == MP4J_ALGO_ONESHOT ? MP4J_ALGO_AUTO : algo, s);
  if (strcmp(last_.algo, "p2p") == 0) last_.algo = "twoshot";
}

void GpuPlane::reduce_scatter(void* const* arrs, int dtype, int op, int64_t from, const int64_t* counts, int algo,
                              cudaStream_t s) {
  entry_check();
  const int p = world();
  check_from_counts(from, counts, p);
  if (p == 1) return;
  int64_t lo[kMaxRanks], hi[kMaxRanks];
  unsigned sm[kMaxRanks], dm[kMaxRanks];
  froms_tos_from_counts(from, counts, p, lo, hi);
  for (int q = 0; q < p; ++q) {
    sm[q] = (1u << p) - 1u;
    dm[q] = 1u << q;
  }
  run_grs("reduce_scatter", arrs, dtype, op, lo, hi, sm, dm, true, algo, s);
}

void GpuPlane::allgather(void* const* arrs, int dtype, const int64_t* froms, const int64_t* tos, int algo,
                         cudaStream_t s) {
  entry_check();
  const int p = world();
  check_froms_tos(froms, tos, p);
  if (p == 1) return;
  unsigned sm[kMaxRanks], dm[kMaxRanks];
  for (int q = 0; q < p; ++q) {
    sm[q] = 1u << q;
    dm[q] = (1u << p) - 1u;
  }
  run_grs("allgather", arrs, dtype, MP4J_COPY, froms, tos, sm, dm, false, algo, s);
}

void GpuPlane::broadcast(void* const* arrs, int dtype, int64_t from, int64_t to, int root, int algo, cudaStream_t s) {
  entry_check();
  const int p = world();
  check_from_to(from, to);
  if (root < 0 || root >= p) throw Error("root rank out of range");
  if (p == 1 || from == to) return;
  // every rank pulls one balanced slice from the root and pushes it to everybody: the root's
  // egress is N(p-1)/p instead of N(p-1), the shape of the reference's scatter+allgather (PCS:750-770)
  int64_t lo[kMaxRanks], hi[kMaxRanks];
  unsigned sm[kMaxRanks], dm[kMaxRanks];
  balanced_split(from, to, p, lo, hi);
  for (int q = 0; q < p; ++q) {
    sm[q] = 1u << root;
    dm[q] = (1u << p) - 1u;
  }
  run_grs("broadcast", arrs, dtype, MP4J_COPY, lo, hi, sm, dm, false, algo, s);
}

void GpuPlane::gather(void* const* arrs, int dtype, const int64_t* froms, const int64_t* tos, int root,
                      cudaStream_t s) {
  entry_check();
  const int p = world();
  check_froms_tos(froms, tos, p);
  if (root < 0 || root >= p) throw Error("root rank out of range");
  if (p == 1) return;
  unsigned sm[kMaxRanks], dm[kMaxRanks];
  for (int q = 0; q < p; ++q) {  // every rank pushes its own segment into the root's array
    sm[q] = 1u << q;
    dm[q] = 1u << root;
  }
  run_grs("gather", arrs, dtype, MP4J_COPY, froms, tos, sm, dm, false, MP4J_ALGO_TWOSHOT, s);
}

void GpuPlane::scatter(void* const* arrs, int dtype, const int64_t* froms, const int64_t* tos, int root,
                       cudaStream_t s) {
  entry_check();
  const int p = world();
  check_froms_tos(froms, tos, p);
  if (root < 0 || root >= p) throw Error("root rank out of range");
  if (p == 1) return;
  unsigned sm[kMaxRanks], dm[kMaxRanks];
  for (int q = 0; q < p; ++q) {  // every rank pulls its own segment out of the root's array
    sm[q] = 1u << root;
    dm[q] = 1u << q;
  }
  run_grs("scatter", arrs, dtype, MP4J_COPY, froms, tos, sm, dm, false, MP4J_ALGO_TWOSHOT, s);
}

void GpuPlane::reduce(void* const* arrs, int dtype, int op, int64_t from, int64_t to, int root, int algo,
                      cudaStream_t s) {
  entry_check();
  const int p = world();
  check_from_to(from, to);
  if (root < 0 || root >= p) throw Error("root rank out of range");
  if (p == 1 || from == to) return;
  int64_t lo[kMaxRanks], hi[kMaxRanks];
  unsigned sm[kMaxRanks], dm[kMaxRanks];
  balanced_split(from, to, p, lo, hi);
  for (int q = 0; q < p; ++q) {  // reduce-scatter whose results land directly in the root
    sm[q] = (1u << p) - 1u;
    dm[q] = 1u << root;
  }
  run_grs("reduce", arrs, dtype, op, lo, hi, sm, dm, true, algo, s);
}

void GpuPlane::thread_allreduce(void* const* bufs, int T, int dtype, int op, int64_t from, int64_t to, int fanout,
                                int dst, bool local_only, int algo, cudaStream_t s) {
  entry_check();
  check_from_to(from, to);
  if (T < 1 || T > MP4J_MAX_THREADS) throw Error("thread count out of range");
  if (dst < 0 || dst >= T) throw Error("dst thread out of range");
  if (!plane_op_ok(dtype, op) || op == MP4J_COPY) throw Error("operator is not defined for this dtype");
  if (from == to) return;
  const size_t es = dtype_size(dtype);
  const size_t nbytes = (size_t)(to - from) * es;
  const int p = local_only ? 1 : world();
  const int L = local_ranks();
  if (is_custom_op(op)) {
    // custom operators: thread-id-ordered fold into one buffer per rank (DOP:441-457), the fused
    // inter-GPU kernel on that buffer, then the fan-out copies (TCS:1915-1963 shape, all on the stream)
    const int acc = fanout ? 0 : dst;
    void* accs[kMaxRanks];
    for (int i = 0; i < L; ++i) {
      char* a = (char*)bufs[(size_t)i * T + acc] + (size_t)from * es;
      for (int t = 0; t < T; ++t)
        if (t != acc) launch_custom_fold(op, a, (const char*)bufs[(size_t)i * T + t] + (size_t)from * es, to - from, sms_, s);
      accs[i] = bufs[(size_t)i * T + acc];
    }
    if (p > 1) allreduce(accs, dtype, op, from, to, algo, s);
    if (fanout)
      for (int i = 0; i < L; ++i)
        for (int t = 1; t < T; ++t)
          MP4J_CUDA(cudaMemcpyAsync((char*)bufs[(size_t)i * T + t] + (size_t)from * es, (char*)accs[i] + (size_t)from * es,
                                    nbytes, cudaMemcpyDeviceToDevice, s));
    last_ = CollStat{"thread_allreduce", "custom-fold+grs", nbytes, 0, false};
    return;
  }
  bool vec = (nbytes % 16) == 0;
  for (int i = 0; i < L * T; ++i)
    if ((((uintptr_t)bufs[i] + (size_t)from * es) & 15) != 0) vec = false;
  const int vec_flag = vec ? 0 : 0x100;
  ThreadArgs a{};
  a.T = T;
  a.fanout = fanout;
  a.dst_index = dst;
  a.from = from;
  a.to = to;
  const long long units = (long long)(nbytes / (vec ? 16 : es));
  if (p == 1) {
    // intra-GPU only; in emulation each emulated rank reduces its own T buffers
    const bool cluster = vec && (algo & 0xff) == 1 && thread_cluster_supported(dtype, op, T, 1);
    bool fast = vec && (algo & 0xff) == 0 && (nbytes % 32) == 0 && thread_local_fast_supported(dtype, op, T);
    for (int i = 0; fast && i < L * T; ++i)
      if ((((uintptr_t)bufs[i] + (size_t)from * es) & 31) != 0) fast = false;
    for (int i = 0; i < L; ++i) {
      a.tab = table();
      a.tab.world = 1;
      a.tab.emulate = 0;
      a.tab.rank = 0;
      for (int t = 0; t < T; ++t) a.bufs[0][t] = bufs[i * T + t];
      long long cap = (long long)sms_ * thread_max_blocks_per_sm();
      if (opt_.max_blocks > 0) cap = std::min<long long>(cap, opt_.max_blocks);
      const int grid = (int)std::max<long long>(1, std::min<long long>(cap, (units + thread_threads() - 1) / thread_threads()));
      last_ = CollStat{"thread_allreduce", cluster ? "cluster-dsmem" : (fast ? "local-256bit" : "register"), nbytes,
                       grid, false};
      launch_thread_allreduce(dtype, op, false, a, grid, (cluster ? 1 : (fast ? 2 : 0)) | vec_flag, s);
    }
    return;
  }
  a.tab = table();
  for (int i = 0; i < L; ++i)
    for (int t = 0; t < T; ++t) a.bufs[i][t] = bufs[i * T + t];
  // ---- preferred: warp-specialised kernels (HBM read / NVLink / HBM write overlap inside each SM):
  // the TMA formulation (any T, 16-byte aligned) or the register formulation (T in {2,4,8}, 32-byte aligned)
  const int ws_ver = opt_.ws_ver, ws_dbg = opt_.ws_dbg;
  // thread_algo: 0 = pick, 4 = TMA formulation, 5 = first register formulation (system-scope release per
  // phase), 6 = pull formulation, 7 = hybrid (pull-style phase A/B, pushed result) (3 = chunk-synchronous
  // fallback below). ws_ver: 1 -> 5, 2 -> 4, 3 -> 6, 4 -> 7, 0 -> 6 at two ranks and 7 beyond.
  const int talgo = algo & 0xff;
  bool ws = vec && (talgo == 0 || talgo == 5 || talgo == 6 || talgo == 7) && (nbytes % 32) == 0 &&
            thread_local_fast_supported(dtype, op, T);
  for (int i = 0; ws && i < L * T; ++i)
    if ((((uintptr_t)bufs[i] + (size_t)from * es) & 31) != 0) ws = false;
  const bool tma = vec && (talgo == 0 || talgo == 4) && thread_fused_tma_supported(dtype, op, T) &&
                   (talgo == 4 || ws_ver == 2 || !ws);
  if (ws || tma) {
    const size_t wslot = (opt_.thread_slot_auto && p == 2 ? (size_t)(48 << 10) : opt_.thread_slot_bytes) / 32 * 32;
    a.slot_bytes = wslot;
    a.slots = opt_.thread_slots;
    a.stg_off = thread_off_;
    a.dbg = ws_dbg;
    a.tail_split = std::max(1, opt_.ws_tail_split);
    a.fa_gpu = opt_.ws_fa_gpu;
    long long wcap = sms_;
    if (emulated()) wcap /= p;
    wcap = std::min<long long>(wcap, (long long)(thread_bytes_ / (wslot * opt_.thread_slots)));
    wcap = std::min<long long>(wcap, thread_ws_max_blocks());
    if (opt_.max_blocks > 0) wcap = std::min<long long>(wcap, opt_.max_blocks);
    const int wgrid = (int)std::max<long long>(1, std::min<long long>(wcap, (long long)((nbytes + wslot - 1) / wslot)));
    // the fused kernels exchange slot-sized pieces whatever the array size: the in-switch reduction pays as soon
    // as it removes traffic (beyond 2 ranks), independently of the dense collectives' size thresholds
    const bool wnvls = has_multicast() && nvls_supported(dtype, op) &&
                       (algo >> 16 == MP4J_ALGO_NVLS || (algo >> 16 == MP4J_ALGO_AUTO && world() > 2));
    if (tma) {
      last_ = CollStat{"thread_allreduce", wnvls ? "fused-tma-nvls" : "fused-tma-p2p", nbytes, wgrid, false};
      launch_thread_fused_tma(dtype, op, wnvls, a, wgrid, s);
    } else if (talgo == 6 || talgo == 7 || (talgo == 0 && (ws_ver == 3 || ws_ver == 4 || ws_ver == 0))) {
      a.push = (talgo == 7 || (talgo == 0 && (ws_ver == 4 || (ws_ver == 0 && p > 2)))) ? 1 : 0;
      last_ = CollStat{"thread_allreduce",
                       a.push ? (wnvls ? "fused-hybrid-nvls" : "fused-hybrid-p2p")
                              : (wnvls ? "fused-pull-nvls" : "fused-pull-p2p"),
                       nbytes, wgrid, false};
      launch_thread_fused_pull(dtype, op, wnvls, a, wgrid, s);
    } else {
      last_ = CollStat{"thread_allreduce", wnvls ? "fused-ws-nvls" : "fused-ws-p2p", nbytes, wgrid, false};
      launch_thread_fused_ws(dtype, op, wnvls, a, wgrid, s);
    }
    return;
  }
  const size_t slot = opt_.thread_slot_bytes / (16 * p) * (16 * p);
  a.slot_bytes = slot;
  a.stg_off = thread_off_;
  long long cap = (long long)sms_ * thread_max_blocks_per_sm();
  if (emulated()) cap /= p;
  cap = std::min<long long>(cap, (long long)(thread_bytes_ / slot));
  cap = std::min<long long>(cap, kMaxBlocks);
  if (opt_.max_blocks > 0) cap = std::min<long long>(cap, opt_.max_blocks);
  const int grid = (int)std::max<long long>(1, std::min<long long>(cap, (long long)((nbytes + slot - 1) / slot)));
  const bool nvls = vec && has_multicast() && nvls_supported(dtype, op) &&
                    (algo >> 16 == MP4J_ALGO_NVLS || (algo >> 16 == MP4J_ALGO_AUTO && world() > 2));
  last_ = CollStat{"thread_allreduce", nvls ? "fused-nvls" : "fused-p2p", nbytes, grid, false};
  launch_thread_allreduce(dtype, op, nvls, a, grid, 0 | vec_flag, s);
}

}  // namespace mp4j
