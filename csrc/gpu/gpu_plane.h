// GpuPlane: host-side owner of the GPU data plane of one communicator - symmetric heap, signal
// pads, staging windows, algorithm choice and kernel launches. One per ProcessComm ("real" mode,
// one process per GPU) or one per test process ("emulated" mode: `world` ranks on one GPU, every
// collective is ONE cooperative launch whose grid rows are the ranks).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "dense.cuh"
#include "symm_heap.h"
#include "trace.h"

namespace mp4j {

class ProcessComm;
bool plane_op_ok(int dtype, int op);  // operator defined for this dtype on the GPU paths

struct GpuOptions {
  size_t heap_bytes = 1ull << 30;       // MP4J_HEAP_BYTES
  size_t staging_bytes = 128ull << 20;  // general staging window for non-symmetric tensors
  // fused thread kernel: per-block slot (MP4J_THREAD_SLOT_BYTES) and ring depth (MP4J_THREAD_SLOTS).
  // Measured (profiles/README.md §1): 64 KiB x 4 is the best trade between pipeline fill/drain and
  // L2 footprint (small) and per-chunk cross-GPU flag latency (large) at 4 and 8 GPUs; at 2 GPUs (P2P, the
  // shortest chain) 48 KiB x 4 is 3 % faster. thread_slot_auto = nobody set the size explicitly.
  size_t thread_slot_bytes = 64 << 10;
  bool thread_slot_auto = true;
  int thread_slots = 4;
  size_t oneshot_max_bytes = 256 << 10; // <= : one-shot allreduce (MP4J_ONE_SHOT_MAX_BYTES)
  size_t nvls_min_bytes = 512 << 10;    // >= : NVLS when available (MP4J_NVLS_MIN_BYTES)
  // the same threshold for the reduce-scatter family (reduce-scatter, reduce) and the allgather family (allgather,
  // broadcast); 0 = follow nvls_min_bytes. Their crossovers differ from allreduce's (tuning/*.json).
  size_t rs_nvls_min_bytes = 0;         // MP4J_RS_NVLS_MIN_BYTES
  size_t ag_nvls_min_bytes = 0;         // MP4J_AG_NVLS_MIN_BYTES
  size_t rs_nvls_max_bytes = 0;         // >= : two-shot again (0 = no upper end; MP4J_RS_NVLS_MAX_BYTES)
  size_t ag_nvls_max_bytes = 0;         // MP4J_AG_NVLS_MAX_BYTES
  // 8-byte reductions: multimem.ld_reduce has no vector form for f64/i64 (one 8-byte request per instruction), so
  // beyond this size the in-switch path is request-bound and two-shot P2P wins (8 GPUs, 1 GiB double: 545 vs
  // ~650 GB/s; NCCL 711). AUTO only. MP4J_NVLS64_MAX_BYTES
  size_t nvls64_max_bytes = 16ull << 20;
  int seg_relay = 1;  // rooted thread collectives (broadcast, reduce) beyond 2 ranks: two launches through all ports
  size_t seg_relay_min_bytes = 4ull << 20;
  int map_nvls = 1;  // map collectives: push the result through the multicast mapping (multimem.st) when there is one
  int nvls_min_world = 2;               // AUTO uses NVLS only when world > this (MP4J_NVLS_MIN_WORLD)
  bool want_multicast = true;           // MP4J_NVLS=0 disables
  int max_blocks = 0;                   // 0 = occupancy-derived (MP4J_MAX_BLOCKS)
  int poll_ns = 0;                      // sleep between device-side flag polls (MP4J_POLL_NS)
  double timeout_s = 60.0;              // device-side flag-wait watchdog (MP4J_GPU_TIMEOUT_S)
  size_t host_chunk_bytes = 8ull << 20; // per-buffer chunk of the host-array pipeline (MP4J_HOST_CHUNK_BYTES)
  int ws_tail_split = 4;                // pull kernel: steps the last chunk of a block is cut into (MP4J_WS_TAIL_SPLIT)
  int ws_fa_gpu = 1;                    // register formulation: device-scope publish of phase A (MP4J_WS_FA_GPU)
  int ws_ver = 1;                       // fused thread kernel: 1 = register formulation, 2 = TMA (MP4J_WS_VER)
  int ws_dbg = 0;                       // diagnostics bit mask for the fused thread kernels (MP4J_WS_DBG)
  void load_env();
};

struct CollStat {  // last collective, for logs / tuning
  const char* name = "";
  const char* algo = "";
  size_t bytes = 0;
  int grid = 0;
  bool staged = false;
};

class GpuPlane {
 public:
  static GpuPlane* create_real(ProcessComm* pc, int device, GpuOptions opt);
  static GpuPlane* create_emulated(int world, int device, GpuOptions opt);
  ~GpuPlane();

  int world() const { return heap_.world(); }
  int rank() const { return heap_.rank(); }
  bool emulated() const { return heap_.emulated(); }
  bool has_multicast() const { return heap_.mc_base() != nullptr; }
  std::string describe() const;
  // run-time tuning knob (sweeps): returns false for an unknown name. All ranks must set the same values.
  bool set_option(const std::string& name, long long value);
  const CollStat& last() const { return last_; }
  CollTrace& trace() { return trace_; }  // per-collective CUDA-event timing + NVTX (MP4J_TRACE=1)

  // symmetric allocation: same offset on every rank. Returns the offset; ptr(rank, off) maps it.
  size_t alloc(size_t bytes);
  void free(size_t off);
  void* ptr(int rank, size_t off) const { return heap_.base(rank) + off; }

  // Dense collectives. `arrs`: element-0 pointers of the user arrays, one per LOCAL rank
  // (1 in real mode, world in emulation). Ranges are element ranges. All in place.
  void allreduce(void* const* arrs, int dtype, int op, int64_t from, int64_t to, int algo, cudaStream_t s);
  void reduce_scatter(void* const* arrs, int dtype, int op, int64_t from, const int64_t* counts, int algo,
                      cudaStream_t s);
  void allgather(void* const* arrs, int dtype, const int64_t* froms, const int64_t* tos, int algo, cudaStream_t s);
  void broadcast(void* const* arrs, int dtype, int64_t from, int64_t to, int root, int algo, cudaStream_t s);
  void gather(void* const* arrs, int dtype, const int64_t* froms, const int64_t* tos, int root, cudaStream_t s);
  void scatter(void* const* arrs, int dtype, const int64_t* froms, const int64_t* tos, int root, cudaStream_t s);
  void reduce(void* const* arrs, int dtype, int op, int64_t from, int64_t to, int root, int algo, cudaStream_t s);
  // Thread-level fused allreduce: bufs[local_rank * T + t]. fanout=0 writes only thread `dst`.
  // local_only: skip the inter-GPU step (intra-process thread reduce, DOP:441-457).
  void thread_allreduce(void* const* bufs, int T, int dtype, int op, int64_t from, int64_t to, int fanout, int dst,
                        bool local_only, int algo, cudaStream_t s);
  // The other six ThreadCommSlave array collectives as ONE launch (thread_seg.cu): a table of segments, each
  // {lo, hi, src_rank, src_thread, dst_rank, dst_thread} (int64; src (-1,-1) = reduce over every thread of
  // every rank with `op`, else copy from that thread; dst -1 = every rank / every thread). bufs[local_rank * T + t].
  void thread_segments(void* const* bufs, int T, int dtype, int op, const int64_t* segs, int nseg, const char* name,
                       cudaStream_t s);
  // Host arrays through the GPU plane (the reference's arrays ARE host arrays, PCS:1733 / TCS:1915): the
  // range is cut into chunks; chunk c+1 is copied in (one cudaMemcpyAsync per buffer on a copy stream)
  // while the fused kernel reduces chunk c and chunk c-1 is copied out on a third stream. hbufs[local_rank *
  // T + t] are host pointers (pinned for full speed). fanout=1: the result lands in every thread's host
  // array, else only in thread `dst`'s. Synchronous: returns when the host arrays hold the result.
  void thread_allreduce_host(void* const* hbufs, int T, int dtype, int op, int64_t from, int64_t to, int fanout,
                             int dst, cudaStream_t s);
  // K8: fp32 SUM allreduce whose payload crosses NVLink as bf16 or block-scaled fp8 (lossy, opt-in).
  void allreduce_wire(float* const* arrs, int64_t from, int64_t to, int wire, cudaStream_t s);
  // Raises if a kernel reported a peer-wait timeout since the last check. After a timeout the plane is
  // unusable (the peers' epochs are out of step): every later collective raises at once.
  void check_error();

  // ---- sparse (map container) collectives: allreduceMap / reduceScatterMap / allgatherMap on
  // (int64 key, value) device arrays. One entry per LOCAL rank in every array argument.
  // mode 0 = allreduce (every rank gets the merged map), 1 = reduce-scatter (every rank gets the
  // keys it owns), 2 = allgather (concatenation in rank order, no reduction).
  // max_n = max over ALL ranks of n (agreed by the caller); out_cap = capacity of the result in
  // pairs, identical on every rank. Results live in plane-owned buffers that stay valid until the
  // next map call: *out_keys/*out_vals receive their addresses, info_host (pinned, 4 x u64 per local
  // rank, valid after the stream is synchronised) = {pairs, receive overflow, output overflow, own}.
  void map_collective(int mode, const long long* const* keys, const void* const* vals, const long long* n,
                      long long max_n, long long out_cap, int dtype, int op, int partition, long long** out_keys,
                      void** out_vals, unsigned long long** info_host, cudaStream_t s);
  // K12: LR gradient G = sum_ranks Xt[F x K] * Rt[C x K]^T as a tcgen05 GEMM whose tiles are added into
  // the owner rank's shard (rows [q*fpr, (q+1)*fpr) of G) as they retire. xt/rt: bf16, K-major, one
  // pointer per local rank. shard_out[i] = this rank's [fpr x C] fp32 shard (plane-owned, valid until
  // the next call).
  void lr_xtr_reduce_scatter(const void* const* xt, const void* const* rt, int F, int C, long long K,
                             float** shard_out, int* feats_per_rank_out, cudaStream_t s);

  // K13: GBDT histogram build fused with the reduce-scatter. bins: feature-major uint8 [F][pitch]
  // (256 bins), grad/hess fp32 [rows]; one pointer per local rank. shard_out[i] = this rank's
  // [fpr][256][2] fp32 (grad, hess) histogram shard of features [rank*fpr, ...), summed over ranks.
  void gbdt_hist_reduce_scatter(const uint8_t* const* bins, const long long* pitch, const float* const* grad,
                                const float* const* hess, const long long* rows, int F, float** shard_out,
                                int* feats_per_rank_out, cudaStream_t s);

  // Local reduce-by-key (the thread-level map merge of DOP:458-477 on device arrays, also usable
  // as pre-aggregation): out = unique keys of (keys, vals) with equal keys reduced by `op`.
  // *out_count (device u64) receives the full number of unique keys (it may exceed out_cap; at most out_cap
  // pairs are stored). Pairs with the reserved key INT64_MIN are ignored. The hash
  // table is plane-owned scratch (cudaMalloc, cached and grown on demand).
  void map_reduce_local(int local_rank, const long long* keys, const void* vals, long long n, int dtype, int op,
                        long long* out_keys, void* out_vals, long long out_cap, unsigned long long* out_count,
                        cudaStream_t s);

 private:
  GpuPlane() {}
  void finish_init(GpuOptions opt);
  int local_ranks() const { return emulated() ? world() : 1; }
  // True while `s` is being captured into a CUDA graph: host-side alternation of buffers is then
  // frozen into the graph, so paths that rely on it add the barrier they otherwise save.
  static bool capturing(cudaStream_t s) {
    cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
    return cudaStreamIsCapturing(s, &st) == cudaSuccess && st != cudaStreamCaptureStatusNone;
  }
  PeerTable table() const;
  struct Plan;
  // Generic driver: windows + staging + one GRS launch per window.
  void run_grs(const char* name, void* const* arrs, int dtype, int op, const int64_t* lo, const int64_t* hi,
               const unsigned* src_mask, const unsigned* dst_mask, bool reduce_like, int algo, cudaStream_t s);
  int pick_grid(long long max_units, int max_blocks_per_sm) const;
  // AUTO's NVLS rule, from measurement (profiles/): in-switch reduction only pays once it removes
  // traffic, i.e. beyond 2 ranks, and for messages large enough to amortise the multicast path.
  bool nvls_auto(size_t bytes, int family = 0) const {  // family: 0 allreduce, 1 reduce-scatter, 2 allgather
    size_t th = opt_.nvls_min_bytes;
    if (family == 1 && opt_.rs_nvls_min_bytes) th = opt_.rs_nvls_min_bytes;
    if (family == 2 && opt_.ag_nvls_min_bytes) th = opt_.ag_nvls_min_bytes;
    const size_t mx = family == 1 ? opt_.rs_nvls_max_bytes : (family == 2 ? opt_.ag_nvls_max_bytes : 0);
    if (mx && bytes >= mx) return false;
    return world() > opt_.nvls_min_world && bytes >= th;
  }
  bool offsets_symmetric(void* const* arrs, size_t* off) const;

  SymmHeap heap_;
  ProcessComm* pc_ = nullptr;  // real mode: for the rare host-side barriers (buffer re-creation)
  void host_barrier();
  GpuOptions opt_;
  int device_ = 0, sms_ = 0;
  uint32_t* err_host_ = nullptr;
  uint32_t* err_dev_ = nullptr;
  // heap layout (byte offsets)
  size_t oneshot_off_[2] = {0, 0};
  size_t oneshot_slot_bytes_ = 0;
  int oneshot_next_ = 0;
  size_t thread_off_ = 0, thread_bytes_ = 0;
  size_t staging_off_ = 0;
  size_t user_off_ = 0;
  // trivial first-fit allocator over [user_off_, heap bytes)
  std::mutex alloc_mu_;
  std::map<size_t, size_t> used_;  // off -> bytes
  CollStat last_;
  bool poisoned_ = false;  // a kernel hit the peer-wait watchdog
  bool in_relay_ = false;  // thread_segments is running one of the two launches of a relay
  int dead_rank_p1_ = 0;   // test hook, see GrsArgs::dead_rank_p1
  void entry_check();      // start of every collective: surfaces a timeout of an earlier (asynchronous) one
  std::string last_algo_store_;  // backing store when last_.algo is composed at run time
  CollTrace trace_;
  size_t lr_shard_off_ = 0, lr_shard_bytes_ = 0;
  size_t lr_recv_off_ = 0, lr_recv_bytes_ = 0;
  int lr_recv_next_ = 0;  // K12: store-forwarded contributions [src rank][fpr x C]
  float* lr_scratch_[kMaxRanks] = {};   // K12: local split-K merge buffers (per local rank)
  size_t lr_scratch_bytes_[kMaxRanks] = {};
  int* lr_tile_cnt_[kMaxRanks] = {};
  int lr_tiles_[kMaxRanks] = {};
  size_t hist_shard_off_ = 0, hist_shard_bytes_ = 0;
  float* hist_amax_[kMaxRanks] = {};
  struct HostPipe;
  HostPipe* hpipe_ = nullptr;
  void hpipe_release();
  struct MapWork;
  MapWork* map_ = nullptr;
  void map_release();
};

}  // namespace mp4j
