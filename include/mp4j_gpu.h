/* GPU half of the C ABI. Included from mp4j.h inside extern "C".
 *
 * A mp4j_gpu is the GPU data plane of one communicator: symmetric heap mapped on every peer,
 * signal pads, staging windows and the sm_100a collective kernels. `arrs` arguments are arrays of
 * element-0 device pointers, one per LOCAL rank: 1 entry for a real communicator, `world` entries
 * for an emulated plane (all ranks on one GPU, used by single-GPU tests). `stream` is a
 * cudaStream_t passed as void* (NULL = default stream). All collectives are in place and
 * asynchronous with respect to the host.
 */
#ifndef MP4J_GPU_H_
#define MP4J_GPU_H_

typedef struct mp4j_gpu mp4j_gpu;

/* Collective over the communicator: every rank calls it with its own device. 0 sizes = defaults /
 * environment (MP4J_HEAP_BYTES, MP4J_STAGING_BYTES, ...). The plane is owned by the communicator. */
int mp4j_gpu_init(mp4j_comm* c, int device, size_t heap_bytes, size_t staging_bytes, mp4j_gpu** out);
/* Single-process plane with `world` emulated ranks on `device`. Free with mp4j_gpu_destroy. */
int mp4j_gpu_emu_create(int world, int device, size_t heap_bytes, size_t staging_bytes, mp4j_gpu** out);
void mp4j_gpu_destroy(mp4j_gpu* g);
int mp4j_gpu_world(mp4j_gpu* g);
int mp4j_gpu_rank(mp4j_gpu* g);
int mp4j_gpu_has_multicast(mp4j_gpu* g);
/* Human-readable description (mode, sizes, notes) and the last collective's record. */
int mp4j_gpu_describe(mp4j_gpu* g, char* buf, size_t n);
int mp4j_gpu_last(mp4j_gpu* g, char* buf, size_t n);
/* Symmetric allocation: the same offset is valid on every rank (collective, deterministic). */
int mp4j_gpu_alloc(mp4j_gpu* g, size_t bytes, size_t* off);
int mp4j_gpu_free(mp4j_gpu* g, size_t off);
void* mp4j_gpu_ptr(mp4j_gpu* g, int rank, size_t off);
/* Throws away nothing: returns MP4J_ERR if a kernel reported a peer-wait timeout. */
int mp4j_gpu_check(mp4j_gpu* g);
/* Run-time tuning knob for sweeps (the environment variables of INTEGRATION.md without the MP4J_ prefix, lower
 * case: thread_slots, thread_slot_bytes, ws_ver, nvls_min_world, nvls_min_bytes, oneshot_max_bytes, max_blocks,
 * poll_ns). Every rank must set the same value before the next collective. No counterpart in the reference. */
int mp4j_gpu_set_option(mp4j_gpu* g, const char* name, long long value);

int mp4j_gpu_allreduce(mp4j_gpu* g, void* const* arrs, int dtype, int op, int64_t from, int64_t to, int algo,
                       void* stream);                                                   /* PCS:1733 */
int mp4j_gpu_reduce_scatter(mp4j_gpu* g, void* const* arrs, int dtype, int op, int64_t from,
                            const int64_t* counts, int algo, void* stream);              /* PCS:1174 */
int mp4j_gpu_allgather(mp4j_gpu* g, void* const* arrs, int dtype, const int64_t* froms, const int64_t* tos,
                       int algo, void* stream);                                          /* PCS:620 */
int mp4j_gpu_broadcast(mp4j_gpu* g, void* const* arrs, int dtype, int64_t from, int64_t to, int root, int algo,
                       void* stream);                                                    /* PCS:750 */
int mp4j_gpu_gather(mp4j_gpu* g, void* const* arrs, int dtype, const int64_t* froms, const int64_t* tos,
                    int root, void* stream);                                             /* PCS:535 */
int mp4j_gpu_scatter(mp4j_gpu* g, void* const* arrs, int dtype, const int64_t* froms, const int64_t* tos,
                     int root, void* stream);                                            /* PCS:900 */
int mp4j_gpu_reduce(mp4j_gpu* g, void* const* arrs, int dtype, int op, int64_t from, int64_t to, int root,
                    int algo, void* stream);                                             /* PCS:1390 */
/* Per-collective statistics (the reference only logs wall-clock deltas in its checks).
 * With tracing on (this call, or MP4J_TRACE=1) every collective is bracketed by a CUDA event pair on
 * the caller's stream and an NVTX range; mp4j_gpu_stats resolves them (synchronising those events)
 * and writes a JSON array: one object per (collective, algorithm) with calls, bytes, total/min/max/avg
 * microseconds and algorithmic GB/s. reset != 0 clears the counters afterwards. */
int mp4j_gpu_trace(mp4j_gpu* g, int on);
int mp4j_gpu_stats(mp4j_gpu* g, char* buf, size_t n, int reset);
/* Custom device operator: the GPU form of implementing I<T>Operator.apply (J/operator/I*Operator.java:29-30).
 * `expr` is a C++ expression over `a` (accumulated so far, ascending rank / thread order) and `b` (next
 * value), both of the element type of `dtype` (one of the six primitive operands); e.g. "a + 2 * b",
 * "a > b ? a : b", "fmaxf(fabsf(a), fabsf(b))". It is compiled with NVRTC for sm_100a into the same
 * gather-reduce-scatter kernel as the predefined operators; *op_out (>= MP4J_CUSTOM_OP_BASE) is accepted by
 * mp4j_gpu_allreduce / _reduce_scatter / _reduce / _thread_allreduce of planes with this world size.
 * Every rank must register the same expression. Non-commutative expressions are evaluated in ascending
 * rank order on every rank, so results stay bit-identical everywhere. */
int mp4j_gpu_custom_op(mp4j_gpu* g, int dtype, const char* expr, int* op_out);
/* Compile only (needs NVRTC but no GPU): 0 if `expr` compiles for `dtype` at this world size; the
 * compiler's diagnostics are in mp4j_last_error() otherwise. *cubin_bytes = size of the sm_100a image. */
int mp4j_rtc_compile(int dtype, const char* expr, int world, long long* cubin_bytes);
/* fp32 SUM allreduce with a compressed wire format (mp4j_wire: BF16 or block-scaled FP8), fp32
 * accumulation. The GPU-side replacement of the reference's Deflate `compress` flag (DOP:267-269);
 * lossy, therefore never chosen implicitly. All ranks end with bit-identical results. */
int mp4j_gpu_allreduce_wire(mp4j_gpu* g, void* const* arrs, int64_t from, int64_t to, int wire, void* stream);

/* ThreadCommSlave.allreduceArray on device buffers (TCS:1915): bufs[local_rank * T + t].
 * fanout = 1 writes the result to all T buffers, 0 only to thread `dst`; local_only = 1 skips the
 * inter-GPU step (the thread-level reduce of TCS:259-303 alone). thread_algo: 0 register, 1
 * cluster/DSMEM (intra-GPU only). */
int mp4j_gpu_thread_allreduce(mp4j_gpu* g, void* const* bufs, int T, int dtype, int op, int64_t from,
                              int64_t to, int fanout, int dst, int local_only, int thread_algo, int algo,
                              void* stream);

/* ThreadCommSlave gatherArray / allgatherArray / broadcastArray / scatterArray / reduceScatterArray /
 * reduceArray on device buffers as ONE launch (TCS:448, 621, 802, 1010, 1174, 1396): bufs[local_rank * T + t],
 * segs = nseg x {lo, hi, src_rank, src_thread, dst_rank, dst_thread} (int64 each). A segment whose source is
 * (-1, -1) is the reduction, with `op`, of that element range over every thread of every rank (thread-id order,
 * then rank order); otherwise it is copied from that one thread. Destination -1 = every rank / every thread.
 * Segments must not overlap. At most 128 non-empty segments per call. name: "thread_gather" ... "thread_reduce"
 * (for mp4j_gpu_last / tracing). */
int mp4j_gpu_thread_segments(mp4j_gpu* g, void* const* bufs, int T, int dtype, int op, const int64_t* segs, int nseg,
                             const char* name, void* stream);

/* ThreadCommSlave.allreduceArray / ProcessCommSlave.allreduceArray (T = 1) on HOST arrays, streamed through
 * the GPU plane (TCS:1915, PCS:1733: the reference's arrays are host arrays): hbufs[local_rank * T + t] are host
 * pointers, pinned for full speed. The range is cut into chunks (MP4J_HOST_CHUNK_BYTES per buffer, default 8 MiB);
 * the copy-in of chunk c+1, the fused kernel of chunk c and the copy-out of chunk c-1 overlap on three streams.
 * fanout = 1 leaves the result in every thread's array, 0 only in thread `dst`'s. Returns when the host arrays
 * hold the result (work enqueued on `stream` before the call is waited for first). */
int mp4j_gpu_thread_allreduce_host(mp4j_gpu* g, void* const* hbufs, int T, int dtype, int op, int64_t from,
                                   int64_t to, int fanout, int dst, void* stream);

/* ThreadCommSlave collectives (TCS:448-1963) called the way the reference's are: by EVERY one of the T threads of a
 * process (threads of `tg`, ids 0..T-1), each with its own buffer. The rendezvous, the election of the launching
 * thread (thread 0; `launch_tid` for the segment form, the reference's root thread rule TCS:486) and the delivery of
 * its error to all T threads are native (mp4j::ThreadComm::delegate), so a binding needs no thread machinery of its
 * own. All threads must pass the same dtype / op / range / segment table; `stream` of the launching thread is used.
 * buf: device pointer (allreduce, segments) or host pointer, pinned for full speed (allreduce_host); a thread
 * whose own argument check failed passes NULL and every thread gets an error. Other arguments as in the
 * mp4j_gpu_thread_* forms above. */
int mp4j_thread_gpu_allreduce(mp4j_tgroup* tg, mp4j_gpu* g, int tid, void* buf, int dtype, int op, int64_t from,
                              int64_t to, int fanout, int dst, int local_only, int thread_algo, int algo,
                              void* stream);
int mp4j_thread_gpu_allreduce_host(mp4j_tgroup* tg, mp4j_gpu* g, int tid, void* hbuf, int dtype, int op, int64_t from,
                                   int64_t to, int fanout, int dst, void* stream);
int mp4j_thread_gpu_segments(mp4j_tgroup* tg, mp4j_gpu* g, int tid, int launch_tid, void* buf, int dtype, int op,
                             const int64_t* segs, int nseg, const char* name, void* stream);

/* Sparse (map container) collectives on (int64 key, value) device arrays: allreduceMap (mode 0,
 * PCS:2053), reduceScatterMap (mode 1, PCS:1299: every rank receives the keys it owns) and
 * allgatherMap (mode 2, PCS:666: concatenation in rank order). One entry per LOCAL rank in every
 * array argument. max_n = max over all ranks of n; out_cap = result capacity in pairs (same on all
 * ranks). partition: 0 = splitmix64(key) % p, 1 = Java String.hashCode placement on the low 32 bits.
 * Results stay in plane-owned device buffers until the next map call: out_keys/out_vals receive the
 * addresses; info[i] points to 4 pinned u64 {pairs, receive overflow, output overflow, owned pairs},
 * valid once `stream` has been synchronised. Key 0x8000000000000000 is reserved. */
int mp4j_gpu_map_collective(mp4j_gpu* g, int mode, const int64_t* const* keys, const void* const* vals,
                            const int64_t* n, int64_t max_n, int64_t out_cap, int dtype, int op, int partition,
                            int64_t** out_keys, void** out_vals, uint64_t** info, void* stream);

/* Logistic-regression gradient fused with its reduce-scatter (consumer of reduceScatterArray in
 * ytk-learn's distributed L-BFGS, reference README.md:268-278): G = sum over ranks of
 * Xt[F x K] * Rt[C x K]^T with bf16 K-major operands (K = local rows), fp32 result. The GEMM runs on
 * the tcgen05 tensor cores (TMA-fed, TMEM accumulator); every output tile is added straight into
 * the shard of the rank that owns its rows, over NVLink. C must be a multiple of 16 in [16, 256],
 * K a multiple of 8. shard_out[i] receives this rank's [feats_per_rank x C] fp32 shard (rows
 * [rank*fpr, ...) of G; plane-owned, valid until the next call). One pointer per LOCAL rank. */
int mp4j_gpu_lr_xtr_reduce_scatter(mp4j_gpu* g, const void* const* xt, const void* const* rt, int F, int C, int64_t K,
                                   float** shard_out, int* feats_per_rank, void* stream);

/* Softmax residual of multinomial LR, fused and transposed - the producer of K12's second operand (consumer:
 * the reference's LR, README.md:268-278). logits: bf16 [rows, C]; y: int64 labels (< 0 = padding row);
 * rt: bf16 [C, rows] receives softmax(logits) - onehot(y) (0 for padding rows); *loss (device fp32) is
 * INCREMENTED by the summed negative log-likelihood of the valid rows. Needs no plane. */
int mp4j_gpu_lr_softmax_residual(const void* logits_bf16, const int64_t* y, int64_t rows, int C, void* rt_bf16,
                                 float* loss, void* stream);

/* GBDT histogram build fused with its reduce-scatter (consumer of reduceScatterArray in ytk-learn's
 * GBDT, reference README.md:278). bins: feature-major uint8 [F][pitch] (256 bins per feature, pitch
 * >= rows and a multiple of 16, base 16-byte aligned), grad/hess: fp32 [rows]. Every CTA adds its
 * shared-memory partial histograms straight into the owner rank's shard over NVLink. shard_out[i]
 * receives this rank's [feats_per_rank][256][2] fp32 (sum grad, sum hess) shard, summed over all
 * ranks (plane-owned, valid until the next call). One pointer / size per LOCAL rank. */
int mp4j_gpu_gbdt_hist_reduce_scatter(mp4j_gpu* g, const uint8_t* const* bins, const int64_t* pitch,
                                      const float* const* grad, const float* const* hess, const int64_t* rows, int F,
                                      float** shard_out, int* feats_per_rank, void* stream);

/* Local reduce-by-key on one GPU: the thread-level map merge (DOP:458-477) on device arrays.
 * out_keys/out_vals (capacity out_cap pairs) receive the unique keys with equal keys reduced by
 * `op` (SUM/MAX/MIN; float/double/int/long values), *out_count (device uint64) their number.
 * *out_count is the FULL number of unique keys: if it exceeds out_cap only out_cap pairs were
 * stored (never more), so compare the two to detect truncation and never read beyond
 * min(*out_count, out_cap). Key INT64_MIN (0x8000000000000000) is reserved as the table's "free
 * slot" marker: pairs carrying it are ignored. The hash table is scratch owned by the plane
 * (allocated with cudaMalloc, cached, grown on demand). Result order is unspecified. */
int mp4j_gpu_map_reduce_local(mp4j_gpu* g, int local_rank, const int64_t* keys, const void* vals, int64_t n,
                              int dtype, int op, int64_t* out_keys, void* out_vals, int64_t out_cap,
                              uint64_t* out_count, void* stream);

/* Reduce-by-key without any device allocation. This entry point has NO counterpart in the
 * reference (ytk-mp4j is host-only Java and never manages device memory); it is an addition for
 * callers of this C ABI that cannot allocate on the GPU when they need the operation: code that
 * must not synchronise the device (cudaMalloc and cudaFree do), processes whose device memory
 * belongs entirely to another allocator or a fixed serving budget, and embedders that have no
 * mp4j_gpu plane whose cached scratch they could borrow.
 *
 * Same contract as mp4j_gpu_map_reduce_local (unique keys of (keys, vals) with equal keys reduced
 * by SUM/MAX/MIN; float/double/int/long values; result order unspecified; *out_count = the full
 * number of unique keys, which may exceed out_cap - at most out_cap pairs are stored, read no more
 * than min(*out_count, out_cap); pairs whose key is INT64_MIN, the reserved free-slot marker, are
 * ignored), on the CURRENT device, with these differences:
 *   - the hash table is a fixed 8 MiB array built into the library (held by every GPU that loads
 *     it until the process exits), so n may not exceed 262,144 pairs; larger n returns MP4J_ERR
 *     with a message naming the limit;
 *   - it calls neither cudaMalloc nor cudaFree nor any allocating library routine; inputs and
 *     outputs are the caller's;
 *   - the first call on each device also looks up the device address of that array
 *     (cudaGetSymbolAddress, checked). Make that first call before starting any stream capture;
 *   - all calls on a device share the array, which keeps the previous call's contents: every call
 *     re-initialises the slots it uses and zeroes *out_count first, and calls that overlap on the
 *     GPU (other streams or threads) corrupt each other - keep them on one stream or serialise. */
int mp4j_map_reduce_local_noalloc(const int64_t* keys, const void* vals, int64_t n, int dtype, int op,
                                  int64_t* out_keys, void* out_vals, int64_t out_cap, uint64_t* out_count,
                                  void* stream);

#endif /* MP4J_GPU_H_ */
