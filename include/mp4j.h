/* mp4j-b200 C ABI.
 *
 * One C++/CUDA core (libmp4j.so) is exposed through this header; the Python package binds it
 * with ctypes and the Java facade (java/) binds it through JNI. Every function returns
 * MP4J_OK (0) or MP4J_ERR (-1) unless stated; after an error mp4j_last_error() returns the
 * message for the calling thread. Element ranges are [from, to) in ELEMENTS.
 *
 * Parity citations use the shorthand of SURVEY.md: PCS = ProcessCommSlave.java,
 * TCS = ThreadCommSlave.java, SRV = rpc/Server.java.
 */
#ifndef MP4J_H_
#define MP4J_H_

#include <stddef.h>
#include <stdint.h>

#include "mp4j_types.h"

#ifdef __cplusplus
extern "C" {
#endif

const char* mp4j_last_error(void);
const char* mp4j_version(void);

/* ------------------------------------------------------------------ master (CommMaster.java:53-132) */
typedef struct mp4j_master mp4j_master;
/* timeouts <= 0 keep the reference defaults (7200 s connect, 600 s heartbeat, 3600 s exchange). */
int mp4j_master_create(int nslaves, int port, int connect_timeout_s, int heartbeat_timeout_s,
                       int exchange_timeout_s, int watchdog_period_ms, const char* log_dir,
                       const char* work_dir, int quiet, mp4j_master** out);
int mp4j_master_start(mp4j_master* m);
int mp4j_master_port(mp4j_master* m);
/* Blocks until the job ended; *code = 0 iff all slaves closed with 0. timeout_ms < 0 = forever.
 * Returns 1 if still running after timeout_ms. */
int mp4j_master_wait(mp4j_master* m, int timeout_ms, int* code);
int mp4j_master_shutdown(mp4j_master* m);
/* Copies the log sink (newline separated) into a malloc'd buffer the caller frees with mp4j_free. */
int mp4j_master_logs(mp4j_master* m, char** out, size_t* len);
void mp4j_master_destroy(mp4j_master* m);
void mp4j_free(void* p);

/* ------------------------------------------------------------------ process communicator (PCS) */
typedef struct mp4j_comm mp4j_comm;
/* rank_hint < 0: rank = index in the sorted "host###port" list (SRV:188,199-206). */
int mp4j_comm_create(const char* login, const char* master_host, int master_port, int rank_hint,
                     const char* my_host, mp4j_comm** out);
void mp4j_comm_destroy(mp4j_comm* c);
int mp4j_comm_rank(mp4j_comm* c);
int mp4j_comm_size(mp4j_comm* c);
/* "shm": the ranks share a machine and host arrays move through shared-memory windows (one
 * gather-reduce-scatter pass per collective); "tcp": ring / tree algorithms over sockets (MP4J_SHM=0). */
const char* mp4j_comm_host_transport(mp4j_comm* c);
int mp4j_comm_barrier(mp4j_comm* c);                                   /* PCS:432 */
int mp4j_comm_log(mp4j_comm* c, int level, const char* msg);            /* PCS:289-353 */
int mp4j_comm_close(mp4j_comm* c, int code);                            /* PCS:234 */
int mp4j_comm_exchange(mp4j_comm* c, int* other);                       /* SRV:358 */
int mp4j_comm_write_file(mp4j_comm* c, const char* content, const char* fname); /* PCS:279 */
int mp4j_comm_shutdown_master(mp4j_comm* c);
/* Rank-ordered gather of one blob per rank through the master mailbox. *out is one malloc'd
 * buffer: for each rank u64 length then bytes. */
int mp4j_comm_gather_all(mp4j_comm* c, const void* data, size_t n, char** out, size_t* out_len);

/* host collectives on primitive arrays in host memory (in place) */
int mp4j_host_reduce_scatter(mp4j_comm* c, void* arr, int dtype, int op, int64_t from,
                             const int64_t* counts);                    /* PCS:1174 */
int mp4j_host_allgather(mp4j_comm* c, void* arr, int dtype, const int64_t* froms,
                        const int64_t* tos);                            /* PCS:620 */
int mp4j_host_gather(mp4j_comm* c, void* arr, int dtype, const int64_t* froms, const int64_t* tos,
                     int root);                                         /* PCS:535 */
int mp4j_host_scatter(mp4j_comm* c, void* arr, int dtype, const int64_t* froms, const int64_t* tos,
                      int root);                                        /* PCS:900 */
int mp4j_host_broadcast(mp4j_comm* c, void* arr, int dtype, int64_t from, int64_t to, int root); /* PCS:750 */
int mp4j_host_reduce(mp4j_comm* c, void* arr, int dtype, int op, int64_t from, int64_t to,
                     int root);                                         /* PCS:1390 */
int mp4j_host_allreduce(mp4j_comm* c, void* arr, int dtype, int op, int64_t from, int64_t to); /* PCS:1733 */
int mp4j_host_allreduce_rpc(mp4j_comm* c, void* arr, int dtype, int op, int64_t n);            /* PCS:1776 */
/* dst[i] = op(dst[i], src[i]) on the host (the operand's local reduce, DOP:441-457) */
/* allreduceMap (mode 0) / reduceMap (mode 1, result on `root`) of maps with String keys and primitive
 * values (PCS:2053-2081, 1490-1516), natively: keys = n NUL-separated UTF-8 strings, vals = n elements of
 * dtype. *out_keys / *out_vals are malloc'd (mp4j_free) in the same layout, *out_n entries. Keys are placed
 * by String.hashCode() % slaveNum like the reference. */
int mp4j_host_map_reduce(mp4j_comm* c, int mode, int root, const char* keys, size_t keys_len, const void* vals,
                         size_t n, int dtype, int op, char** out_keys, size_t* out_keys_len, char** out_vals,
                         size_t* out_n);
/* reduceScatterMap (PCS:1299-1327) on (String, primitive) maps: the slaveNum maps of this rank back to back in keys /
 * vals, counts[q] entries in the q-th (the one addressed to rank q). Result (malloc'd, mp4j_free): this rank's map,
 * equal keys reduced with `op` during the native ring reduce-scatter. */
int mp4j_host_map_reduce_scatter(mp4j_comm* c, const char* keys, size_t keys_len, const void* vals, const int64_t* counts,
                                 int dtype, int op, char** out_keys, size_t* out_keys_len, char** out_vals,
                                 size_t* out_n);
int mp4j_host_reduce_into(int dtype, int op, void* dst, const void* src, int64_t n);

/* Collectives on opaque payloads (String / Object operands, ObjectOperand.java:114-156; maps with values the core
 * cannot reduce): the ring reduce-scatter / ring allgather / binomial gather / binomial scatter of the reference
 * (PCS:1329, 694, 440, 975) run natively; the binding supplies its codec.
 *   get(user, part, &data, &n): serialized bytes of local part `part`, valid until the next callback
 *   put(user, part, data, n)  : deserialize and merge into (reduce-scatter) / store as (others) part `part`
 * kind: MP4J_REDUCE_SCATTER (p parts per rank, result = part[rank]), MP4J_ALL_GATHER (part[rank] is this rank's,
 * put(q) delivers rank q's), MP4J_GATHER (put(q) on the root only), MP4J_SCATTER (root: get(q) for every q;
 * others: put(rank)). Callbacks return 0, or non-zero to abort the collective with an error. */
typedef int (*mp4j_blob_get)(void* user, int part, const char** data, size_t* n);
typedef int (*mp4j_blob_put)(void* user, int part, const char* data, size_t n);
int mp4j_host_blob_collective(mp4j_comm* c, int kind, int root, mp4j_blob_get get, mp4j_blob_put put, void* user);

/* byte-blob transport between slaves (String/Object/map payloads; *out freed with mp4j_free) */
int mp4j_comm_send_bytes(mp4j_comm* c, int dest, const void* data, size_t n);
int mp4j_comm_recv_bytes(mp4j_comm* c, int src, char** out, size_t* out_len);
int mp4j_comm_sendrecv_bytes(mp4j_comm* c, int dest, const void* data, size_t n, int src, char** out,
                             size_t* out_len);

/* ------------------------------------------------------------------ thread rendezvous (TCS:139-157) */
typedef struct mp4j_tgroup mp4j_tgroup;
int mp4j_tgroup_create(int nthreads, mp4j_tgroup** out);
void mp4j_tgroup_destroy(mp4j_tgroup* g);
int mp4j_tgroup_barrier(mp4j_tgroup* g);
int mp4j_tgroup_set_slot(mp4j_tgroup* g, int tid, int which, uint64_t v);
uint64_t mp4j_tgroup_get_slot(mp4j_tgroup* g, int tid, int which);
/* ThreadCommSlave.allreduceArray on host arrays (TCS:1915-1963): called by ALL threads of the group with
 * their own array. Each thread reduces its 1/T slice over the T arrays (thread-id order), thread 0 runs
 * the process allreduce, each thread copies the result into its array: both thread-level steps use T cores. */
int mp4j_thread_allreduce(mp4j_tgroup* g, mp4j_comm* c, int tid, void* arr, int dtype, int op, int64_t from,
                          int64_t to);
/* The other six ThreadCommSlave array collectives on host arrays (TCS:448-1430), called by ALL threads:
 * kind = mp4j_collective (GATHER, SCATTER, ALL_GATHER, BROADCAST, REDUCE_SCATTER, REDUCE). lo / hi: 2-D
 * ranges flattened [slaveNum][threadNum] (gather / scatter / allgather); counts likewise (reduce-scatter);
 * from / to for broadcast, reduce and the start of reduce-scatter. The thread-level combine and fan-out
 * run on all T threads in parallel; the delegate (root thread on the root process, thread 0 elsewhere)
 * runs the process collective. */
int mp4j_thread_collective(mp4j_tgroup* g, mp4j_comm* c, int tid, int kind, void* arr, int dtype, int op,
                           int64_t from, int64_t to, const int64_t* lo, const int64_t* hi, const int64_t* counts,
                           int root_rank, int root_tid);

/* ------------------------------------------------------------------ packed LOC helpers (Operators.java:75-88,187-200) */
double mp4j_composite_double(float val, int32_t loc);
float mp4j_composite_double_val(double c);
int32_t mp4j_composite_double_loc(double c);
int64_t mp4j_composite_long(int32_t val, int32_t loc);
int32_t mp4j_composite_long_val(int64_t c);
int32_t mp4j_composite_long_loc(int64_t c);
/* java.lang.String.hashCode of a UTF-16 code-unit sequence given as UTF-8 (PCS:2066-2070). */
int32_t mp4j_java_string_hash(const char* utf8, size_t n);
/* n NUL-separated UTF-8 keys -> 64-bit keys for the GPU map collectives: low word String.hashCode (placement as in
 * the reference, PCS:2063-2071), high word FNV-1a; never INT64_MIN (the kernels' free-slot marker). */
int mp4j_map_keys64(const char* blob, size_t blob_len, int64_t n, int64_t* out);

#include "mp4j_gpu.h"

#ifdef __cplusplus
}
#endif
#endif /* MP4J_H_ */
