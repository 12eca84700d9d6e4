// Host-side launch entry points of the kernel translation units.
#pragma once
#include <cuda_runtime.h>

#include "dense.cuh"

namespace mp4j {

// dense.cu
void launch_grs(int dtype, int op, bool nvls, const GrsArgs& args, int grid, cudaStream_t stream);
int grs_max_blocks_per_sm();
int grs_threads();

// thread_kernels.cu
// algo: 0 = register (each thread sums the T buffers in registers), 1 = cluster/DSMEM
void launch_thread_allreduce(int dtype, int op, bool nvls, const ThreadArgs& args, int grid, int algo,
                             cudaStream_t stream);
int thread_max_blocks_per_sm();
int thread_threads();
bool thread_cluster_supported(int dtype, int op, int T, int world);
bool thread_local_fast_supported(int dtype, int op, int T);  // algo 2: static-T, 256-bit accesses

// thread_fused_ws.cu: warp-specialised fused kernel (world > 1, 32-byte aligned, T in {2,4,8},
// SUM/MAX/MIN on 4/8-byte types). One CTA per SM; needs grid * slots * slot_bytes of staging.
void launch_thread_fused_ws(int dtype, int op, bool nvls, const ThreadArgs& args, int grid, cudaStream_t stream);
// thread_fused_tma.cu: the same pipeline with TMA bulk copies for the HBM streams (any T <= 64, 16-byte aligned)
void launch_thread_fused_tma(int dtype, int op, bool nvls, const ThreadArgs& args, int grid, cudaStream_t stream);
bool thread_fused_tma_supported(int dtype, int op, int T);
// thread_fused_pull.cu: pull formulation (no remote stores, no system-scope release); same preconditions as _ws
void launch_thread_fused_pull(int dtype, int op, bool nvls, const ThreadArgs& args, int grid, cudaStream_t stream);
int thread_ws_slots();
int thread_ws_max_blocks();

}  // namespace mp4j

namespace mp4j {
// lr_gemm.cu: one-block cross-rank barrier enqueued on a stream (between the phases of K12 / K13)
void launch_peer_barrier(const PeerTable& tab, cudaStream_t s);
}  // namespace mp4j

#include "map_kernels.cuh"

namespace mp4j {
// map_kernels.cu (type/op-erased launchers of the sparse-map kernels)
int map_threads();
void map_launch_partition(int dtype, const MapExchangeArgs& a, int grid, cudaStream_t s);
void map_launch_clear(int dtype, int op, const MapTable& t, int grid, cudaStream_t s);
// clear + insert + compact of a local array in one cooperative launch
void map_launch_local(int dtype, int op, const MapTable& t, const long long* keys, const void* vals, long long n,
                      long long* out_keys, void* out_vals, unsigned long long* count, unsigned long long out_cap,
                      int sms, cudaStream_t s);
void map_launch_merge_local(int dtype, int op, const MapTable& t, const long long* keys, const void* vals,
                            long long n, int grid, cudaStream_t s);
void map_launch_compact(int dtype, const MapTable& t, long long* out_keys, void* out_vals,
                        unsigned long long* count, unsigned long long out_cap, int grid, cudaStream_t s);
void map_launch_gather(int dtype, const PeerTable& tab, unsigned long long ocnt_off, unsigned long long okeys_off,
                       unsigned long long ovals_off, unsigned long long out_cap, const long long* const* own_keys,
                       const void* const* own_vals, const unsigned long long* const* own_count,
                       unsigned long long* const* info, unsigned long long overflow_off, int nlocal, bool publish,
                       bool push, bool use_mc, int grid, cudaStream_t s);
// clear + merge + compact + publish of a map collective in one cooperative kernel (t etc.: one entry per local rank)
void map_launch_table(int dtype, int op, const PeerTable& tab, const MapTable* t, long long* const* own_keys,
                      void* const* own_vals, unsigned long long* const* own_count, int nlocal,
                      unsigned long long own_cap, unsigned long long rkeys_off, unsigned long long rvals_off,
                      unsigned long long rcnt_off, unsigned long long ocnt_off, long long cap, int sms, cudaStream_t s);
}  // namespace mp4j
