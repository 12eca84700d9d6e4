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
int thread_ws_slots();
int thread_ws_max_blocks();

}  // namespace mp4j
