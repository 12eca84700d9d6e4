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

}  // namespace mp4j
