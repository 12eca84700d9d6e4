// Custom device operators compiled at run time with NVRTC (custom_op.cu).
#pragma once
#include <cuda_runtime.h>

#include <string>

#include "dense.cuh"

namespace mp4j {

// Compiles `expr` (a C++ expression over `a`, `b` of the element type) for `dtype` and the plane's
// world size, loads it on the current device and returns its operator code (>= MP4J_CUSTOM_OP_BASE).
// The same (dtype, expr, world variant) returns the same code.
int custom_op_register(int dtype, const char* expr, int world);
// Compile only (no GPU needed): returns the lowered kernel name, *cubin_bytes = size of the SASS image.
std::string custom_op_compile_only(int dtype, const char* expr, int world, size_t* cubin_bytes);
bool is_custom_op(int op);
int custom_op_dtype(int op);
void launch_custom_grs(int op, const GrsArgs& args, int grid, cudaStream_t stream);
// dst[i] = op(dst[i], src[i]) for i in [0, n): the thread-level combine (DOP:441-457) for custom operators
void launch_custom_fold(int op, void* dst, const void* src, long long n, int sms, cudaStream_t stream);

}  // namespace mp4j
