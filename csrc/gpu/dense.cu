// Dense array collectives on the GPU: ONE kernel shape, "gather-reduce-scatter" (GRS).
//
//   for every element i of this rank's work item:
//       acc  = reduce over ranks q in src_mask (ascending)  of  heap[q][src_off + i]
//       heap[q][dst_off + i] = acc   for ranks q in dst_mask      (+ optional local output)
//
// with a per-block cross-GPU barrier on entry (peers' inputs are ready) and on exit (all pushes
// landed, nobody still reads my buffer). Instances:
//   allreduce two-shot : item = balanced slice r, src = all,    dst = all         (PCS:1733-1763)
//   allreduce one-shot : item = whole range,     src = all,    dst = local only  (PCS:1776, small)
//   reduce-scatter     : item = slice r,         src = all,    dst = {r}         (PCS:1174)
//   allgather          : item = slice r,         src = {r},    dst = all         (PCS:620)
//   broadcast          : item = balanced slice r, src = {root}, dst = all        (PCS:750)
//   gather             : item = slice r,         src = {r},    dst = {root}      (PCS:535)
//   scatter            : item = slice r,         src = {root}, dst = {r}         (PCS:900)
//   reduce             : item = balanced slice r, src = all,    dst = {root}      (PCS:1390)
// The reference runs these as p-1 serial ring steps or lg p tree rounds over TCP, each followed by a
// master barrier RPC; on NVSwitch every GPU reaches every peer at full bandwidth, so all peers are
// pulled/pushed concurrently inside one launch. With NVLS the reduce is a multimem.ld_reduce (the
// switch adds) and the scatter a multimem.st (the switch replicates).
//
// The operator reduction and the dtype are template parameters: the collective, the Operator and
// the element type are fused in the same loop (S1-S9 of SURVEY §2.6).
#include <algorithm>

#include "cuda_check.h"
#include "custom_op.h"
#include "grs_kernel.cuh"
#include "half_types.cuh"
#include "launch.h"

namespace mp4j {

using namespace grs;

namespace {

template <typename T, int OP, bool NVLS, int W = 0>
void launch_one(const GrsArgs& args, int grid, cudaStream_t stream) {
  auto kfn = grs_kernel<T, OP, NVLS, W>;
  if (args.tab.emulate) {
    void* params[] = {(void*)&args};
    MP4J_CUDA(cudaLaunchCooperativeKernel((const void*)kfn, dim3(grid, args.tab.world), dim3(kThreads), params, 0,
                                          stream));
  } else {
    kfn<<<grid, kThreads, 0, stream>>>(args);
    MP4J_CUDA(cudaGetLastError());
  }
}

template <typename T, int OP>
void launch_nvls_or_not(const GrsArgs& args, bool nvls, int grid, cudaStream_t stream) {
  if constexpr (Nvls<T, OP>::supported || OP == MP4J_COPY) {
    if (nvls) return launch_one<T, OP, true>(args, grid, stream);
  }
  // compile-time world size for the operators and types that carry the bulk of the traffic
  if constexpr (sizeof(T) >= 4 && (OP == MP4J_SUM || OP == MP4J_MAX || OP == MP4J_MIN)) {
    switch (args.tab.world) {
      case 2: return launch_one<T, OP, false, 2>(args, grid, stream);
      case 4: return launch_one<T, OP, false, 4>(args, grid, stream);
      case 8: return launch_one<T, OP, false, 8>(args, grid, stream);
    }
  }
  launch_one<T, OP, false>(args, grid, stream);
}

template <typename T, bool INTEGRAL, bool LOC>
void by_op(int op, const GrsArgs& args, bool nvls, int grid, cudaStream_t stream) {
  switch (op) {
    case MP4J_SUM: return launch_nvls_or_not<T, MP4J_SUM>(args, nvls, grid, stream);
    case MP4J_MAX: return launch_nvls_or_not<T, MP4J_MAX>(args, nvls, grid, stream);
    case MP4J_MIN: return launch_nvls_or_not<T, MP4J_MIN>(args, nvls, grid, stream);
    case MP4J_PROD: return launch_nvls_or_not<T, MP4J_PROD>(args, false, grid, stream);
  }
  if constexpr (INTEGRAL) {
    switch (op) {
      case MP4J_BAND: return launch_nvls_or_not<T, MP4J_BAND>(args, nvls, grid, stream);
      case MP4J_BOR: return launch_nvls_or_not<T, MP4J_BOR>(args, nvls, grid, stream);
      case MP4J_BXOR: return launch_nvls_or_not<T, MP4J_BXOR>(args, nvls, grid, stream);
    }
  }
  if constexpr (LOC) {
    switch (op) {
      case MP4J_MAXLOC: return launch_one<T, MP4J_MAXLOC, false>(args, grid, stream);
      case MP4J_MINLOC: return launch_one<T, MP4J_MINLOC, false>(args, grid, stream);
    }
  }
  throw Error("operator " + std::to_string(op) + " is not defined for this dtype on the GPU path");
}

}  // namespace

int grs_max_blocks_per_sm() {
  static int cached = -1;
  if (cached < 0) {
    int nb = 0;
    MP4J_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, grs_kernel<float, MP4J_SUM, false, 0>, kThreads, 0));
    cached = std::max(nb, 1);
  }
  return cached;
}

int grs_threads() { return kThreads; }

void launch_grs(int dtype, int op, bool nvls, const GrsArgs& args, int grid, cudaStream_t stream) {
  if (op == MP4J_COPY) {  // ranges are in bytes
    if (nvls) return launch_one<int8_t, MP4J_COPY, true>(args, grid, stream);
    return launch_one<int8_t, MP4J_COPY, false>(args, grid, stream);
  }
  if (is_custom_op(op)) {  // run-time compiled instance of the same kernel template
    if (custom_op_dtype(op) != dtype) throw Error("custom operator was compiled for another element type");
    return launch_custom_grs(op, args, grid, stream);
  }
  switch (dtype) {
    case MP4J_F64: return by_op<double, false, true>(op, args, nvls, grid, stream);
    case MP4J_F32: return by_op<float, false, false>(op, args, nvls, grid, stream);
    case MP4J_I64: return by_op<int64_t, true, true>(op, args, nvls, grid, stream);
    case MP4J_I32: return by_op<int32_t, true, false>(op, args, nvls, grid, stream);
    case MP4J_I16: return by_op<int16_t, true, false>(op, args, false, grid, stream);
    case MP4J_I8: return by_op<int8_t, true, false>(op, args, false, grid, stream);
    case MP4J_BF16: return by_op<bf16_t, false, false>(op, args, nvls, grid, stream);  // half_types.cuh
    case MP4J_F16: return by_op<f16_t, false, false>(op, args, nvls, grid, stream);
  }
  throw Error("dtype " + std::to_string(dtype) + " is not supported on the GPU path");
}

}  // namespace mp4j
