// Softmax residual of multinomial logistic regression, fused and written TRANSPOSED: the producer of the
// operand the tcgen05 X^T R kernel (lr_gemm.cu, K12) wants.
//
//   logits [rows, C] bf16 (the forward GEMM's output), y [rows] int64 (label, < 0 = padding row)
//   Rt[c, r]   = softmax(logits[r, :])[c] - (y[r] == c)          bf16, [C, rows] (K-major for K12), 0 for padding
//   loss      += - sum_r log softmax(logits[r, :])[y[r]]          fp32 atomic
//
// One pass over the logits instead of log_softmax, exp, one-hot scatter, subtract, mask, gather, transpose and
// cast as separate tensor operations (the consumer is the reference's LR on ytk-learn, README.md:268-278).
// A block owns 32 consecutive rows: each warp takes 4 of them (lanes stride over the classes, shuffles for
// max / sum), the residuals go through a shared [C][33] tile so that the global writes are 64-byte runs of 32
// consecutive rows of one class.
#include <cuda_bf16.h>

#include "../capi/capi_util.h"
#include "cuda_check.h"

namespace mp4j {
namespace {

constexpr int kRows = 32;
constexpr int kThreads = 256;

__global__ void __launch_bounds__(kThreads) lr_residual_kernel(const __nv_bfloat16* __restrict__ logits,
                                                              const long long* __restrict__ y, long long rows, int C,
                                                              __nv_bfloat16* __restrict__ rt, float* __restrict__ loss) {
  extern __shared__ __nv_bfloat16 tile[];  // [C][kRows + 2]
  constexpr int kPitch = kRows + 2;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long r0 = (long long)blockIdx.x * kRows;
  float block_loss = 0.f;
  for (int i = 0; i < kRows / (kThreads / 32); ++i) {
    const int lr = warp * (kRows / (kThreads / 32)) + i;
    const long long r = r0 + lr;
    if (r >= rows) {
      for (int c = lane; c < C; c += 32) tile[c * kPitch + lr] = __float2bfloat16(0.f);
      continue;
    }
    const long long label = y[r];
    const __nv_bfloat16* row = logits + r * C;
    float m = -INFINITY;
    for (int c = lane; c < C; c += 32) m = fmaxf(m, __bfloat162float(row[c]));
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    float s = 0.f;
    for (int c = lane; c < C; c += 32) s += __expf(__bfloat162float(row[c]) - m);
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    const float inv = 1.f / s, lse = m + __logf(s);
    for (int c = lane; c < C; c += 32) {
      const float x = __bfloat162float(row[c]);
      float res = __expf(x - m) * inv - (c == label ? 1.f : 0.f);
      if (label < 0) res = 0.f;
      tile[c * kPitch + lr] = __float2bfloat16(res);
      if (c == label) block_loss += lse - x;
    }
  }
  __syncthreads();
  // transposed store: thread -> (class, row) with the row fastest
  for (int idx = threadIdx.x; idx < C * kRows; idx += kThreads) {
    const int c = idx / kRows, lr = idx % kRows;
    if (r0 + lr < rows) rt[(long long)c * rows + r0 + lr] = tile[c * kPitch + lr];
  }
  for (int o = 16; o > 0; o >>= 1) block_loss += __shfl_xor_sync(0xffffffffu, block_loss, o);
  if (lane == 0 && block_loss != 0.f) atomicAdd(loss, block_loss);
}

}  // namespace
}  // namespace mp4j

extern "C" int mp4j_gpu_lr_softmax_residual(const void* logits_bf16, const int64_t* y, int64_t rows, int C, void* rt_bf16,
                                            float* loss, void* stream) {
  using namespace mp4j;
  MP4J_API_BEGIN
  if (rows < 0 || C < 1 || C > 3072) throw Error("lr_softmax_residual: 1 <= classes <= 3072 (one [C][32] tile per block in shared memory)");
  if (rows == 0) return 0;
  const size_t smem = (size_t)C * (kRows + 2) * sizeof(__nv_bfloat16);
  static bool configured = false;
  if (!configured) {
    MP4J_CUDA(cudaFuncSetAttribute(lr_residual_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    configured = true;
  }
  const long long blocks = (rows + kRows - 1) / kRows;
  lr_residual_kernel<<<(unsigned)blocks, kThreads, smem, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)logits_bf16, (const long long*)y, rows, C, (__nv_bfloat16*)rt_bf16, loss);
  MP4J_CUDA(cudaGetLastError());
  MP4J_API_END
}
