// K13 - GBDT histogram build fused with the histogram reduce-scatter (consumer:
// ytk-learn GBDT, R/README.md:278, which reduce-scatters per-feature (grad, hess) histograms with
// reduceScatterArray, PCS:1174-1209).
//
//   hist[f][b] = sum over local rows with bins[f][row] == b of (grad[row], hess[row])
//   rank q ends up owning features [q*fpr, (q+1)*fpr) of the sum over ALL ranks
//
// bins are feature-major uint8 [F][rows] (256 bins). One CTA owns a block of kFB features and a row
// range: every thread streams 16 rows per step (one 16-byte load of bins per feature, grad/hess
// loaded once and reused for the kFB features), accumulates into a shared-memory histogram with
// shared atomics, and at the end ADDS its partial histogram straight into the owner rank's shard
// in the symmetric heap (red.global.add over NVLink). No local full histogram, no separate
// collective: the reduce-scatter happens in the flush.
#include <algorithm>

#include "cuda_check.h"
#include "gpu_plane.h"
#include "launch.h"

namespace mp4j {

namespace {

constexpr int kFB = 8;          // features per CTA (grad/hess reuse factor)
constexpr int kBins = 256;
constexpr int kHistThreads = 256;

struct HistArgs {
  PeerTable tab;
  unsigned long long out_off;   // heap offset of this rank's shard [fpr][256][2] fp32
  const uint8_t* bins[kMaxRanks];
  const float* grad[kMaxRanks];
  const float* hess[kMaxRanks];
  long long rows[kMaxRanks];
  long long pitch[kMaxRanks];   // bytes between features in bins (>= rows, multiple of 16)
  int F, fpr;
  long long rows_per_split;     // multiple of 16
  const float* amax[kMaxRanks]; // device float[2]: max |grad|, max |hess| of this rank's rows
};

// max |x| of grad and hess (non-negative floats order like their bit patterns)
__global__ void gbdt_amax_kernel(const float* grad, const float* hess, long long n, float* out) {
  float mg = 0.f, mh = 0.f;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    mg = fmaxf(mg, fabsf(grad[i]));
    mh = fmaxf(mh, fabsf(hess[i]));
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    mg = fmaxf(mg, __shfl_xor_sync(0xffffffffu, mg, o));
    mh = fmaxf(mh, __shfl_xor_sync(0xffffffffu, mh, o));
  }
  if ((threadIdx.x & 31) == 0) {
    atomicMax((unsigned*)&out[0], __float_as_uint(mg));
    atomicMax((unsigned*)&out[1], __float_as_uint(mh));
  }
}

// power-of-two quantisation step such that rows_per_split values of magnitude amax cannot overflow
// an int32 partial sum; integers (and dyadic fractions) above the step are represented exactly
__device__ __forceinline__ float quant_step(float amax, long long rows_per_split) {
  if (!(amax > 0.f)) return 1.f;
  const float qmax = 2147483000.f / (float)rows_per_split;
  return exp2f(ceilf(log2f(amax / qmax)));
}

__global__ void __launch_bounds__(kHistThreads) gbdt_hist_rs_kernel(const __grid_constant__ HistArgs a) {
  // Partial histograms are accumulated as fixed-point int32: shared-memory integer atomics are
  // native (ATOMS.ADD) while float ones compile to a CAS spin loop (ATOMS.CAST.SPIN), which made the
  // first version 4x slower (profiles/README.md §5). Sums of the quantised values are exact, so the
  // result does not depend on the order of the atomics.
  __shared__ int2 hist[kFB][kBins];
  const PeerTable& tab = a.tab;
  const int rank = tab.emulate ? (int)blockIdx.z : tab.rank;
  const int row = tab.emulate ? rank : 0;
  const int f0 = blockIdx.x * kFB;
  const long long rows = a.rows[row];
  const long long r0 = (long long)blockIdx.y * a.rows_per_split;
  const long long r1 = min(rows, r0 + a.rows_per_split);
  for (int i = threadIdx.x; i < kFB * kBins; i += blockDim.x) ((int2*)hist)[i] = make_int2(0, 0);
  __syncthreads();
  const uint8_t* bins = a.bins[row];
  const float* grad = a.grad[row];
  const float* hess = a.hess[row];
  const long long pitch = a.pitch[row];
  const float step_g = quant_step(a.amax[row][0], a.rows_per_split), step_h = quant_step(a.amax[row][1], a.rows_per_split);
  const float inv_g = 1.f / step_g, inv_h = 1.f / step_h;
  for (long long i = r0 + (long long)threadIdx.x * 16; i < r1; i += (long long)blockDim.x * 16) {
    float g[16], h[16];
    if (i + 16 <= rows) {
#pragma unroll
      for (int j = 0; j < 16; j += 4) {
        const float4 gv = *(const float4*)(grad + i + j);
        const float4 hv = *(const float4*)(hess + i + j);
        g[j] = gv.x; g[j + 1] = gv.y; g[j + 2] = gv.z; g[j + 3] = gv.w;
        h[j] = hv.x; h[j + 1] = hv.y; h[j + 2] = hv.z; h[j + 3] = hv.w;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        g[j] = (i + j < rows) ? grad[i + j] : 0.f;
        h[j] = (i + j < rows) ? hess[i + j] : 0.f;
      }
    }
    const int nvalid = (int)min(16ll, r1 - i);
    int qg[16], qh[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      qg[j] = __float2int_rn(g[j] * inv_g);
      qh[j] = __float2int_rn(h[j] * inv_h);
    }
#pragma unroll
    for (int ff = 0; ff < kFB; ++ff) {
      if (f0 + ff >= a.F) break;
      // pitch is a multiple of 16 and i too: the 16-byte load stays inside this feature's row
      const Vec16 bv = ld_stream(bins + (long long)(f0 + ff) * pitch + i);
      const uint32_t w[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        if (j < nvalid) {
          const int b = (w[j >> 2] >> ((j & 3) * 8)) & 0xff;
          atomicAdd(&hist[ff][b].x, qg[j]);
          atomicAdd(&hist[ff][b].y, qh[j]);
        }
      }
    }
  }
  __syncthreads();
  // flush = the reduce-scatter: add this CTA's partial histograms into the owner's shard
  for (int i = threadIdx.x; i < kFB * kBins; i += blockDim.x) {
    const int ff = i / kBins, b = i % kBins;
    const int f = f0 + ff;
    if (f >= a.F) break;
    const int2 q = hist[ff][b];
    if (q.x == 0 && q.y == 0) continue;
    const float2 v = make_float2((float)q.x * step_g, (float)q.y * step_h);
    const int owner = min(f / a.fpr, tab.world - 1);
    float* dst = (float*)(tab.heap[owner] + a.out_off) + ((size_t)(f - owner * a.fpr) * kBins + b) * 2;
    asm volatile("red.relaxed.sys.global.add.v2.f32 [%0], {%1, %2};" ::"l"(dst), "f"(v.x), "f"(v.y) : "memory");
  }
}

}  // namespace

void GpuPlane::gbdt_hist_reduce_scatter(const uint8_t* const* bins, const long long* pitch, const float* const* grad,
                                        const float* const* hess, const long long* rows, int F, float** shard_out,
                                        int* feats_per_rank_out, cudaStream_t s) {
  entry_check();
  const int p = world();
  const int L = local_ranks();
  if (F <= 0) throw Error("gbdt_hist_reduce_scatter: bad shape");
  const int fpr = (F + p - 1) / p;
  const size_t shard_bytes = (size_t)fpr * kBins * 2 * 4;
  if (hist_shard_bytes_ < shard_bytes) {
    MP4J_CUDA(cudaStreamSynchronize(s));
    if (hist_shard_bytes_) free(hist_shard_off_);
    hist_shard_off_ = alloc(shard_bytes);
    hist_shard_bytes_ = shard_bytes;
  }
  HistArgs a{};
  a.tab = table();
  a.out_off = hist_shard_off_;
  a.F = F;
  a.fpr = fpr;
  long long max_rows = 0;
  for (int i = 0; i < L; ++i) {
    if (rows[i] < 0 || pitch[i] < rows[i] || pitch[i] % 16 != 0 || ((uintptr_t)bins[i] & 15) ||
        ((uintptr_t)grad[i] & 15) || ((uintptr_t)hess[i] & 15))
      throw Error("gbdt_hist_reduce_scatter: bins need a 16-byte aligned base and pitch (>= rows); grad/hess 16-byte aligned");
    a.bins[i] = bins[i];
    a.grad[i] = grad[i];
    a.hess[i] = hess[i];
    a.rows[i] = rows[i];
    a.pitch[i] = pitch[i];
    max_rows = std::max(max_rows, rows[i]);
  }
  const int fblocks = (F + kFB - 1) / kFB;
  // enough row splits for ~4 waves of CTAs, each split a multiple of 16 rows and >= 4096 rows
  long long splits = std::max<long long>(1, (4ll * sms_ * 4 + fblocks - 1) / fblocks);
  splits = std::min<long long>(splits, std::max<long long>(1, max_rows / 4096));
  long long rps = (max_rows + splits - 1) / std::max<long long>(splits, 1);
  rps = std::max<long long>(16, (rps + 15) / 16 * 16);
  splits = std::max<long long>(1, (max_rows + rps - 1) / rps);
  a.rows_per_split = rps;
  for (int i = 0; i < L; ++i) {  // quantisation range of this call's gradients
    if (!hist_amax_[i]) MP4J_CUDA(cudaMalloc(&hist_amax_[i], 64));
    MP4J_CUDA(cudaMemsetAsync(hist_amax_[i], 0, 8, s));
    if (rows[i] > 0) gbdt_amax_kernel<<<std::max(1, std::min(sms_ * 2, (int)((rows[i] + 255) / 256))), 256, 0, s>>>(
        grad[i], hess[i], rows[i], hist_amax_[i]);
    a.amax[i] = hist_amax_[i];
  }
  for (int i = 0; i < L; ++i) {
    const int r = emulated() ? i : rank();
    MP4J_CUDA(cudaMemsetAsync(heap_.base(r) + hist_shard_off_, 0, shard_bytes, s));
    shard_out[i] = (float*)(heap_.base(r) + hist_shard_off_);
  }
  *feats_per_rank_out = fpr;
  if (p > 1) launch_peer_barrier(a.tab, s);  // every shard is zero before anyone adds into it
  gbdt_hist_rs_kernel<<<dim3(fblocks, (unsigned)splits, emulated() ? p : 1), kHistThreads, 0, s>>>(a);
  MP4J_CUDA(cudaGetLastError());
  if (p > 1) launch_peer_barrier(a.tab, s);  // every contribution landed before anyone reads its shard
  last_ = CollStat{"gbdt_hist_reduce_scatter", "smem-atomics+red.add", (size_t)F * max_rows, (int)(fblocks * splits),
                   false};
}

}  // namespace mp4j
