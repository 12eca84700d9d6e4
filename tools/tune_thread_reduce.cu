// Stand-alone tuner for the intra-GPU T-way reduce + fan-out (the N=1 flagship shape):
//   out[t][i] = sum_t in[t][i]  for T = 8 buffers of 64 Mi fp32, in place.
// Sweeps block size / blocks per SM / vectors per thread / cache policies and prints GB/s of
// algorithmic HBM traffic (read T*S + write T*S). Build + run (GPU box):
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo tools/tune_thread_reduce.cu -o /tmp/tune && /tmp/tune
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <algorithm>
#include <vector>

#define CK(x)                                                                    \
  do {                                                                           \
    cudaError_t e = (x);                                                         \
    if (e != cudaSuccess) {                                                      \
      printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); \
      exit(1);                                                                   \
    }                                                                            \
  } while (0)

struct Ptrs {
  float* p[8];
};

template <int LD>
__device__ __forceinline__ float4 ld16(const float* p) {
  float4 v;
  if (LD == 0) asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  if (LD == 1) asm volatile("ld.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  if (LD == 2) asm volatile("ld.global.cs.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
template <int ST>
__device__ __forceinline__ void st16(float* p, float4 v) {
  if (ST == 0) asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
  if (ST == 1) asm volatile("st.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
  if (ST == 2) asm volatile("st.global.cs.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

template <int NT, int U, int THREADS, int MINB, int LD, int ST>
__global__ void __launch_bounds__(THREADS, MINB) k(const __grid_constant__ Ptrs P, long long nvec) {
  const long long stride = (long long)gridDim.x * THREADS;
  long long v = (long long)blockIdx.x * THREADS + threadIdx.x;
  for (; v + (U - 1) * stride < nvec; v += U * stride) {
    float4 x[U][NT];
#pragma unroll
    for (int u = 0; u < U; ++u)
#pragma unroll
      for (int t = 0; t < NT; ++t) x[u][t] = ld16<LD>(P.p[t] + (v + u * stride) * 4);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      float4 a = x[u][0];
#pragma unroll
      for (int t = 1; t < NT; ++t) {
        a.x += x[u][t].x; a.y += x[u][t].y; a.z += x[u][t].z; a.w += x[u][t].w;
      }
#pragma unroll
      for (int t = 0; t < NT; ++t) st16<ST>(P.p[t] + (v + u * stride) * 4, a);
    }
  }
  for (; v < nvec; v += stride) {
    float4 a = ld16<LD>(P.p[0] + v * 4);
#pragma unroll
    for (int t = 1; t < NT; ++t) {
      float4 b = ld16<LD>(P.p[t] + v * 4);
      a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
    }
#pragma unroll
    for (int t = 0; t < NT; ++t) st16<ST>(P.p[t] + v * 4, a);
  }
}

struct F8 {
  float v[8];
};
template <int EF>
__device__ __forceinline__ F8 ld32(const float* p) {
  F8 r;
  if (EF)
    asm volatile("ld.global.nc.L1::no_allocate.L2::evict_first.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(r.v[0]), "=f"(r.v[1]), "=f"(r.v[2]), "=f"(r.v[3]), "=f"(r.v[4]), "=f"(r.v[5]), "=f"(r.v[6]), "=f"(r.v[7])
                 : "l"(p));
  else
    asm volatile("ld.global.nc.L1::no_allocate.L2::evict_normal.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(r.v[0]), "=f"(r.v[1]), "=f"(r.v[2]), "=f"(r.v[3]), "=f"(r.v[4]), "=f"(r.v[5]), "=f"(r.v[6]), "=f"(r.v[7])
                 : "l"(p));
  return r;
}
template <int EF>
__device__ __forceinline__ void st32(float* p, const F8& r) {
  if (EF)
    asm volatile("st.global.L1::no_allocate.L2::evict_first.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "f"(r.v[0]),
                 "f"(r.v[1]), "f"(r.v[2]), "f"(r.v[3]), "f"(r.v[4]), "f"(r.v[5]), "f"(r.v[6]), "f"(r.v[7])
                 : "memory");
  else
    asm volatile("st.global.L1::no_allocate.L2::evict_normal.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "f"(r.v[0]),
                 "f"(r.v[1]), "f"(r.v[2]), "f"(r.v[3]), "f"(r.v[4]), "f"(r.v[5]), "f"(r.v[6]), "f"(r.v[7])
                 : "memory");
}

template <int NT, int THREADS, int MINB, int EF>
__global__ void __launch_bounds__(THREADS, MINB) k256(const __grid_constant__ Ptrs P, long long nvec32) {
  const long long stride = (long long)gridDim.x * THREADS;
  for (long long v = (long long)blockIdx.x * THREADS + threadIdx.x; v < nvec32; v += stride) {
    F8 x[NT];
#pragma unroll
    for (int t = 0; t < NT; ++t) x[t] = ld32<EF>(P.p[t] + v * 8);
#pragma unroll
    for (int t = 1; t < NT; ++t)
#pragma unroll
      for (int i = 0; i < 8; ++i) x[0].v[i] += x[t].v[i];
#pragma unroll
    for (int t = 0; t < NT; ++t) st32<EF>(P.p[t] + v * 8, x[0]);
  }
}

static int g_sms = 148;
static Ptrs g_p;
static long long g_nvec;
static cudaEvent_t e0, e1;

template <int NT, int U, int THREADS, int MINB, int LD, int ST>
void run(const char* tag) {
  auto fn = k<NT, U, THREADS, MINB, LD, ST>;
  int occ = 0;
  CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fn, THREADS, 0));
  cudaFuncAttributes fa;
  CK(cudaFuncGetAttributes(&fa, fn));
  for (int mult : {1, 2, 4}) {
    int grid = g_sms * occ * mult;
    float best = 1e9;
    for (int it = 0; it < 6; ++it) {
      CK(cudaEventRecord(e0));
      fn<<<grid, THREADS>>>(g_p, g_nvec);
      CK(cudaEventRecord(e1));
      CK(cudaEventSynchronize(e1));
      float ms;
      CK(cudaEventElapsedTime(&ms, e0, e1));
      if (it >= 2) best = std::min(best, ms);
    }
    CK(cudaGetLastError());
    double bytes = 2.0 * NT * g_nvec * 16;
    printf("%-10s NT=%d U=%d thr=%d minb=%d occ=%d regs=%d spill=%zu ld=%d st=%d grid=%5d  %.4f ms  %.1f GB/s\n", tag, NT, U,
           THREADS, MINB, occ, fa.numRegs, (size_t)fa.localSizeBytes, LD, ST, grid, best, bytes / best / 1e6);
  }
}

template <int NT, int THREADS, int MINB, int EF>
void run256(const char* tag) {
  auto fn = k256<NT, THREADS, MINB, EF>;
  int occ = 0;
  CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fn, THREADS, 0));
  cudaFuncAttributes fa;
  CK(cudaFuncGetAttributes(&fa, fn));
  for (int mult : {1, 2, 4}) {
    int grid = g_sms * occ * mult;
    float best = 1e9;
    for (int it = 0; it < 6; ++it) {
      CK(cudaEventRecord(e0));
      fn<<<grid, THREADS>>>(g_p, g_nvec / 2);
      CK(cudaEventRecord(e1));
      CK(cudaEventSynchronize(e1));
      float ms;
      CK(cudaEventElapsedTime(&ms, e0, e1));
      if (it >= 2) best = std::min(best, ms);
    }
    CK(cudaGetLastError());
    double bytes = 2.0 * NT * g_nvec * 16;
    printf("%-10s NT=%d 256-bit thr=%d minb=%d occ=%d regs=%d spill=%zu ef=%d grid=%5d  %.4f ms  %.1f GB/s\n", tag, NT, THREADS,
           MINB, occ, fa.numRegs, (size_t)fa.localSizeBytes, EF, grid, best, bytes / best / 1e6);
  }
}

int main() {
  int dev = 0;
  CK(cudaSetDevice(dev));
  CK(cudaDeviceGetAttribute(&g_sms, cudaDevAttrMultiProcessorCount, dev));
  const long long n = 64ll << 20;
  g_nvec = n / 4;
  for (int t = 0; t < 8; ++t) {
    CK(cudaMalloc(&g_p.p[t], n * 4));
    CK(cudaMemset(g_p.p[t], 0, n * 4));
  }
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  // reference point: plain device-to-device copy of the same total traffic
  {
    float best = 1e9;
    for (int it = 0; it < 5; ++it) {
      CK(cudaEventRecord(e0));
      for (int t = 0; t < 8; t += 2) CK(cudaMemcpyAsync(g_p.p[t], g_p.p[t + 1], n * 4, cudaMemcpyDeviceToDevice));
      CK(cudaEventRecord(e1));
      CK(cudaEventSynchronize(e1));
      float ms;
      CK(cudaEventElapsedTime(&ms, e0, e1));
      if (it >= 1) best = std::min(best, ms);
    }
    printf("cudaMemcpy D2D 4 x 256MiB: %.4f ms  %.1f GB/s (read+write)\n", best, 8.0 * n * 4 / best / 1e6);
  }
  run<8, 1, 256, 2, 0, 0>("base");
  run<8, 1, 256, 4, 0, 0>("occ4");
  run<8, 1, 512, 2, 0, 0>("t512");
  run<8, 1, 128, 8, 0, 0>("t128");
  run<8, 2, 256, 2, 0, 0>("u2");
  run<8, 2, 256, 1, 0, 0>("u2b1");
  run<8, 4, 256, 1, 0, 0>("u4b1");
  run<8, 1, 1024, 1, 0, 0>("t1024");
  run<8, 1, 256, 2, 1, 1>("plain");
  run<8, 1, 256, 2, 2, 2>("cs");
  run256<8, 256, 2, 1>("v8ef");
  run256<8, 256, 2, 0>("v8");
  run256<8, 128, 4, 1>("v8t128");
  run256<8, 512, 1, 1>("v8t512");
  run<8, 1, 256, 2, 0, 2>("nc+cs");
  run<8, 1, 256, 2, 0, 1>("nc+plain");
  run<8, 1, 256, 4, 0, 2>("occ4+cs");
  run<8, 2, 256, 2, 0, 2>("u2+cs");
  run<4, 2, 256, 2, 0, 0>("T4u2");
  run<2, 4, 256, 2, 0, 0>("T2u4");
  return 0;
}
