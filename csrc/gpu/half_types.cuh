// 16-bit floating point element types for the dense GPU collectives (bf16 / fp16 tensors of torch
// users; the reference has no such operand). The kernels move them as opaque 16-bit scalars - scoped
// enums, so that unions, volatile accesses and zero-initialisation behave as for the integer types -
// and every reduction step converts to fp32, applies the operator and rounds back to nearest-even.
// A chain over p ranks therefore rounds p-1 times in ascending rank order: deterministic and
// identical on every rank. The NVLS path lets the switch accumulate in fp32 (.acc::f32).
#pragma once
#include <cuda_fp16.h>

#include "peer.cuh"

namespace mp4j {

enum class bf16_t : uint16_t {};
enum class f16_t : uint16_t {};

__device__ __forceinline__ float half_to_float(bf16_t x) { return __uint_as_float((uint32_t)(uint16_t)x << 16); }
__device__ __forceinline__ float half_to_float(f16_t x) { return __half2float(__ushort_as_half((uint16_t)x)); }

template <typename H>
__device__ __forceinline__ H float_to_half(float f);
template <>
__device__ __forceinline__ bf16_t float_to_half<bf16_t>(float f) {
  uint32_t u = __float_as_uint(f);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (bf16_t)(uint16_t)((u >> 16) | 0x0040u);  // quiet NaN
  u += 0x7fffu + ((u >> 16) & 1u);                                                      // round to nearest even
  return (bf16_t)(uint16_t)(u >> 16);
}
template <>
__device__ __forceinline__ f16_t float_to_half<f16_t>(float f) {
  return (f16_t)__half_as_ushort(__float2half_rn(f));
}

#define MP4J_HALF_REDUCE(H)                                                                      \
  template <>                                                                                    \
  struct Reduce<H, MP4J_SUM> {                                                                   \
    static constexpr bool valid = true;                                                          \
    __device__ __forceinline__ static H apply(H a, H b) {                                        \
      return float_to_half<H>(half_to_float(a) + half_to_float(b));                              \
    }                                                                                            \
  };                                                                                             \
  template <>                                                                                    \
  struct Reduce<H, MP4J_PROD> {                                                                  \
    static constexpr bool valid = true;                                                          \
    __device__ __forceinline__ static H apply(H a, H b) {                                        \
      return float_to_half<H>(half_to_float(a) * half_to_float(b));                              \
    }                                                                                            \
  };                                                                                             \
  template <>                                                                                    \
  struct Reduce<H, MP4J_MAX> {                                                                   \
    static constexpr bool valid = true;                                                          \
    __device__ __forceinline__ static H apply(H a, H b) {                                        \
      return float_to_half<H>(java_max<float>(half_to_float(a), half_to_float(b)));              \
    }                                                                                            \
  };                                                                                             \
  template <>                                                                                    \
  struct Reduce<H, MP4J_MIN> {                                                                   \
    static constexpr bool valid = true;                                                          \
    __device__ __forceinline__ static H apply(H a, H b) {                                        \
      return float_to_half<H>(java_min<float>(half_to_float(a), half_to_float(b)));              \
    }                                                                                            \
  };
MP4J_HALF_REDUCE(bf16_t)
MP4J_HALF_REDUCE(f16_t)
#undef MP4J_HALF_REDUCE

// in-switch sum with fp32 accumulation: 8 elements per 16-byte request
template <>
struct Nvls<bf16_t, MP4J_SUM> {
  static constexpr bool supported = true;
  __device__ __forceinline__ static Vec16 ld_reduce(const void* mc) {
    Vec16 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(mc)
                 : "memory");
    return v;
  }
};
template <>
struct Nvls<f16_t, MP4J_SUM> {
  static constexpr bool supported = true;
  __device__ __forceinline__ static Vec16 ld_reduce(const void* mc) {
    Vec16 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(mc)
                 : "memory");
    return v;
  }
};

}  // namespace mp4j
