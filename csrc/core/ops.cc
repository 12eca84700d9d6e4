// Host reduction dispatch (dtype x op) for the TCP data plane and the CPU oracle.
#include "ops.h"
#include "error.h"

namespace mp4j {
namespace {

inline float bf16_to_f32(uint16_t h) { return bit_cast<float>((uint32_t)h << 16); }
inline uint16_t f32_to_bf16(float f) {
  uint32_t u = bit_cast<uint32_t>(f);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);  // quiet NaN
  u += 0x7fffu + ((u >> 16) & 1);  // round to nearest even
  return (uint16_t)(u >> 16);
}
inline float f16_to_f32(uint16_t h) {
  uint32_t sign = (uint32_t)(h & 0x8000) << 16, exp = (h >> 10) & 0x1f, man = h & 0x3ff;
  uint32_t u;
  if (exp == 0) {
    if (man == 0) u = sign;
    else {
      int e = -1;
      do { man <<= 1; ++e; } while (!(man & 0x400));
      u = sign | ((uint32_t)(127 - 15 - e) << 23) | ((man & 0x3ff) << 13);
    }
  } else if (exp == 31) u = sign | 0x7f800000u | (man << 13);
  else u = sign | ((exp + 112) << 23) | (man << 13);
  return bit_cast<float>(u);
}
inline uint16_t f32_to_f16(float f) {
  uint32_t u = bit_cast<uint32_t>(f);
  uint32_t sign = (u >> 16) & 0x8000;
  int32_t exp = (int32_t)((u >> 23) & 0xff) - 127 + 15;
  uint32_t man = u & 0x7fffff;
  if (((u >> 23) & 0xff) == 0xff) return (uint16_t)(sign | 0x7c00 | (man ? 0x200 : 0));
  if (exp >= 31) return (uint16_t)(sign | 0x7c00);
  if (exp <= 0) {
    if (exp < -10) return (uint16_t)sign;
    man |= 0x800000;
    int shift = 14 - exp;
    uint32_t r = man >> shift;
    uint32_t rem = man & ((1u << shift) - 1), half = 1u << (shift - 1);
    if (rem > half || (rem == half && (r & 1))) ++r;
    return (uint16_t)(sign | r);
  }
  uint32_t r = (uint32_t)exp << 10 | man >> 13;
  uint32_t rem = man & 0x1fff;
  if (rem > 0x1000 || (rem == 0x1000 && (r & 1))) ++r;
  return (uint16_t)(sign | r);
}

template <typename T, int OP>
void reduce_loop(void* dst, const void* src, size_t n) {
  T* d = (T*)dst;
  const T* s = (const T*)src;
  for (size_t i = 0; i < n; ++i) d[i] = Reduce<T, OP>::apply(d[i], s[i]);
}

template <int OP, float (*LD)(uint16_t), uint16_t (*ST)(float)>
void reduce_loop_half(void* dst, const void* src, size_t n) {
  uint16_t* d = (uint16_t*)dst;
  const uint16_t* s = (const uint16_t*)src;
  for (size_t i = 0; i < n; ++i) d[i] = ST(Reduce<float, OP>::apply(LD(d[i]), LD(s[i])));
}

template <typename T>
void by_op(int op, void* dst, const void* src, size_t n) {
  switch (op) {
    case MP4J_SUM: return reduce_loop<T, MP4J_SUM>(dst, src, n);
    case MP4J_MAX: return reduce_loop<T, MP4J_MAX>(dst, src, n);
    case MP4J_MIN: return reduce_loop<T, MP4J_MIN>(dst, src, n);
    case MP4J_PROD: return reduce_loop<T, MP4J_PROD>(dst, src, n);
    case MP4J_BAND: return reduce_loop<T, MP4J_BAND>(dst, src, n);
    case MP4J_BOR: return reduce_loop<T, MP4J_BOR>(dst, src, n);
    case MP4J_BXOR: return reduce_loop<T, MP4J_BXOR>(dst, src, n);
    case MP4J_MAXLOC: return reduce_loop<T, MP4J_MAXLOC>(dst, src, n);
    case MP4J_MINLOC: return reduce_loop<T, MP4J_MINLOC>(dst, src, n);
  }
}

template <float (*LD)(uint16_t), uint16_t (*ST)(float)>
void by_op_half(int op, void* dst, const void* src, size_t n) {
  switch (op) {
    case MP4J_SUM: return reduce_loop_half<MP4J_SUM, LD, ST>(dst, src, n);
    case MP4J_MAX: return reduce_loop_half<MP4J_MAX, LD, ST>(dst, src, n);
    case MP4J_MIN: return reduce_loop_half<MP4J_MIN, LD, ST>(dst, src, n);
    case MP4J_PROD: return reduce_loop_half<MP4J_PROD, LD, ST>(dst, src, n);
  }
}

}  // namespace

void host_reduce_into(int dtype, int op, void* dst, const void* src, size_t n) {
  if (!op_valid_for(dtype, op) || op == MP4J_COPY)
    throw Error("operator " + std::to_string(op) + " is not defined for dtype " + std::to_string(dtype));
  switch (dtype) {
    case MP4J_F64: return by_op<double>(op, dst, src, n);
    case MP4J_F32: return by_op<float>(op, dst, src, n);
    case MP4J_I64: return by_op<int64_t>(op, dst, src, n);
    case MP4J_I32: return by_op<int32_t>(op, dst, src, n);
    case MP4J_I16: return by_op<int16_t>(op, dst, src, n);
    case MP4J_I8: return by_op<int8_t>(op, dst, src, n);
    case MP4J_BF16: return by_op_half<bf16_to_f32, f32_to_bf16>(op, dst, src, n);
    case MP4J_F16: return by_op_half<f16_to_f32, f32_to_f16>(op, dst, src, n);
  }
  throw Error("unknown dtype");
}

}  // namespace mp4j
