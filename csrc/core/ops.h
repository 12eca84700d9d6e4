// Reduction functors shared by host collectives and CUDA kernels.
//
// Semantics pinned to the reference's predefined operators
// (J/operator/Operators.java:30-351) including the quirks of SURVEY §2.7:
//  * float/double MAX/MIN follow java.lang.Math.max/min: NaN-propagating, -0.0 < +0.0
//    (Operators.java:40,46,101,107) - CUDA fmax/fmin are NOT, so they are spelled out;
//  * short/byte SUM/PROD wrap on narrowing (Operators.java:258,298,308,348);
//  * packed LOC operators (Operators.java:56-88,168-200). Deviation, documented: ties are
//    broken by the LOWER location instead of "first argument wins", which makes the operator
//    commutative+associative so every rank/reduction order yields the same bits.
#pragma once
#ifndef __CUDACC_RTC__
#include <stdint.h>
#include <string.h>
#include "../../include/mp4j_types.h"
#endif  // NVRTC (custom device operators, csrc/gpu/custom_op.cu) gets these from a generated prelude

#if defined(__CUDACC__)
#define MP4J_HD __host__ __device__ __forceinline__
#else
#define MP4J_HD inline
#endif

namespace mp4j {

template <typename To, typename From>
MP4J_HD To bit_cast(const From& f) {
  static_assert(sizeof(To) == sizeof(From), "bit_cast size");
  To t;
  memcpy(&t, &f, sizeof(To));
  return t;
}

// ---- packed (value, loc) helpers: Operators.Double.compositeDouble / Operators.Long.compositeLong
MP4J_HD double composite_double(float val, int32_t loc) {
  uint64_t bits = (uint64_t)bit_cast<uint32_t>(val) << 32 | (uint32_t)loc;
  return bit_cast<double>(bits);
}
MP4J_HD float composite_double_val(double c) {
  return bit_cast<float>((uint32_t)(bit_cast<uint64_t>(c) >> 32));
}
MP4J_HD int32_t composite_double_loc(double c) { return (int32_t)(uint32_t)bit_cast<uint64_t>(c); }
MP4J_HD int64_t composite_long(int32_t val, int32_t loc) {
  return (int64_t)((uint64_t)(uint32_t)val << 32 | (uint32_t)loc);
}
MP4J_HD int32_t composite_long_val(int64_t c) { return (int32_t)((uint64_t)c >> 32); }
MP4J_HD int32_t composite_long_loc(int64_t c) { return (int32_t)(uint32_t)(uint64_t)c; }

template <typename T>
struct bits_of { typedef uint32_t type; };
template <> struct bits_of<double> { typedef uint64_t type; };

template <typename T>
struct is_fp { static constexpr bool value = false; };
template <> struct is_fp<float> { static constexpr bool value = true; };
template <> struct is_fp<double> { static constexpr bool value = true; };

// java.lang.Math.max for floating point: NaN wins, +0.0 > -0.0.
template <typename T>
MP4J_HD T java_max(T a, T b) {
  if (a != a) return a;
  if (b != b) return b;
  if (a == T(0) && b == T(0)) {  // distinguish signed zeros
    return (bit_cast<typename bits_of<T>::type>(a) >> (sizeof(T) * 8 - 1))
               ? b
               : a;
  }
  return a > b ? a : b;
}
template <typename T>
MP4J_HD T java_min(T a, T b) {
  if (a != a) return a;
  if (b != b) return b;
  if (a == T(0) && b == T(0)) {
    return (bit_cast<typename bits_of<T>::type>(a) >> (sizeof(T) * 8 - 1))
               ? a
               : b;
  }
  return a < b ? a : b;
}

template <typename T, int OP>
struct Reduce;

template <typename T> struct Reduce<T, MP4J_SUM> {
  static constexpr bool valid = true;
  MP4J_HD static T apply(T a, T b) { return (T)(a + b); }
};
template <typename T> struct Reduce<T, MP4J_PROD> {
  static constexpr bool valid = true;
  MP4J_HD static T apply(T a, T b) { return (T)(a * b); }
};
template <typename T> struct Reduce<T, MP4J_MAX> {
  static constexpr bool valid = true;
  MP4J_HD static T apply(T a, T b) {
    if constexpr (is_fp<T>::value) return java_max<T>(a, b);
    else return a > b ? a : b;
  }
};
template <typename T> struct Reduce<T, MP4J_MIN> {
  static constexpr bool valid = true;
  MP4J_HD static T apply(T a, T b) {
    if constexpr (is_fp<T>::value) return java_min<T>(a, b);
    else return a < b ? a : b;
  }
};
template <typename T> struct Reduce<T, MP4J_BAND> {
  static constexpr bool valid = !is_fp<T>::value;
  MP4J_HD static T apply(T a, T b) {
    if constexpr (is_fp<T>::value) return a; else return (T)(a & b);
  }
};
template <typename T> struct Reduce<T, MP4J_BOR> {
  static constexpr bool valid = !is_fp<T>::value;
  MP4J_HD static T apply(T a, T b) {
    if constexpr (is_fp<T>::value) return a; else return (T)(a | b);
  }
};
template <typename T> struct Reduce<T, MP4J_BXOR> {
  static constexpr bool valid = !is_fp<T>::value;
  MP4J_HD static T apply(T a, T b) {
    if constexpr (is_fp<T>::value) return a; else return (T)(a ^ b);
  }
};
// LOC operators exist only for the two 64-bit packed carriers.
template <typename T> struct Reduce<T, MP4J_MAXLOC> {
  static constexpr bool valid = false;
  MP4J_HD static T apply(T a, T) { return a; }
};
template <typename T> struct Reduce<T, MP4J_MINLOC> {
  static constexpr bool valid = false;
  MP4J_HD static T apply(T a, T) { return a; }
};
template <> struct Reduce<double, MP4J_MAXLOC> {
  static constexpr bool valid = true;
  MP4J_HD static double apply(double a, double b) {
    float va = composite_double_val(a), vb = composite_double_val(b);
    if (va > vb) return a;
    if (vb > va) return b;
    if (va != va) return a;  // NaN payloads: keep a deterministic side
    if (vb != vb) return b;
    return composite_double_loc(a) <= composite_double_loc(b) ? a : b;
  }
};
template <> struct Reduce<double, MP4J_MINLOC> {
  static constexpr bool valid = true;
  MP4J_HD static double apply(double a, double b) {
    float va = composite_double_val(a), vb = composite_double_val(b);
    if (va < vb) return a;
    if (vb < va) return b;
    if (va != va) return a;
    if (vb != vb) return b;
    return composite_double_loc(a) <= composite_double_loc(b) ? a : b;
  }
};
template <> struct Reduce<int64_t, MP4J_MAXLOC> {
  static constexpr bool valid = true;
  MP4J_HD static int64_t apply(int64_t a, int64_t b) {
    int32_t va = composite_long_val(a), vb = composite_long_val(b);
    if (va != vb) return va > vb ? a : b;
    return composite_long_loc(a) <= composite_long_loc(b) ? a : b;
  }
};
template <> struct Reduce<int64_t, MP4J_MINLOC> {
  static constexpr bool valid = true;
  MP4J_HD static int64_t apply(int64_t a, int64_t b) {
    int32_t va = composite_long_val(a), vb = composite_long_val(b);
    if (va != vb) return va < vb ? a : b;
    return composite_long_loc(a) <= composite_long_loc(b) ? a : b;
  }
};

inline size_t dtype_size(int dt) {
  switch (dt) {
    case MP4J_F64: case MP4J_I64: return 8;
    case MP4J_F32: case MP4J_I32: return 4;
    case MP4J_I16: case MP4J_BF16: case MP4J_F16: return 2;
    case MP4J_I8: return 1;
  }
  return 0;
}

inline bool op_valid_for(int dt, int op) {
  if (op == MP4J_COPY) return true;
  switch (op) {
    case MP4J_SUM: case MP4J_MAX: case MP4J_MIN: case MP4J_PROD: return true;
    case MP4J_BAND: case MP4J_BOR: case MP4J_BXOR:
      return dt == MP4J_I64 || dt == MP4J_I32 || dt == MP4J_I16 || dt == MP4J_I8;
    case MP4J_MAXLOC: case MP4J_MINLOC: return dt == MP4J_F64 || dt == MP4J_I64;
  }
  return false;
}

// Host-side dynamic dispatch: dst[i] = op(dst[i], src[i]) for i in [0,n).
// (Restates the recv-and-reduce hot loop DOP:187-205 without the per-element interface call.)
void host_reduce_into(int dtype, int op, void* dst, const void* src, size_t n);

}  // namespace mp4j
