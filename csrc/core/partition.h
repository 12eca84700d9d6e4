// Range validators and balanced partitioners (parity: J/utils/CommUtils.java:34-222).
// Rule: avg = size / n; participants 0..n-2 get avg, the last takes the rest.
// Indices are int64 here (the reference is int-indexed, <= 2^31-1 elements).
#pragma once
#include <stdint.h>
#include <vector>
#include "error.h"

namespace mp4j {

inline void check_froms_tos(const int64_t* froms, const int64_t* tos, int n) {
  for (int i = 0; i < n; ++i) {
    if (froms[i] < 0) throw Error("froms[i] must be >= 0!");
    if (tos[i] < 0) throw Error("tos[i] must be >= 0!");
    if (froms[i] > tos[i]) throw Error("froms[i] must be <= tos[i]!");
    if (i >= 1 && froms[i] < tos[i - 1]) throw Error("froms[i] must be >= to[i - 1]");
  }
}

inline void check_from_counts(int64_t from, const int64_t* counts, int n) {
  if (from < 0) throw Error("from must be >= 0!");
  for (int i = 0; i < n; ++i)
    if (counts[i] < 0) throw Error("counts[i] must be >= 0!");
}

inline void check_from_to(int64_t from, int64_t to) {
  if (from < 0) throw Error("from must be >= 0!");
  if (to < 0) throw Error("to must be >= 0!");
  if (from > to) throw Error("from must be <= to!");
}

// [from, to) split over n participants, reference rule. Fills froms[n], tos[n].
inline void balanced_split(int64_t from, int64_t to, int n, int64_t* froms, int64_t* tos) {
  int64_t avg = (to - from) / n;
  for (int r = 0; r < n; ++r) {
    froms[r] = from + r * avg;
    tos[r] = (r == n - 1) ? to : from + (r + 1) * avg;
  }
}

inline void froms_tos_from_counts(int64_t from, const int64_t* counts, int n, int64_t* froms,
                                  int64_t* tos) {
  for (int i = 0; i < n; ++i) {
    froms[i] = from;
    from += counts[i];
    tos[i] = from;
  }
}

}  // namespace mp4j
