// sort_bins.cuh -- the arithmetic of Sort / DistributedSort that has to agree with the reference bit for bit: order-preserving
// sort keys and RangePartitionKernel::get_bin_pos (arrow/arrow_partition_kernels.cpp:436-440) in each column type's own C++
// arithmetic. Host-and-device so that tests/test_sort_bins_host.py can run the very same source on the CPU (g++) against the
// reference's results.
#pragma once

#include <stdint.h>
#include <string.h>

#include <type_traits>

#include "../../include/cylon_b200_types.h"

#ifdef __CUDACC__
#define CYLB_HD __host__ __device__
#else
#define CYLB_HD
#endif

namespace cylb {

enum KeyClass : int32_t { KC_UNSIGNED = 0, KC_SIGNED = 1, KC_FLOAT = 2 };

CYLB_HD inline int32_t key_class_of(int32_t type) {
  if (type == CYL_FLOAT || type == CYL_DOUBLE) return KC_FLOAT;
  switch (type) {
    case CYL_INT8: case CYL_INT16: case CYL_INT32: case CYL_INT64: case CYL_DATE32: case CYL_DATE64: case CYL_TIMESTAMP: case CYL_TIME32:
    case CYL_TIME64:
      return KC_SIGNED;
    default:
      return KC_UNSIGNED;
  }
}

// raw bits (zero-extended, `width` bytes) -> key whose unsigned order is the value order; stays inside `width` bytes
CYLB_HD inline uint64_t order_key(uint64_t bits, int32_t kclass, int32_t width) {
  const uint64_t top = 1ull << (8 * width - 1);
  const uint64_t all = width == 8 ? ~0ull : ((1ull << (8 * width)) - 1);
  if (kclass == KC_SIGNED) return bits ^ top;
  if (kclass == KC_FLOAT) return (bits & top) ? (~bits & all) : (bits | top);
  return bits;
}
CYLB_HD inline uint64_t order_key_inverse(uint64_t key, int32_t kclass, int32_t width) {
  const uint64_t top = 1ull << (8 * width - 1);
  const uint64_t all = width == 8 ? ~0ull : ((1ull << (8 * width)) - 1);
  if (kclass == KC_SIGNED) return key ^ top;
  if (kclass == KC_FLOAT) return (key & top) ? (key ^ top) : (~key & all);
  return key;
}

template <typename T>
struct RangeOf {
  T mn, mx, range;  // range = (T)(max - min), narrowed to the column type as the reference's member is
  uint32_t nb;
};

// get_bin_pos, arrow_partition_kernels.cpp:436-440:
//   (val >= min) + ((val >= min) & (val < max)) * ((val - min) * num_bins / range) + (val >= max) * num_bins
// with num_bins a uint32_t. Where the reference is undefined (range == 0, NaN, a quotient outside the table) the quotient counts as
// 0, NaN goes to the last bin and the position is clamped to the num_bins + 2 entries of bin_to_partition.
template <typename T>
CYLB_HD inline uint32_t bin_pos(T val, const RangeOf<T> &r) {
  const bool ge_min = val >= r.mn, ge_max = val >= r.mx, mask = ge_min && val < r.mx;
  if constexpr (std::is_floating_point<T>::value) {
    if (val != val) return r.nb + 1;
    T q = T(0);
    if (mask) q = (val - r.mn) * (T)r.nb / r.range;
    T total = (T)(ge_min ? 1 : 0) + q;
    total = total + (T)(ge_max ? r.nb : 0u);
    if (!(total >= T(0))) return 0;
    if (total >= (T)(r.nb + 1)) return r.nb + 1;
    return (uint32_t)total;  // truncation, as the conversion to size_t
  } else if constexpr (sizeof(T) <= 4) {
    // int / unsigned int arithmetic: the product with the uint32_t num_bins and the division are unsigned 32-bit
    uint32_t a = (uint32_t)((int64_t)val - (int64_t)r.mn);
    a *= r.nb;
    const uint32_t b = (uint32_t)r.range;  // sign-extends a narrowed negative range, as the usual arithmetic conversions do
    const uint32_t q = (mask && b) ? a / b : 0u;
    const uint32_t pos = (ge_min ? 1u : 0u) + q + (ge_max ? r.nb : 0u);
    return pos > r.nb + 1 ? r.nb + 1 : pos;
  } else if constexpr (std::is_signed<T>::value) {
    int64_t a = (int64_t)((uint64_t)val - (uint64_t)r.mn);
    a = (int64_t)((uint64_t)a * (uint64_t)r.nb);
    int64_t q = 0;
    if (mask && r.range != 0) q = r.range == -1 ? (int64_t)(0ull - (uint64_t)a) : a / (int64_t)r.range;
    const int64_t pos = (ge_min ? 1 : 0) + q + (ge_max ? (int64_t)r.nb : 0);
    return pos < 0 ? 0u : (pos > (int64_t)r.nb + 1 ? r.nb + 1 : (uint32_t)pos);
  } else {
    const uint64_t a = ((uint64_t)val - (uint64_t)r.mn) * (uint64_t)r.nb;
    const uint64_t q = (mask && r.range != 0) ? a / (uint64_t)r.range : 0ull;
    const uint64_t pos = (ge_min ? 1ull : 0ull) + q + (ge_max ? (uint64_t)r.nb : 0ull);
    return pos > (uint64_t)r.nb + 1 ? r.nb + 1 : (uint32_t)pos;
  }
}

template <typename T>
inline RangeOf<T> make_range(uint64_t min_key, uint64_t max_key, int32_t kclass, uint32_t num_bins) {
  RangeOf<T> r;
  const uint64_t lo = order_key_inverse(min_key, kclass, (int32_t)sizeof(T)), hi = order_key_inverse(max_key, kclass, (int32_t)sizeof(T));
  memcpy(&r.mn, &lo, sizeof(T));  // little endian: the low sizeof(T) bytes are the value
  memcpy(&r.mx, &hi, sizeof(T));
  if constexpr (std::is_floating_point<T>::value) {
    r.range = r.mx - r.mn;
  } else if constexpr (std::is_signed<T>::value) {
    r.range = (T)(uint64_t)((uint64_t)(int64_t)r.mx - (uint64_t)(int64_t)r.mn);  // wraps like the narrowing store in the reference
  } else {
    r.range = (T)((uint64_t)r.mx - (uint64_t)r.mn);
  }
  r.nb = num_bins;
  return r;
}

}  // namespace cylb
