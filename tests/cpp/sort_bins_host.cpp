// Host build of cylon_b200/csrc/sort_bins.cuh (the same source the CUDA kernels compile) for tests/test_sort_bins_host.py.
#include "../../cylon_b200/csrc/sort_bins.cuh"

using namespace cylb;

template <typename T>
static void run(const void *data, int64_t n, uint64_t min_key, uint64_t max_key, int32_t kclass, uint32_t nb, uint32_t *out) {
  const RangeOf<T> r = make_range<T>(min_key, max_key, kclass, nb);
  const T *v = static_cast<const T *>(data);
  for (int64_t i = 0; i < n; i++) out[i] = bin_pos<T>(v[i], r);
}

extern "C" int host_bin_pos(int32_t type, const void *data, int64_t n, uint64_t min_key, uint64_t max_key, uint32_t nb, uint32_t *out) {
  const int32_t kc = key_class_of(type);
  switch (type) {
    case CYL_UINT8: run<uint8_t>(data, n, min_key, max_key, kc, nb, out); break;
    case CYL_INT8: run<int8_t>(data, n, min_key, max_key, kc, nb, out); break;
    case CYL_UINT16: run<uint16_t>(data, n, min_key, max_key, kc, nb, out); break;
    case CYL_INT16: run<int16_t>(data, n, min_key, max_key, kc, nb, out); break;
    case CYL_UINT32: run<uint32_t>(data, n, min_key, max_key, kc, nb, out); break;
    case CYL_INT32: run<int32_t>(data, n, min_key, max_key, kc, nb, out); break;
    case CYL_UINT64: run<uint64_t>(data, n, min_key, max_key, kc, nb, out); break;
    case CYL_INT64: run<int64_t>(data, n, min_key, max_key, kc, nb, out); break;
    case CYL_FLOAT: run<float>(data, n, min_key, max_key, kc, nb, out); break;
    case CYL_DOUBLE: run<double>(data, n, min_key, max_key, kc, nb, out); break;
    default: return 1;
  }
  return 0;
}

// order-preserving key of raw bits and its inverse
extern "C" uint64_t host_order_key(int32_t type, int32_t width, uint64_t bits) { return order_key(bits, key_class_of(type), width); }
extern "C" uint64_t host_order_key_inverse(int32_t type, int32_t width, uint64_t key) { return order_key_inverse(key, key_class_of(type), width); }
