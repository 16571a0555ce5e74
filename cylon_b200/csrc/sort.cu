// sort.cu -- local Sort (argsort by one or more columns) and the range partitioner of DistributedSort.
//
// Replaces (reference, relative to cpp/src/cylon/):
//   * Sort, table.cpp:291-328 -> util::SortTable / SortTableMultiColumns (util/arrow_utils.cpp:30-118) -> SortIndices /
//     SortIndicesMultiColumns (arrow/arrow_kernels.cpp:294-303,412-457: std::sort of row indices through a comparator chain)
//     followed by arrow Take of every column;
//   * MapToSortPartitions, partition/partition.cpp:168-198 -> RangePartitionKernel, arrow/arrow_partition_kernels.cpp:334-455
//     (sample -> min/max over all ranks -> histogram of the sample over num_bins + 2 bins -> all-reduce -> quantile split of
//     the bins into partitions -> bin of every row).
//
// The argsort is the stable 8-bit LSD radix sort of join_sort.cu over order-preserving 64-bit keys, one column at a time from
// the least significant sort column to the most significant (stability composes the columns); descending = complemented key.
// The range partitioner evaluates the reference's bin expression in the column's own C++ type -- (val - min) * num_bins / range
// promotes differently for 8/16/32-bit, 64-bit and floating-point columns and the per-rank result depends on it
// (sort_bins.cuh; tests/test_sort_bins_host.py runs that header on the host).
#include <type_traits>

#include "common.cuh"
#include "sort_bins.cuh"

namespace cylb {

int radix_sort_pairs(uint64_t **tags_io, uint32_t **ids_io, uint64_t **tags_tmp, uint32_t **ids_tmp, int64_t n, cudaStream_t s);

namespace {

inline int sort_grid(int64_t n) {
  int64_t b = div_up(n, 256);
  const int64_t cap = (int64_t)num_sms() * 16;
  return (int)(b < 1 ? 1 : (b < cap ? b : cap));
}

// tags[i] = key of the column's value in row ids[i] (or row i when ids is null; then the identity permutation is written too)
__global__ void __launch_bounds__(256) sort_tags_kernel(const DCol c, const uint32_t *__restrict__ ids, int64_t n, int32_t kclass, int32_t descending,
                                                        uint64_t *__restrict__ tags, uint32_t *__restrict__ ids_out) {
  const uint64_t all = c.width == 8 ? ~0ull : ((1ull << (8 * c.width)) - 1);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t row = ids ? (int64_t)ids[i] : i;
    uint64_t k = order_key(load_bits(c, row), kclass, c.width);
    if (descending) k ^= all;
    tags[i] = k;
    if (!ids) ids_out[i] = (uint32_t)i;
  }
}

__global__ void __launch_bounds__(256) widen_ids_kernel(const uint32_t *__restrict__ ids, int64_t n, int64_t *__restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) out[i] = (int64_t)ids[i];
}

bool sortable(const DCol &c) { return c.kind != HK_MURMUR_VAR && c.type != CYL_FIXED_SIZE_BINARY && c.width >= 1 && c.width <= 8; }

}  // namespace
}  // namespace cylb

using namespace cylb;

extern "C" int cyl_sort_indices(const cyl_column_t *keys, int32_t nkeys, const int32_t *ascending, int64_t nrows, int64_t *out_idx, void *stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (nkeys < 1 || !keys || !ascending) return fail(CYL_Invalid, "cyl_sort_indices: needs at least one sort column and its direction");
  if (nrows < 0 || nrows > 0xFFFFFFFEll) return fail(CYL_Invalid, "cyl_sort_indices: row count %lld out of range", (long long)nrows);
  if (nrows > 0 && !out_idx) return fail(CYL_Invalid, "cyl_sort_indices: null output");
  DCols k;
  CYL_TRY(make_dcols(keys, nkeys, nrows, &k, "cyl_sort_indices"));
  for (int c = 0; c < nkeys; c++)
    if (!sortable(k.c[c])) return fail(CYL_NotImplemented, "cyl_sort_indices: sort column %d is not a fixed-width value of at most 8 bytes", c);
  if (nrows == 0) return CYL_OK;
  Scratch ta, tb, ia, ib;
  CYL_TRY(ta.alloc((size_t)(nrows + 1) * 8, s));
  CYL_TRY(tb.alloc((size_t)(nrows + 1) * 8, s));
  CYL_TRY(ia.alloc((size_t)(nrows + 1) * 4, s));
  CYL_TRY(ib.alloc((size_t)(nrows + 1) * 4, s));
  uint64_t *tags = ta.as<uint64_t>(), *tags_tmp = tb.as<uint64_t>();
  uint32_t *ids = ia.as<uint32_t>(), *ids_tmp = ib.as<uint32_t>();
  ProfScope ps("sort_indices", s);
  for (int c = nkeys - 1; c >= 0; c--) {  // least significant sort column first; every pass is stable
    const bool first = c == nkeys - 1;
    sort_tags_kernel<<<sort_grid(nrows), 256, 0, s>>>(k.c[c], first ? nullptr : ids, nrows, key_class_of(k.c[c].type), ascending[c] ? 0 : 1, tags, ids);
    CYL_LAUNCH_CHECK();
    CYL_TRY(radix_sort_pairs(&tags, &ids, &tags_tmp, &ids_tmp, nrows, s));
  }
  widen_ids_kernel<<<sort_grid(nrows), 256, 0, s>>>(ids, nrows, out_idx);
  CYL_LAUNCH_CHECK();
  return CYL_OK;
}

// ---------------------------------------------------------------------------------------------------------------------------
// range partitioning
// ---------------------------------------------------------------------------------------------------------------------------
namespace cylb {
namespace {

// Rows the partitioner looks at: every row when num_samples >= n, else rows floor(i * n / num_samples). The reference draws a
// random sample (util::SampleArray) unless num_samples == n; a fixed rule makes results repeat.
__device__ __forceinline__ int64_t sample_row(int64_t i, int64_t n, int64_t num_samples) {
  return num_samples >= n ? i : (int64_t)(((unsigned __int128)i * (unsigned __int128)n) / (unsigned __int128)num_samples);
}

__global__ void __launch_bounds__(256) sample_minmax_kernel(const DCol c, int64_t n, int64_t num_samples, int32_t kclass,
                                                            unsigned long long *__restrict__ out /* [min key, max key, count] */) {
  const int64_t m = num_samples >= n ? n : num_samples;
  unsigned long long lo = ~0ull, hi = 0ull, cnt = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (int64_t)gridDim.x * blockDim.x) {
    const uint64_t bits = load_bits(c, sample_row(i, n, num_samples));
    if (key_bits_is_nan(c.type, bits)) continue;  // as arrow's min_max, which the reference calls
    const unsigned long long k = order_key(bits, kclass, c.width);
    lo = k < lo ? k : lo;
    hi = k > hi ? k : hi;
    cnt++;
  }
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long l2 = __shfl_xor_sync(0xFFFFFFFFu, lo, o), h2 = __shfl_xor_sync(0xFFFFFFFFu, hi, o);
    lo = l2 < lo ? l2 : lo;
    hi = h2 > hi ? h2 : hi;
    cnt += __shfl_xor_sync(0xFFFFFFFFu, cnt, o);
  }
  if ((threadIdx.x & 31) == 0 && cnt) {
    atomicMin(&out[0], lo);
    atomicMax(&out[1], hi);
    atomicAdd(&out[2], cnt);
  }
}

constexpr int kSharedBins = 4096;

// histogram of the sample over the num_bins + 2 bins
template <typename T>
__global__ void __launch_bounds__(256) range_hist_kernel(const T *__restrict__ data, int64_t n, int64_t num_samples, const RangeOf<T> r,
                                                         unsigned long long *__restrict__ hist) {
  __shared__ uint32_t sh[kSharedBins];
  const uint32_t bins = r.nb + 2;
  const bool use_shared = bins <= (uint32_t)kSharedBins;
  if (use_shared) {
    for (uint32_t t = threadIdx.x; t < bins; t += blockDim.x) sh[t] = 0;
    __syncthreads();
  }
  const int64_t m = num_samples >= n ? n : num_samples;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (int64_t)gridDim.x * blockDim.x) {
    const uint32_t pos = bin_pos<T>(data[sample_row(i, n, num_samples)], r);
    if (use_shared) atomicAdd(&sh[pos], 1u);
    else atomicAdd(&hist[pos], 1ull);
  }
  if (use_shared) {
    __syncthreads();
    for (uint32_t t = threadIdx.x; t < bins; t += blockDim.x)
      if (sh[t]) atomicAdd(&hist[t], (unsigned long long)sh[t]);
  }
}

// target partition of every row (+ rows per partition)
template <typename T>
__global__ void __launch_bounds__(256) range_pid_kernel(const T *__restrict__ data, int64_t n, const RangeOf<T> r, const uint32_t *__restrict__ bin_to_partition,
                                                        uint32_t num_partitions, int32_t ascending, int32_t no_sample, uint32_t *__restrict__ pid,
                                                        unsigned long long *__restrict__ part_hist) {
  __shared__ uint32_t sh[kSharedBins];
  const bool use_shared = part_hist != nullptr && num_partitions <= (uint32_t)kSharedBins;
  if (use_shared) {
    for (uint32_t t = threadIdx.x; t < num_partitions; t += blockDim.x) sh[t] = 0;
    __syncthreads();
  }
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    // nothing was sampled anywhere: every value is "below min" and falls into bin 0 (what the reference's empty min/max leads to)
    const uint32_t pos = no_sample ? 0u : bin_pos<T>(data[i], r);
    uint32_t p = bin_to_partition[pos];
    p = p >= num_partitions ? num_partitions - 1 : p;
    if (!ascending) p = num_partitions - 1 - p;
    pid[i] = p;
    if (use_shared) atomicAdd(&sh[p], 1u);
    else if (part_hist) atomicAdd(&part_hist[p], 1ull);
  }
  if (use_shared) {
    __syncthreads();
    for (uint32_t t = threadIdx.x; t < num_partitions; t += blockDim.x)
      if (sh[t]) atomicAdd(&part_hist[t], (unsigned long long)sh[t]);
  }
}

// runs BODY with T = the C++ type the reference's RangePartitionKernel<ARROW_T> computes in
#define CYL_DISPATCH_VALUE_TYPE(type, ...)                                                                       \
  switch (type) {                                                                                                \
    case CYL_UINT8: { using T = uint8_t; __VA_ARGS__ } break;                                                           \
    case CYL_INT8: { using T = int8_t; __VA_ARGS__ } break;                                                             \
    case CYL_UINT16: { using T = uint16_t; __VA_ARGS__ } break;                                                         \
    case CYL_INT16: { using T = int16_t; __VA_ARGS__ } break;                                                           \
    case CYL_UINT32: { using T = uint32_t; __VA_ARGS__ } break;                                                         \
    case CYL_INT32: case CYL_DATE32: case CYL_TIME32: { using T = int32_t; __VA_ARGS__ } break;                         \
    case CYL_UINT64: { using T = uint64_t; __VA_ARGS__ } break;                                                         \
    case CYL_INT64: case CYL_DATE64: case CYL_TIMESTAMP: case CYL_TIME64: { using T = int64_t; __VA_ARGS__ } break;     \
    case CYL_FLOAT: { using T = float; __VA_ARGS__ } break;                                                             \
    case CYL_DOUBLE: { using T = double; __VA_ARGS__ } break;                                                           \
    default: return fail(CYL_NotImplemented, "range partitioning: column type %d is not numeric", (int)(type)); \
  }

int range_column(const cyl_column_t *col, int64_t nrows, DCol *out, const char *who) {
  if (!col) return fail(CYL_Invalid, "%s: null column", who);
  DCols d;
  CYL_TRY(make_dcols(col, 1, nrows, &d, who));
  if (!sortable(d.c[0])) return fail(CYL_NotImplemented, "%s: the sort column must be a numeric or temporal column", who);
  *out = d.c[0];
  return CYL_OK;
}

}  // namespace
}  // namespace cylb

extern "C" int cyl_range_sample_minmax(const cyl_column_t *col, int64_t nrows, int64_t num_samples, uint64_t *min_key, uint64_t *max_key,
                                       int64_t *count, void *stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (!min_key || !max_key || !count) return fail(CYL_Invalid, "cyl_range_sample_minmax: null output");
  if (nrows < 0 || num_samples < 0) return fail(CYL_Invalid, "cyl_range_sample_minmax: negative size");
  DCol c;
  CYL_TRY(range_column(col, nrows, &c, "cyl_range_sample_minmax"));
  *min_key = ~0ull;
  *max_key = 0;
  *count = 0;
  const int64_t m = num_samples >= nrows ? nrows : num_samples;
  if (m == 0) return CYL_OK;
  Scratch out;
  CYL_TRY(out.alloc(3 * sizeof(unsigned long long), s));
  const unsigned long long init[3] = {~0ull, 0ull, 0ull};
  CYL_CUDA(cudaMemcpyAsync(out.p, init, sizeof(init), cudaMemcpyHostToDevice, s));
  sample_minmax_kernel<<<sort_grid(m), 256, 0, s>>>(c, nrows, num_samples, key_class_of(c.type), out.as<unsigned long long>());
  CYL_LAUNCH_CHECK();
  unsigned long long host[3];
  CYL_CUDA(cudaMemcpyAsync(host, out.p, sizeof(host), cudaMemcpyDeviceToHost, s));
  CYL_CUDA(cudaStreamSynchronize(s));
  *min_key = host[0];
  *max_key = host[1];
  *count = (int64_t)host[2];
  return CYL_OK;
}

extern "C" int cyl_range_sample_histogram(const cyl_column_t *col, int64_t nrows, int64_t num_samples, uint64_t min_key, uint64_t max_key,
                                          uint32_t num_bins, uint64_t *hist, void *stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (!hist) return fail(CYL_Invalid, "cyl_range_sample_histogram: null output");
  if (nrows < 0 || num_samples < 0) return fail(CYL_Invalid, "cyl_range_sample_histogram: negative size");
  if (num_bins < 1 || num_bins > (1u << 24)) return fail(CYL_Invalid, "cyl_range_sample_histogram: num_bins %u out of range", num_bins);
  DCol c;
  CYL_TRY(range_column(col, nrows, &c, "cyl_range_sample_histogram"));
  const size_t bins = (size_t)num_bins + 2;
  for (size_t b = 0; b < bins; b++) hist[b] = 0;
  const int64_t m = num_samples >= nrows ? nrows : num_samples;
  if (m == 0) return CYL_OK;
  Scratch dev;
  CYL_TRY(dev.alloc(bins * sizeof(unsigned long long), s));
  CYL_CUDA(cudaMemsetAsync(dev.p, 0, bins * sizeof(unsigned long long), s));
  const int32_t kclass = key_class_of(c.type);
  CYL_DISPATCH_VALUE_TYPE(c.type, {
    const RangeOf<T> r = make_range<T>(min_key, max_key, kclass, num_bins);
    range_hist_kernel<T><<<sort_grid(m), 256, 0, s>>>(reinterpret_cast<const T *>(c.data), nrows, num_samples, r, dev.as<unsigned long long>());
  })
  CYL_LAUNCH_CHECK();
  CYL_CUDA(cudaMemcpyAsync(hist, dev.p, bins * sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
  CYL_CUDA(cudaStreamSynchronize(s));
  return CYL_OK;
}

extern "C" int cyl_range_partition(const cyl_column_t *col, int64_t nrows, uint64_t min_key, uint64_t max_key, int64_t total_samples, uint32_t num_bins,
                                   const uint32_t *bin_to_partition, uint32_t num_partitions, int32_t ascending, uint32_t *pid, uint64_t *partition_hist,
                                   void *stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (!bin_to_partition || (nrows > 0 && !pid)) return fail(CYL_Invalid, "cyl_range_partition: null argument");
  if (nrows < 0) return fail(CYL_Invalid, "cyl_range_partition: negative size");
  if (num_partitions < 1) return fail(CYL_Invalid, "cyl_range_partition: num_partitions must be positive");
  if (num_bins < 1 || num_bins > (1u << 24)) return fail(CYL_Invalid, "cyl_range_partition: num_bins %u out of range", num_bins);
  DCol c;
  CYL_TRY(range_column(col, nrows, &c, "cyl_range_partition"));
  if (partition_hist) CYL_CUDA(cudaMemsetAsync(partition_hist, 0, (size_t)num_partitions * sizeof(uint64_t), s));
  if (nrows == 0) return CYL_OK;
  const size_t bins = (size_t)num_bins + 2;
  Scratch map;
  CYL_TRY(map.alloc(bins * sizeof(uint32_t), s));
  CYL_CUDA(cudaMemcpyAsync(map.p, bin_to_partition, bins * sizeof(uint32_t), cudaMemcpyHostToDevice, s));
  const int32_t kclass = key_class_of(c.type);
  const int32_t no_sample = total_samples <= 0 ? 1 : 0;
  CYL_DISPATCH_VALUE_TYPE(c.type, {
    const RangeOf<T> r = no_sample ? RangeOf<T>{T(0), T(0), T(0), num_bins} : make_range<T>(min_key, max_key, kclass, num_bins);
    range_pid_kernel<T><<<sort_grid(nrows), 256, 0, s>>>(reinterpret_cast<const T *>(c.data), nrows, r, map.as<uint32_t>(), num_partitions, ascending,
                                                         no_sample, pid, reinterpret_cast<unsigned long long *>(partition_hist));
  })
  CYL_LAUNCH_CHECK();
  CYL_CUDA(cudaStreamSynchronize(s));  // bin_to_partition is the caller's host memory (pageable: the copy above may still be reading it)
  return CYL_OK;
}

// bin_to_partition of RangePartitionKernel::build_bin_to_partition (arrow_partition_kernels.cpp:409-427): the histogram's prefix
// frequency crossing k / num_partitions moves on to the next partition. float_t arithmetic, as there. Host only.
extern "C" int cyl_range_quantile_split(const uint64_t *global_counts, uint32_t num_bins, uint64_t total_samples, uint32_t num_partitions,
                                        uint32_t *bin_to_partition) {
  if (!global_counts || !bin_to_partition) return fail(CYL_Invalid, "cyl_range_quantile_split: null argument");
  if (num_partitions < 1) return fail(CYL_Invalid, "cyl_range_quantile_split: num_partitions must be positive");
  const float quantile = (float)(1.0 / num_partitions);
  float prefix_sum = 0, target_quantile = quantile;
  uint32_t curr_partition = 0;
  for (size_t b = 0; b < (size_t)num_bins + 2; b++) {
    bin_to_partition[b] = curr_partition;
    const float freq = total_samples ? (float)global_counts[b] / (float)total_samples : 0.0f;
    prefix_sum += freq;
    if (prefix_sum > target_quantile) {
      curr_partition += (curr_partition < num_partitions - 1);
      target_quantile += quantile;
    }
  }
  return CYL_OK;
}

