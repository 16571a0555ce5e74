// groupby.cu -- hash group-by: dense group ids in first-seen order, then SUM / COUNT / MIN / MAX by atomics.
//
// Replaces (reference, relative to cpp/src/cylon/): groupby/hash_groupby.cpp:92-126 (make_groups: bytell_hash_map
// row -> group id, first-occurrence filter), :142-201 (aggregate<op,T>) and the kernels of
// compute/aggregate_kernels.hpp:410-485. Semantics kept bit for bit: group ids are numbered in first-seen order,
// the key of a group is its first row, SUM accumulates in the column's own C type (integers wrap), COUNT counts
// every row into int64, MIN starts from numeric_limits<T>::max() and MAX from numeric_limits<T>::min() -- which
// for float/double is the smallest positive normal number (reference quirk).
// Floating-point SUM uses atomicAdd, so the order of additions differs from the reference's row order:
// results agree within the 1e-12 relative tolerance BASELINE.json states, not bit for bit.
#include <float.h>

#include <limits>
#include <type_traits>

#include "rowtable.cuh"
#include "scan.cuh"

namespace cylb {

static inline int grid1d(int64_t n) {
  int64_t b = div_up(n, 256);
  const int64_t cap = (int64_t)num_sms() * 16;
  return (int)(b < 1 ? 1 : (b < cap ? b : cap));
}

struct GroupArgs {
  DCols keys;
  RowTable t;
  int64_t n;
  int32_t exact;
  const uint32_t *pslot;
  uint8_t *first;
  const int64_t *pos;
  int64_t *gid;
};

// first[i] = row i is its group's first occurrence; hashed mode verifies the row against the representative
__global__ void __launch_bounds__(256) group_flag_kernel(const __grid_constant__ GroupArgs a) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < a.n; i += (int64_t)gridDim.x * blockDim.x) {
    const unsigned long long rep = a.t.slots[a.pslot[i]].rep;
    const bool is_first = rep == (unsigned long long)i;
    a.first[i] = is_first ? 1 : 0;
    if (!a.exact && !is_first && !rows_equal(a.keys, (int64_t)rep, a.keys, i)) a.t.flags[1] = 1;
  }
}

__global__ void __launch_bounds__(256) group_publish_kernel(const __grid_constant__ GroupArgs a) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < a.n; i += (int64_t)gridDim.x * blockDim.x)
    if (a.first[i]) a.t.slots[a.pslot[i]].rep = (unsigned long long)a.pos[i];
}

__global__ void __launch_bounds__(256) group_assign_kernel(const __grid_constant__ GroupArgs a) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < a.n; i += (int64_t)gridDim.x * blockDim.x)
    a.gid[i] = (int64_t)a.t.slots[a.pslot[i]].rep;
}

// ---- aggregation ----------------------------------------------------------------------------------------------
template <typename T>
struct AccWord {
  using type = typename std::conditional<sizeof(T) == 8, unsigned long long, unsigned int>::type;
};

// order-preserving map T -> unsigned word (for atomicMin / atomicMax), and back
template <typename T>
__host__ __device__ inline typename AccWord<T>::type enc_ordered(T v) {
  using W = typename AccWord<T>::type;
  if constexpr (std::is_floating_point<T>::value) {
    W b;
    memcpy(&b, &v, sizeof(W));
    const W sign = (W)1 << (8 * sizeof(W) - 1);
    return (b & sign) ? ~b : (b | sign);
  } else if constexpr (std::is_signed<T>::value) {
    using S = typename std::make_signed<W>::type;
    return (W)(S)v ^ ((W)1 << (8 * sizeof(W) - 1));
  } else {
    return (W)v;
  }
}
template <typename T>
__host__ __device__ inline T dec_ordered(typename AccWord<T>::type w) {
  using W = typename AccWord<T>::type;
  if constexpr (std::is_floating_point<T>::value) {
    const W sign = (W)1 << (8 * sizeof(W) - 1);
    const W b = (w & sign) ? (w & ~sign) : ~w;
    T v;
    memcpy(&v, &b, sizeof(W));
    return v;
  } else if constexpr (std::is_signed<T>::value) {
    using S = typename std::make_signed<W>::type;
    return (T)(S)(w ^ ((W)1 << (8 * sizeof(W) - 1)));
  } else {
    return (T)w;
  }
}

constexpr int kAggSumSq = 100;  // internal: sum of squares (VAR / STDDEV)

template <typename W>
__global__ void fill_words_kernel(W *__restrict__ p, int64_t n, W v) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = v;
}

template <typename T, int OP>
__global__ void __launch_bounds__(256) aggregate_kernel(const T *__restrict__ v, const int64_t *__restrict__ gid, int64_t n,
                                                        typename AccWord<T>::type *__restrict__ acc) {
  using W = typename AccWord<T>::type;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t g = gid[i];
    const T x = v[i];
    if constexpr (OP == kAggSumSq) {  // VarianceKernel's first state: sum of value * value in the column's own type
      if constexpr (std::is_floating_point<T>::value) {
        atomicAdd(reinterpret_cast<T *>(acc) + g, x * x);
      } else {
        const W w = (W)(typename std::make_unsigned<T>::type)x;  // the product wraps like T's after truncation whatever the sign
        atomicAdd(acc + g, w * w);
      }
    } else if constexpr (OP == CYL_AGG_SUM) {
      if constexpr (std::is_floating_point<T>::value) {
        atomicAdd(reinterpret_cast<T *>(acc) + g, x);
      } else if constexpr (std::is_signed<T>::value) {
        using S = typename std::make_signed<W>::type;
        atomicAdd(acc + g, (W)(S)x);  // two's complement: wraps exactly like the C type after truncation
      } else {
        atomicAdd(acc + g, (W)x);
      }
    } else if constexpr (OP == CYL_AGG_MIN) {
      atomicMin(acc + g, enc_ordered<T>(x));
    } else {
      atomicMax(acc + g, enc_ordered<T>(x));
    }
  }
}

template <typename T, int OP>
__global__ void finalize_kernel(const typename AccWord<T>::type *acc, int64_t g, T *out) {  // acc may alias out
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < g; i += (int64_t)gridDim.x * blockDim.x) {
    const typename AccWord<T>::type w = acc[i];
    if constexpr (OP == CYL_AGG_SUM) {
      out[i] = (T)w;  // only reached for integer types narrower than the accumulator word
    } else {
      out[i] = dec_ordered<T>(w);
    }
  }
}

__global__ void __launch_bounds__(256) count_kernel(const int64_t *__restrict__ gid, int64_t n, unsigned long long *__restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) atomicAdd(out + gid[i], 1ull);
}

template <typename T, int OP>
static int aggregate_typed(const void *data, const int64_t *gid, int64_t n, int64_t g, void *out, cudaStream_t s) {
  using W = typename AccWord<T>::type;
  // accumulate in the output buffer when the word sizes agree, else in scratch words and narrow at the end
  Scratch tmp;
  W *acc = reinterpret_cast<W *>(out);
  const bool in_place = sizeof(W) == sizeof(T);
  if (!in_place) {
    CYL_TRY(tmp.alloc((size_t)(g + 1) * sizeof(W), s));
    acc = tmp.as<W>();
  }
  W init = 0;
  if (OP == CYL_AGG_MIN) init = enc_ordered<T>(std::numeric_limits<T>::max());   // aggregate_kernels.hpp:459
  if (OP == CYL_AGG_MAX) init = enc_ordered<T>(std::numeric_limits<T>::min());   // aggregate_kernels.hpp:477 (DBL_MIN for double)
  if (g > 0) {
    fill_words_kernel<W><<<grid1d(g), 256, 0, s>>>(acc, g, init);
    CYL_LAUNCH_CHECK();
  }
  if (n > 0) {
    ProfScope ps("groupby_aggregate", s);
    aggregate_kernel<T, OP><<<grid1d(n), 256, 0, s>>>(static_cast<const T *>(data), gid, n, acc);
    CYL_LAUNCH_CHECK();
  }
  const bool sum_in_place = OP == CYL_AGG_SUM && in_place;
  if (g > 0 && !sum_in_place) {
    finalize_kernel<T, OP><<<grid1d(g), 256, 0, s>>>(acc, g, static_cast<T *>(out));  // in place: same index read then written
    CYL_LAUNCH_CHECK();
  }
  return CYL_OK;
}

// MeanKernel / VarianceKernel finalisers (aggregate_kernels.hpp:344-405): sum, sum of squares (both in T) and count -> result
template <typename T>
__global__ void finalize_mean_kernel(const typename AccWord<T>::type *sum, const unsigned long long *cnt, int64_t g, T *out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < g; i += (int64_t)gridDim.x * blockDim.x) {
    T s;
    if constexpr (std::is_floating_point<T>::value) s = reinterpret_cast<const T *>(sum)[i];
    else s = (T)sum[i];
    const long long c = (long long)cnt[i];
    out[i] = c ? (T)(s / c) : (T)0;  // T / int64 with the usual arithmetic conversions, stored as T
  }
}

template <typename T>
__global__ void finalize_var_kernel(const typename AccWord<T>::type *sumsq, const typename AccWord<T>::type *sum, const unsigned long long *cnt,
                                    int64_t g, int do_std, double *out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < g; i += (int64_t)gridDim.x * blockDim.x) {
    T s, q;
    if constexpr (std::is_floating_point<T>::value) {
      s = reinterpret_cast<const T *>(sum)[i];
      q = reinterpret_cast<const T *>(sumsq)[i];
    } else {
      s = (T)sum[i];
      q = (T)sumsq[i];
    }
    const long long c = (long long)cnt[i];
    double r = 0;
    if (c > 1) {
      const double mean = (double)s / (double)c, mean_sq = (double)q / (double)c;
      r = (double)c * (mean_sq - mean * mean) / (double)(c - 1);  // ddof = 1: VarKernelOptions' default (:72)
      if (do_std) r = sqrt(r);
    }
    out[i] = r;
  }
}

template <typename T>
static int aggregate_moments(int op, const void *data, const int64_t *gid, int64_t n, int64_t g, void *out, cudaStream_t s) {
  using W = typename AccWord<T>::type;
  Scratch sum, sq, cnt;
  CYL_TRY(sum.alloc((size_t)(g + 1) * sizeof(W), s));
  CYL_TRY(cnt.alloc((size_t)(g + 1) * sizeof(unsigned long long), s));
  CYL_CUDA(cudaMemsetAsync(sum.p, 0, (size_t)(g + 1) * sizeof(W), s));
  CYL_CUDA(cudaMemsetAsync(cnt.p, 0, (size_t)(g + 1) * sizeof(unsigned long long), s));
  const bool var = op != CYL_AGG_MEAN;
  if (var) {
    CYL_TRY(sq.alloc((size_t)(g + 1) * sizeof(W), s));
    CYL_CUDA(cudaMemsetAsync(sq.p, 0, (size_t)(g + 1) * sizeof(W), s));
  }
  if (n > 0) {
    ProfScope ps("groupby_aggregate", s);
    aggregate_kernel<T, CYL_AGG_SUM><<<grid1d(n), 256, 0, s>>>(static_cast<const T *>(data), gid, n, sum.as<W>());
    CYL_LAUNCH_CHECK();
    count_kernel<<<grid1d(n), 256, 0, s>>>(gid, n, cnt.as<unsigned long long>());
    CYL_LAUNCH_CHECK();
    if (var) {
      aggregate_kernel<T, kAggSumSq><<<grid1d(n), 256, 0, s>>>(static_cast<const T *>(data), gid, n, sq.as<W>());
      CYL_LAUNCH_CHECK();
    }
  }
  if (g > 0) {
    if (var) finalize_var_kernel<T><<<grid1d(g), 256, 0, s>>>(sq.as<W>(), sum.as<W>(), cnt.as<unsigned long long>(), g, op == CYL_AGG_STDDEV, static_cast<double *>(out));
    else finalize_mean_kernel<T><<<grid1d(g), 256, 0, s>>>(sum.as<W>(), cnt.as<unsigned long long>(), g, static_cast<T *>(out));
    CYL_LAUNCH_CHECK();
  }
  return CYL_OK;
}

template <typename T>
static int aggregate_op(int op, const void *data, const int64_t *gid, int64_t n, int64_t g, void *out, cudaStream_t s) {
  if (op == CYL_AGG_MEAN || op == CYL_AGG_VAR || op == CYL_AGG_STDDEV) return aggregate_moments<T>(op, data, gid, n, g, out, s);
  switch (op) {
    case CYL_AGG_SUM: return aggregate_typed<T, CYL_AGG_SUM>(data, gid, n, g, out, s);
    case CYL_AGG_MIN: return aggregate_typed<T, CYL_AGG_MIN>(data, gid, n, g, out, s);
    case CYL_AGG_MAX: return aggregate_typed<T, CYL_AGG_MAX>(data, gid, n, g, out, s);
    default: return fail(CYL_NotImplemented, "cyl_aggregate: aggregation op %d is not on the accelerated path (NUNIQUE / QUANTILE are not)", op);
  }
}

}  // namespace cylb

using namespace cylb;

extern "C" {

int cyl_groupby_hash(const cyl_column_t *keys, int32_t nkeys, int64_t nrows, int64_t *gid, uint8_t *first_mask, int64_t *num_groups,
                     void *stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (nkeys < 1) return fail(CYL_Invalid, "cyl_groupby_hash: needs at least one key column");
  if (!num_groups) return fail(CYL_Invalid, "cyl_groupby_hash: null output");
  GroupArgs a;
  memset(&a, 0, sizeof(a));
  CYL_TRY(make_dcols(keys, nkeys, nrows, &a.keys, "cyl_groupby_hash"));
  *num_groups = 0;
  if (nrows == 0) return CYL_OK;
  if (nrows > (1ll << 30)) return fail(CYL_CapacityError, "cyl_groupby_hash: %lld rows (max 2^30 per GPU: 32-bit slot indices)", (long long)nrows);
  a.n = nrows;
  a.exact = is_exact_key(a.keys) ? 1 : 0;
  a.first = first_mask;
  a.gid = gid;
  Scratch pslot, pos;
  CYL_TRY(pslot.alloc((size_t)nrows * sizeof(uint32_t), s));
  CYL_TRY(pos.alloc((size_t)(nrows + 1) * sizeof(int64_t), s));
  a.pslot = pslot.as<uint32_t>();
  RowTableHost rt;
  const int64_t max_cap = next_pow2(nrows * 2 < 1024 ? 1024 : nrows * 2);
  int64_t cap = rowtable_initial_capacity(nrows);
  unsigned long long seed = 0;
  for (int attempt = 0;; attempt++) {
    if (attempt > 64) return fail(CYL_ExecutionError, "cyl_groupby_hash: could not place all keys");
    CYL_TRY(rt.alloc(cap, s));
    {
      ProfScope ps("groupby_insert", s);
      CYL_TRY(rowtable_insert(rt, a.keys, nrows, a.exact != 0, seed, 0, pslot.as<uint32_t>()));
    }
    int overflow = 0, collision = 0;
    CYL_TRY(rt.read_flags(&overflow, &collision));
    if (overflow) {
      if (cap >= max_cap) return fail(CYL_ExecutionError, "cyl_groupby_hash: table overflow at full capacity");
      cap = cap * 4 > max_cap ? max_cap : cap * 4;
      continue;
    }
    a.t = rt.t;
    {
      ProfScope ps("groupby_flag", s);
      group_flag_kernel<<<grid1d(nrows), 256, 0, s>>>(a);
      CYL_LAUNCH_CHECK();
    }
    if (!a.exact) {
      CYL_TRY(rt.read_flags(&overflow, &collision));
      if (collision) {  // two unequal key tuples share a 64-bit tag: redo with another seed
        seed = seed * 6364136223846793005ULL + 1442695040888963407ULL;
        continue;
      }
    }
    break;
  }
  {
    ProfScope ps("groupby_scan", s);
    CYL_TRY(exclusive_scan<uint8_t>(first_mask, nrows, pos.as<int64_t>(), s));
  }
  a.pos = pos.as<int64_t>();
  {
    ProfScope ps("groupby_assign_ids", s);
    group_publish_kernel<<<grid1d(nrows), 256, 0, s>>>(a);
    CYL_LAUNCH_CHECK();
    group_assign_kernel<<<grid1d(nrows), 256, 0, s>>>(a);
    CYL_LAUNCH_CHECK();
  }
  int64_t g = 0;
  CYL_CUDA(cudaMemcpyAsync(&g, pos.as<int64_t>() + nrows, sizeof(int64_t), cudaMemcpyDeviceToHost, s));
  CYL_CUDA(cudaStreamSynchronize(s));
  *num_groups = g;
  return CYL_OK;
}

int cyl_aggregate(const cyl_column_t *col, const int64_t *gid, int64_t num_groups, int32_t op, void *out, void *stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (!col) return fail(CYL_Invalid, "cyl_aggregate: null column");
  const int64_t n = col->length;
  if (num_groups < 0) return fail(CYL_Invalid, "cyl_aggregate: negative group count");
  if (op == CYL_AGG_COUNT) {  // CountKernel: counts every row, nulls included (aggregate_kernels.hpp:432-447)
    CYL_CUDA(cudaMemsetAsync(out, 0, (size_t)(num_groups ? num_groups : 1) * sizeof(int64_t), s));
    if (n > 0) {
      ProfScope ps("groupby_aggregate", s);
      count_kernel<<<grid1d(n), 256, 0, s>>>(gid, n, static_cast<unsigned long long *>(out));
      CYL_LAUNCH_CHECK();
    }
    return CYL_OK;
  }
  switch (col->type) {  // value types of groupby/hash_groupby.cpp:234-247 (BOOL is bit-packed: not supported)
    case CYL_UINT8: return aggregate_op<uint8_t>(op, col->data, gid, n, num_groups, out, s);
    case CYL_INT8: return aggregate_op<int8_t>(op, col->data, gid, n, num_groups, out, s);
    case CYL_UINT16: return aggregate_op<uint16_t>(op, col->data, gid, n, num_groups, out, s);
    case CYL_INT16: return aggregate_op<int16_t>(op, col->data, gid, n, num_groups, out, s);
    case CYL_UINT32: return aggregate_op<uint32_t>(op, col->data, gid, n, num_groups, out, s);
    case CYL_INT32: return aggregate_op<int32_t>(op, col->data, gid, n, num_groups, out, s);
    case CYL_UINT64: return aggregate_op<unsigned long long>(op, col->data, gid, n, num_groups, out, s);
    case CYL_INT64: return aggregate_op<long long>(op, col->data, gid, n, num_groups, out, s);
    case CYL_FLOAT: return aggregate_op<float>(op, col->data, gid, n, num_groups, out, s);
    case CYL_DOUBLE: return aggregate_op<double>(op, col->data, gid, n, num_groups, out, s);
    default: return fail(CYL_ExecutionError, "unable to find aggregation fn (type id %d)", col->type);  // hash_groupby.cpp:295
  }
}

}  // extern "C"
