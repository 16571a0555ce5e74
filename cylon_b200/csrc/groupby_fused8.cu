// groupby_fused8.cu -- one-pass hash group-by for a single 8-byte key and 8-byte value columns (BASELINE config 4:
// int64 key, double value, SUM / COUNT / MIN / MAX).
//
// Replaces (reference, relative to cpp/src/cylon/): groupby/hash_groupby.cpp:92-126 (make_groups), :142-201
// (aggregate<op,T>) and the kernels of compute/aggregate_kernels.hpp:410-485 for that shape, in ONE sweep over the rows
// instead of the general path's six (row table insert, first-occurrence flags, scan, id publish, id assign, one atomic
// pass per aggregate -- all against a table in HBM/L2).
//
// Rows are packed as records [key, value columns...] and radix-partitioned on mix64(key) with the staged passes of the
// fused join (join_fused8.cu) into bins of ~2000 rows. One CTA then owns a bin: its group table -- key + one 8-byte
// state per aggregate, 4096 slots -- lives in SHARED memory, every row of the bin finds or claims its slot there and
// folds all requested aggregates into it, and the finished groups leave as coalesced column writes. No group id per
// row, no per-aggregate relaunch, no global atomic per row; a bin cannot hold more groups than rows, so the table always
// fits (a bin made long by one hot key has few groups). States use order-preserving 64-bit encodings so that MIN / MAX
// are unsigned compare-and-swaps whatever the type; SUM(double) is a compare-and-swap loop on the bits, SUM(int) wraps
// like the C type. MAX starts from numeric_limits<T>::min() -- DBL_MIN for doubles -- as the reference's does
// (aggregate_kernels.hpp:477).
//
// Group order: bins in hash order, arbitrary inside a bin (the reference numbers groups in first-seen order; results
// are compared as row multisets). The number of groups is only known afterwards: optimistic capacity, exact count
// always returned, the caller repeats the aggregation with the exact size if it did not fit.
#include <float.h>

#include "common.cuh"
#include "join_fused.cuh"

namespace cylb {

constexpr int kGbThreads = 1024;
constexpr int kGbSlots = 4096;      // power of two; a bin is sized to <= ~2000 rows on average, so load stays below 0.6
constexpr int kGbMaxAggs = 4;
constexpr unsigned long long kGbEmpty = 0xFFFFFFFFFFFFFFFFull;

enum GbOp : int32_t { GB_SUM_F64 = 0, GB_SUM_INT = 1, GB_COUNT = 2, GB_MIN = 3, GB_MAX = 4 };
enum GbEnc : int32_t { GB_ENC_U64 = 0, GB_ENC_I64 = 1, GB_ENC_F64 = 2 };

struct GbArgs {
  const uint64_t *rec;      // partition-ordered records [key, v0, v1, ...], W words each
  const int64_t *bin_off;   // [nbins + 1]
  int64_t nbins;
  int32_t naggs;
  int32_t op[kGbMaxAggs];   // GbOp
  int32_t src[kGbMaxAggs];  // word of the record the aggregate reads (unused for COUNT)
  int32_t enc[kGbMaxAggs];  // GbEnc of that word (MIN / MAX)
  unsigned long long init[kGbMaxAggs];
  uint64_t *out_key;
  uint64_t *out_agg[kGbMaxAggs];
  int64_t capacity;
  unsigned long long *out_cursor;  // groups written so far (exact count even when capacity is exceeded)
  unsigned int *work_counter;
  int *overflow;                   // set when a bin held more groups than the table has slots
};

// order-preserving map of an 8-byte value onto unsigned 64-bit
__device__ __forceinline__ unsigned long long gb_encode(uint64_t bits, int enc) {
  if (enc == GB_ENC_I64) return bits ^ 0x8000000000000000ull;
  if (enc == GB_ENC_F64) return (bits & 0x8000000000000000ull) ? ~bits : (bits | 0x8000000000000000ull);
  return bits;
}
__device__ __forceinline__ uint64_t gb_decode(unsigned long long w, int enc) {
  if (enc == GB_ENC_I64) return w ^ 0x8000000000000000ull;
  if (enc == GB_ENC_F64) return (w & 0x8000000000000000ull) ? (w & 0x7FFFFFFFFFFFFFFFull) : ~w;
  return w;
}
static unsigned long long gb_encode_host(uint64_t bits, int enc) {
  if (enc == GB_ENC_I64) return bits ^ 0x8000000000000000ull;
  if (enc == GB_ENC_F64) return (bits & 0x8000000000000000ull) ? ~bits : (bits | 0x8000000000000000ull);
  return bits;
}

// v[idx] without a dynamic register-array index (which would send the record to local memory)
template <int W>
__device__ __forceinline__ uint64_t gb_pick(const uint64_t (&v)[W], int idx) {
  uint64_t r = v[0];
#pragma unroll
  for (int w = 1; w < W; w++)
    if (w == idx) r = v[w];
  return r;
}

template <int W>
__global__ void __launch_bounds__(kGbThreads, 1) groupby_fused8_kernel(const __grid_constant__ GbArgs a) {
  extern __shared__ __align__(16) unsigned char gb_smem[];
  unsigned long long *keys = reinterpret_cast<unsigned long long *>(gb_smem);   // [kGbSlots + 1] (+1: the key equal to kGbEmpty)
  unsigned long long *state = keys + kGbSlots + 2;                              // [naggs][kGbSlots + 2]
  unsigned short *list = reinterpret_cast<unsigned short *>(state + (size_t)a.naggs * (kGbSlots + 2));  // [kGbSlots + 1] occupied slots
  __shared__ unsigned int s_bin, s_ngroups, s_special;
  __shared__ unsigned long long s_base;
  const int tid = threadIdx.x;
  constexpr int STRIDE = kGbSlots + 2;

  // the table starts clean and every bin leaves it clean (its emit phase resets exactly the slots it used)
  for (int i = tid; i < kGbSlots + 1; i += kGbThreads) keys[i] = kGbEmpty;
  for (int ag = 0; ag < a.naggs; ag++)
    for (int i = tid; i < kGbSlots + 1; i += kGbThreads) state[ag * STRIDE + i] = a.init[ag];
  if (tid == 0) {
    s_ngroups = 0;
    s_special = 0;
  }

  while (true) {
    __syncthreads();
    if (tid == 0) s_bin = atomicAdd(a.work_counter, 1u);
    __syncthreads();
    const int64_t bin = s_bin;
    if (bin >= a.nbins) break;
    const int64_t r0 = a.bin_off[bin], r1 = a.bin_off[bin + 1];
    if (r1 == r0) continue;

    for (int64_t r = r0 + tid; r < r1; r += kGbThreads) {
      uint64_t v[W];
      if constexpr (W == 2) {
        const ulonglong2 t = __ldcs(reinterpret_cast<const ulonglong2 *>(a.rec) + r);
        v[0] = t.x;
        v[1] = t.y;
      } else if constexpr (W == 4) {
        const ulonglong2 t0 = __ldcs(reinterpret_cast<const ulonglong2 *>(a.rec + r * 4));
        const ulonglong2 t1 = __ldcs(reinterpret_cast<const ulonglong2 *>(a.rec + r * 4 + 2));
        v[0] = t0.x; v[1] = t0.y; v[2] = t1.x; v[3] = t1.y;
      } else {
#pragma unroll
        for (int w = 0; w < W; w++) v[w] = ld_stream_u64(a.rec + r * W + w);
      }
      const unsigned long long key = v[0];
      // find or claim the key's slot
      int slot = -1;
      if (key == kGbEmpty) {
        slot = kGbSlots;
        s_special = 1;  // benign race: every writer stores 1
      } else {
        unsigned int h = (unsigned int)((key * 0x9E3779B97F4A7C15ULL) >> 40) & (kGbSlots - 1);
        for (int probes = 0; probes < kGbSlots; probes++) {
          unsigned long long cur = keys[h];
          if (cur == kGbEmpty) cur = atomicCAS(&keys[h], kGbEmpty, key);
          if (cur == kGbEmpty || cur == key) {
            slot = (int)h;
            break;
          }
          h = (h + 1) & (kGbSlots - 1);
        }
      }
      if (slot < 0) {
        *a.overflow = 1;
        continue;
      }
#pragma unroll
      for (int ag = 0; ag < kGbMaxAggs; ag++) {
        if (ag >= a.naggs) break;
        unsigned long long *st = &state[ag * STRIDE + slot];
        const int op = a.op[ag];
        if (op == GB_COUNT) {
          atomicAdd(st, 1ull);
        } else if (op == GB_SUM_INT) {
          atomicAdd(st, (unsigned long long)gb_pick<W>(v, a.src[ag]));
        } else if (op == GB_SUM_F64) {
          const double x = __longlong_as_double((long long)gb_pick<W>(v, a.src[ag]));
          unsigned long long seen = *st, prev;
          do {
            prev = seen;
            seen = atomicCAS(st, prev, (unsigned long long)__double_as_longlong(__longlong_as_double((long long)prev) + x));
          } while (seen != prev);
        } else {
          const unsigned long long e = gb_encode(gb_pick<W>(v, a.src[ag]), a.enc[ag]);
          unsigned long long seen = *st;
          while (op == GB_MIN ? e < seen : e > seen) {  // only a better value ever reaches the atomic
            const unsigned long long prev = atomicCAS(st, seen, e);
            if (prev == seen) break;
            seen = prev;
          }
        }
      }
    }
    __syncthreads();
    // emit: compact the occupied slots, reserve the bin's output range once, write the columns, reset the slots
    for (int i = tid; i < kGbSlots; i += kGbThreads)
      if (keys[i] != kGbEmpty) list[atomicAdd(&s_ngroups, 1u)] = (unsigned short)i;
    __syncthreads();
    if (tid == 0) {
      if (s_special) list[s_ngroups++] = (unsigned short)kGbSlots;
      s_base = atomicAdd(a.out_cursor, (unsigned long long)s_ngroups);
    }
    __syncthreads();
    const unsigned int ng = s_ngroups;
    const bool fits = (int64_t)(s_base + ng) <= a.capacity;
    for (unsigned int j = tid; j < ng; j += kGbThreads) {
      const int slot = list[j];
      const int64_t o = (int64_t)s_base + j;
      if (fits) a.out_key[o] = slot == kGbSlots ? kGbEmpty : keys[slot];
      keys[slot] = kGbEmpty;
#pragma unroll
      for (int ag = 0; ag < kGbMaxAggs; ag++) {
        if (ag >= a.naggs) break;
        const unsigned long long w = state[ag * STRIDE + slot];
        if (fits) a.out_agg[ag][o] = (a.op[ag] == GB_MIN || a.op[ag] == GB_MAX) ? gb_decode(w, a.enc[ag]) : w;
        state[ag * STRIDE + slot] = a.init[ag];
      }
    }
    __syncthreads();
    if (tid == 0) {
      s_ngroups = 0;
      s_special = 0;
    }
  }
}

static size_t gb_smem_bytes(int naggs) {
  return (size_t)(kGbSlots + 2) * 8 * (1 + naggs) + (size_t)(kGbSlots + 2) * 2 + 16;
}

template <int W>
static int launch_groupby(const GbArgs &a, int grid, cudaStream_t s) {
  static PerDeviceOnce attr_once;
  if (attr_once.first()) CYL_CUDA(cudaFuncSetAttribute(groupby_fused8_kernel<W>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gb_smem_bytes(kGbMaxAggs)));
  groupby_fused8_kernel<W><<<grid, kGbThreads, gb_smem_bytes(a.naggs), s>>>(a);
  CYL_LAUNCH_CHECK();
  return CYL_OK;
}

static bool gb_key_type_ok(int32_t t) {
  return t == CYL_INT64 || t == CYL_UINT64 || t == CYL_DOUBLE || t == CYL_DATE64 || t == CYL_TIMESTAMP || t == CYL_TIME64;
}
static bool gb_value_type_ok(int32_t t) { return t == CYL_INT64 || t == CYL_UINT64 || t == CYL_DOUBLE; }

}  // namespace cylb

using namespace cylb;

extern "C" {

int cyl_groupby_fused_supported(const cyl_column_t *key, const cyl_column_t *values, int32_t nvalues, const int32_t *agg_value, const int32_t *agg_op,
                                int32_t naggs) {
  if (!key || naggs < 1 || naggs > kGbMaxAggs || nvalues < 0 || nvalues > kFused8MaxCols - 1) return 0;
  if (!gb_key_type_ok(key->type) || key->byte_width != 8 || key->validity) return 0;
  for (int v = 0; v < nvalues; v++)
    if (!gb_value_type_ok(values[v].type) || values[v].byte_width != 8 || values[v].validity || values[v].length != key->length) return 0;
  for (int a = 0; a < naggs; a++) {
    if (agg_op[a] < CYL_AGG_SUM || agg_op[a] > CYL_AGG_COUNT) return 0;
    if (agg_value[a] < 0 || agg_value[a] >= nvalues) return 0;
  }
  return 1;
}

int cyl_groupby_fused(const cyl_column_t *key, const cyl_column_t *values, int32_t nvalues, const int32_t *agg_value, const int32_t *agg_op,
                      int32_t naggs, int64_t capacity, void *out_key, void *const *out_aggs, int64_t *num_groups, void *stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (!num_groups) return fail(CYL_Invalid, "cyl_groupby_fused: null output");
  *num_groups = 0;
  if (!cyl_groupby_fused_supported(key, values, nvalues, agg_value, agg_op, naggs))
    return fail(CYL_Invalid, "cyl_groupby_fused: unsupported shape (one 8-byte key, 8-byte int64/uint64/double values without nulls, <= %d aggregates)", kGbMaxAggs);
  const int64_t n = key->length;
  if (n > 0xFFFFFFFEll) return fail(CYL_CapacityError, "cyl_groupby_fused: %lld rows (max 2^32-2 per GPU)", (long long)n);
  if (n == 0) return CYL_OK;
  const int W = 1 + nvalues;
  cyl_column_t cols[kFused8MaxCols];
  cols[0] = *key;
  for (int v = 0; v < nvalues; v++) cols[1 + v] = values[v];

  // bins of ~2000 rows on average: at most that many groups per bin, half the table's slots
  int64_t nbins = 1;
  while (nbins * 2000 < n && nbins < (1ll << 22)) nbins *= 2;
  int bits = 0;
  while ((1ll << bits) < nbins) bits++;
  const int shift = 64 - bits;

  Scratch rec, off, counters;
  CYL_TRY(rec.alloc((size_t)n * W * 8 + 16, s));
  CYL_TRY(off.alloc((size_t)(nbins + 1) * sizeof(int64_t), s));
  CYL_TRY(f8_partition(cols, W, 0, n, nbins, shift, rec.as<uint64_t>(), off.as<int64_t>(), s));

  GbArgs a;
  memset(&a, 0, sizeof(a));
  a.rec = rec.as<uint64_t>();
  a.bin_off = off.as<int64_t>();
  a.nbins = nbins;
  a.naggs = naggs;
  for (int g = 0; g < naggs; g++) {
    const cyl_column_t &vc = values[agg_value[g]];
    a.src[g] = 1 + agg_value[g];
    a.enc[g] = vc.type == CYL_DOUBLE ? GB_ENC_F64 : (vc.type == CYL_INT64 ? GB_ENC_I64 : GB_ENC_U64);
    a.out_agg[g] = static_cast<uint64_t *>(out_aggs[g]);
    uint64_t lo = 0, hi = 0;  // numeric_limits<T>::min() / max() as bits
    if (vc.type == CYL_DOUBLE) {
      const double dmin = DBL_MIN, dmax = DBL_MAX;  // aggregate_kernels.hpp:459,477: min() is the smallest positive normal
      memcpy(&lo, &dmin, 8);
      memcpy(&hi, &dmax, 8);
    } else if (vc.type == CYL_INT64) {
      lo = 0x8000000000000000ull;
      hi = 0x7FFFFFFFFFFFFFFFull;
    } else {
      lo = 0;
      hi = 0xFFFFFFFFFFFFFFFFull;
    }
    switch (agg_op[g]) {
      case CYL_AGG_SUM:
        a.op[g] = vc.type == CYL_DOUBLE ? GB_SUM_F64 : GB_SUM_INT;
        a.init[g] = 0;
        break;
      case CYL_AGG_COUNT:
        a.op[g] = GB_COUNT;
        a.init[g] = 0;
        break;
      case CYL_AGG_MIN:
        a.op[g] = GB_MIN;
        a.init[g] = gb_encode_host(hi, a.enc[g]);
        break;
      default:
        a.op[g] = GB_MAX;
        a.init[g] = gb_encode_host(lo, a.enc[g]);
        break;
    }
  }
  a.out_key = static_cast<uint64_t *>(out_key);
  a.capacity = capacity;
  CYL_TRY(counters.alloc(64, s));
  CYL_CUDA(cudaMemsetAsync(counters.p, 0, 64, s));
  a.out_cursor = counters.as<unsigned long long>();
  a.work_counter = reinterpret_cast<unsigned int *>(counters.as<char>() + 16);
  a.overflow = reinterpret_cast<int *>(counters.as<char>() + 32);
  int64_t grid = num_sms();
  if (grid > nbins) grid = nbins;
  {
    ProfScope ps("groupby_fused_aggregate", s);
    int rc;
    switch (W) {
      case 1: rc = launch_groupby<1>(a, (int)grid, s); break;
      case 2: rc = launch_groupby<2>(a, (int)grid, s); break;
      case 3: rc = launch_groupby<3>(a, (int)grid, s); break;
      default: rc = launch_groupby<4>(a, (int)grid, s); break;
    }
    CYL_TRY(rc);
  }
  unsigned long long h[8] = {0};
  CYL_CUDA(cudaMemcpyAsync(h, counters.p, 64, cudaMemcpyDeviceToHost, s));
  CYL_CUDA(cudaStreamSynchronize(s));
  if (reinterpret_cast<int *>(h)[8]) return fail(CYL_CapacityError, "cyl_groupby_fused: a bin held more groups than the shared-memory table has slots");
  *num_groups = (int64_t)h[0];
  return CYL_OK;
}

}  // extern "C"
