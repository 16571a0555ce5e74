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

// Several small CTAs per SM rather than one large one: a bin is a short job (load its rows, two passes, write its groups),
// each step waiting on the one before, so a single 1024-thread CTA per SM spent most of its time in exposed latencies
// (6 ms for 125M rows); 256 threads and a 1024-slot table leave room for 4-5 CTAs per SM whose phases overlap.
static int g_gb_threads = 1024;  // CTA size (cyl_tune_set key 6): 256 / 512 / 1024
static int g_gb_bins_log2 = 0;  // > 0: force 2^bits bins (key 8)
static int g_gb_slots = 4096;   // table slots, a power of two (key 7); bins average 0.47 * slots rows, so load stays below 0.6
void groupby_fused8_tune(int key, int value) {
  if (key == 6) g_gb_threads = value;
  if (key == 7) g_gb_slots = value;
  if (key == 8) g_gb_bins_log2 = value;
}
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
  int32_t slots;                   // table slots (power of two)
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

// Folds a combined contribution `c` (already encoded for MIN / MAX) into a shared-memory state word that other warps may
// be updating too.
__device__ __forceinline__ void gb_update_enc(unsigned long long *st, int op, unsigned long long c) {
  if (op == GB_COUNT || op == GB_SUM_INT) {
    atomicAdd(st, c);
  } else if (op == GB_SUM_F64) {
    const double x = __longlong_as_double((long long)c);
    unsigned long long seen = *st, prev;
    do {
      prev = seen;
      seen = atomicCAS(st, prev, (unsigned long long)__double_as_longlong(__longlong_as_double((long long)prev) + x));
    } while (seen != prev);
  } else {
    unsigned long long seen = *st;
    while (op == GB_MIN ? c < seen : c > seen) {  // only a better value ever reaches the atomic
      const unsigned long long prev = atomicCAS(st, seen, c);
      if (prev == seen) break;
      seen = prev;
    }
  }
}

// folds one replica's state word into an accumulator of the same aggregate
__device__ __forceinline__ unsigned long long gb_merge(unsigned long long acc, unsigned long long w, int op) {
  if (op == GB_SUM_F64) return (unsigned long long)__double_as_longlong(__longlong_as_double((long long)acc) + __longlong_as_double((long long)w));
  if (op == GB_MIN) return w < acc ? w : acc;
  if (op == GB_MAX) return w > acc ? w : acc;
  return acc + w;  // COUNT, integer SUM (wraps)
}

constexpr int kGbKeep = 4;  // rows per thread whose record and slot stay in registers between the two passes over a bin

// Two passes over a bin. Pass 1 claims a slot and a dense group id for every distinct key. Pass 2 folds the rows into
// state words indexed by (replica, group id): the bin's ng groups are replicated R = min(32, slots / ng) times and warp w
// updates replica w mod R, so a bin with few groups (config 4: ~15 per bin, ~125 rows each) does not serialise 1024
// threads on 15 addresses -- measured 15.8 ms for 125M rows with one copy per group; the replicas are merged when the
// groups are written out.
template <int W, int kGbThreads>
__global__ void __launch_bounds__(kGbThreads, 1024 / kGbThreads) groupby_fused8_kernel(const __grid_constant__ GbArgs a) {
  const int kGbSlots = a.slots;
  extern __shared__ __align__(16) unsigned char gb_smem[];
  unsigned long long *keys = reinterpret_cast<unsigned long long *>(gb_smem);   // [kGbSlots + 2]; slot kGbSlots serves the key equal to kGbEmpty
  unsigned long long *state = keys + kGbSlots + 2;                              // [naggs][kGbSlots]: entry replica * ng + gid
  unsigned short *gid_of = reinterpret_cast<unsigned short *>(state + (size_t)a.naggs * kGbSlots);  // [kGbSlots + 2] slot -> group id
  unsigned short *slot_of = gid_of + kGbSlots + 2;                                                  // [kGbSlots + 2] group id -> slot
  unsigned int *locks = reinterpret_cast<unsigned int *>(slot_of + kGbSlots + 2);                   // [kGbSlots] one per state entry
  __shared__ unsigned int s_bin, s_ng, s_special;
  __shared__ unsigned long long s_base;
  const int tid = threadIdx.x;
  const unsigned int warp = tid >> 5;

  // the table starts clean and every bin leaves it clean (its emit phase resets exactly what it used)
  for (int i = tid; i < kGbSlots + 1; i += kGbThreads) keys[i] = kGbEmpty;
  for (int ag = 0; ag < a.naggs; ag++)
    for (int i = tid; i < kGbSlots; i += kGbThreads) state[ag * kGbSlots + i] = a.init[ag];
  for (int i = tid; i < kGbSlots; i += kGbThreads) locks[i] = 0u;
  if (tid == 0) {
    s_ng = 0;
    s_special = 0;
  }

  auto load_row = [&](int64_t r, uint64_t (&v)[W]) {
    if constexpr (W == 2) {
      const ulonglong2 t = *(reinterpret_cast<const ulonglong2 *>(a.rec) + r);
      v[0] = t.x;
      v[1] = t.y;
    } else if constexpr (W == 4) {
      const ulonglong2 t0 = *reinterpret_cast<const ulonglong2 *>(a.rec + r * 4);
      const ulonglong2 t1 = *reinterpret_cast<const ulonglong2 *>(a.rec + r * 4 + 2);
      v[0] = t0.x; v[1] = t0.y; v[2] = t1.x; v[3] = t1.y;
    } else {
#pragma unroll
      for (int w = 0; w < W; w++) v[w] = a.rec[r * W + w];
    }
  };
  // slot of `key`; claim == true may insert it (and then assigns the next group id)
  auto find_slot = [&](unsigned long long key, bool claim) -> int {
    if (key == kGbEmpty) {
      if (claim && atomicCAS(&s_special, 0u, 1u) == 0u) {
        const unsigned int g = atomicAdd(&s_ng, 1u);
        gid_of[kGbSlots] = (unsigned short)g;
        slot_of[g] = (unsigned short)kGbSlots;
      }
      return kGbSlots;
    }
    unsigned int h = (unsigned int)((key * 0x9E3779B97F4A7C15ULL) >> 40) & (kGbSlots - 1);
    for (int probes = 0; probes < kGbSlots; probes++) {
      unsigned long long cur = keys[h];
      if (cur == kGbEmpty && claim) {
        cur = atomicCAS(&keys[h], kGbEmpty, key);
        if (cur == kGbEmpty) {  // this thread inserted the key: it names the group
          const unsigned int g = atomicAdd(&s_ng, 1u);
          gid_of[h] = (unsigned short)g;
          slot_of[g] = (unsigned short)h;
          return (int)h;
        }
      }
      if (cur == key) return (int)h;
      if (cur == kGbEmpty) return -1;
      h = (h + 1) & (kGbSlots - 1);
    }
    return -1;
  };

  if (tid == 0) s_bin = atomicAdd(a.work_counter, 1u);
  while (true) {
    __syncthreads();
    const int64_t bin = s_bin;
    if (bin >= a.nbins || *reinterpret_cast<volatile int *>(a.overflow)) break;  // a full table anywhere voids the attempt
    // the next bin is requested now and the answer stored at the end of this one: its latency hides behind the work
    unsigned int next_bin = 0;
    if (tid == 0) next_bin = atomicAdd(a.work_counter, 1u);
    const int64_t r0 = a.bin_off[bin], r1 = a.bin_off[bin + 1];
    unsigned long long base_reg = 0;

    // ---- pass 1: every distinct key gets a slot and a group id
    uint64_t kept[kGbKeep][W];
    int kept_slot[kGbKeep];
#pragma unroll
    for (int k = 0; k < kGbKeep; k++) {
      const int64_t r = r0 + tid + (int64_t)k * kGbThreads;
      kept_slot[k] = -2;
      if (r < r1) {
        load_row(r, kept[k]);
        kept_slot[k] = find_slot(kept[k][0], true);
      }
    }
    for (int64_t r = r0 + tid + (int64_t)kGbKeep * kGbThreads; r < r1; r += kGbThreads) {  // long bins only (one hot key)
      uint64_t v[W];
      load_row(r, v);
      if (find_slot(v[0], true) < 0) *a.overflow = 1;
    }
    __syncthreads();
    unsigned int ng = s_ng;
    if (ng > (unsigned int)kGbSlots) {  // cannot happen while a bin has fewer rows than the table has slots; never index past the states
      if (tid == 0) *a.overflow = 1;
      ng = kGbSlots;
    }
    unsigned int R = ng ? kGbSlots / ng : 1;
    if (R > kGbThreads / 32) R = kGbThreads / 32;
    if (R < 1) R = 1;
    const unsigned int rep_base = (warp % R) * ng;
    if (tid == 0 && ng) base_reg = atomicAdd(a.out_cursor, (unsigned long long)ng);  // consumed after pass 2

    // ---- pass 2: fold every row into its group's state (replica of this warp)
    // The lanes of a warp that hold the same group (match.any) first combine their contributions in a tree over their
    // ranks -- log2(multiplicity) rounds of one shuffle per aggregate -- and only the lowest lane of each group touches
    // the state word: with PLAIN loads and stores when the warp owns its replica (R == warps per CTA), else with one
    // atomic per group instead of one per row. Letting every row update the state cost 12 cycles per row in
    // compare-and-swap traffic (5.3 ms for 125M rows with ~15 groups per bin), taking turns lane by lane 45 warp
    // instructions per row (ncu, profiles/r02_groupby_notes.md).
    const bool private_replica = R == kGbThreads / 32;
    const unsigned int lane = tid & 31, lt = (1u << lane) - 1;
    auto fold = [&](const uint64_t (&v)[W], int slot) {  // warp-uniform call; slot < 0: this lane has no row
      const bool valid = slot >= 0;
      const unsigned int e = rep_base + (valid ? gid_of[slot] : 0u);
      const unsigned int peers = __match_any_sync(0xffffffffu, valid ? e : (0xFFFF0000u | lane));
      const unsigned int rank = __popc(peers & lt), m = __popc(peers);
      const unsigned int maxm = __reduce_max_sync(0xffffffffu, valid ? m : 1u);
      unsigned long long c[kGbMaxAggs];
#pragma unroll
      for (int ag = 0; ag < kGbMaxAggs; ag++) {
        c[ag] = 0;
        if (ag < a.naggs) {
          const uint64_t bits = gb_pick<W>(v, a.src[ag]);
          c[ag] = a.op[ag] == GB_COUNT ? 1ull : ((a.op[ag] == GB_MIN || a.op[ag] == GB_MAX) ? gb_encode(bits, a.enc[ag]) : bits);
        }
      }
      for (unsigned int step = 1; step < maxm; step <<= 1) {
        const bool taker = valid && (rank & (2 * step - 1)) == 0 && rank + step < m;
        const int from = taker ? (int)__fns(peers, 0, (int)(rank + step + 1)) : (int)lane;
#pragma unroll
        for (int ag = 0; ag < kGbMaxAggs; ag++) {
          if (ag >= a.naggs) break;
          const unsigned long long other = __shfl_sync(0xffffffffu, c[ag], from);
          if (taker) c[ag] = gb_merge(c[ag], other, a.op[ag]);
        }
      }
      if (valid && rank == 0) {
        if (private_replica || a.naggs == 1) {
#pragma unroll
          for (int ag = 0; ag < kGbMaxAggs; ag++) {
            if (ag >= a.naggs) break;
            unsigned long long *st = &state[ag * kGbSlots + e];
            if (private_replica) *st = gb_merge(*st, c[ag], a.op[ag]);
            else gb_update_enc(st, a.op[ag], c[ag]);
          }
        } else {
          // Several aggregates on a replica that other warps share: ONE returning atomic (the entry's lock) and plain
          // read-modify-writes under it, instead of one atomic or compare-and-swap loop per aggregate -- returning
          // shared-memory atomics cost the SM ~2 cycles per lane each, which at four per row was the kernel's bound
          // (3.8 ms for 125M rows). Lanes of one warp never want the same lock here (one leader per group), and
          // independent thread scheduling lets a waiting lane spin while the holder finishes.
          bool done = false;
          while (!done) {
            if (atomicExch(&locks[e], 1u) == 0u) {
#pragma unroll
              for (int ag = 0; ag < kGbMaxAggs; ag++) {
                if (ag >= a.naggs) break;
                volatile unsigned long long *st = &state[ag * kGbSlots + e];
                *st = gb_merge(*st, c[ag], a.op[ag]);
              }
              __threadfence_block();
              *reinterpret_cast<volatile unsigned int *>(&locks[e]) = 0u;
              done = true;
            }
          }
        }
      }
    };
#pragma unroll
    for (int k = 0; k < kGbKeep; k++) {
      if (kept_slot[k] == -1) *a.overflow = 1;
      if (__any_sync(0xffffffffu, kept_slot[k] >= 0)) fold(kept[k], kept_slot[k]);
    }
    for (int64_t rb = r0 + (int64_t)kGbKeep * kGbThreads + (tid & ~31); rb < r1; rb += kGbThreads) {  // warp-uniform trip count
      const int64_t r = rb + lane;
      uint64_t v[W];
      int slot = -1;
      if (r < r1) {
        load_row(r, v);
        slot = find_slot(v[0], false);
      } else {
#pragma unroll
        for (int w = 0; w < W; w++) v[w] = 0;
      }
      fold(v, slot);
    }
    __syncthreads();

    // ---- emit: merge the replicas, write the columns, clean the table (the bin's output range was reserved after pass 1)
    if (tid == 0) s_base = base_reg;
    __syncthreads();
    const bool fits = (int64_t)(s_base + ng) <= a.capacity;
    for (unsigned int g = tid; g < ng; g += kGbThreads) {
      const int slot = slot_of[g];
      const int64_t o = (int64_t)s_base + g;
      if (fits) a.out_key[o] = slot == kGbSlots ? kGbEmpty : keys[slot];
      keys[slot] = kGbEmpty;
#pragma unroll
      for (int ag = 0; ag < kGbMaxAggs; ag++) {
        if (ag >= a.naggs) break;
        unsigned long long acc = state[ag * kGbSlots + g];
        for (unsigned int rp = 1; rp < R; rp++) acc = gb_merge(acc, state[ag * kGbSlots + rp * ng + g], a.op[ag]);
        if (fits) a.out_agg[ag][o] = (a.op[ag] == GB_MIN || a.op[ag] == GB_MAX) ? gb_decode(acc, a.enc[ag]) : acc;
      }
    }
    __syncthreads();
    for (int ag = 0; ag < a.naggs; ag++)
      for (unsigned int i = tid; i < R * ng; i += kGbThreads) state[ag * kGbSlots + i] = a.init[ag];
    if (tid == 0) {
      s_ng = 0;
      s_special = 0;
      s_bin = next_bin;
    }
  }
}

static size_t gb_smem_bytes(int naggs, int slots) {
  return (size_t)(slots + 2) * 8 + (size_t)slots * 8 * naggs + (size_t)(slots + 2) * 2 * 2 + (size_t)slots * 4 + 16;
}

template <int W, int THREADS>
static int launch_groupby_cfg(const GbArgs &a, int grid, cudaStream_t s) {
  static PerDeviceOnce attr_once;
  if (auto once_guard = attr_once.first()) CYL_CUDA(cudaFuncSetAttribute(groupby_fused8_kernel<W, THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024));
  groupby_fused8_kernel<W, THREADS><<<grid, THREADS, gb_smem_bytes(a.naggs, a.slots), s>>>(a);
  CYL_LAUNCH_CHECK();
  return CYL_OK;
}

template <int W>
static int launch_groupby(const GbArgs &a, int grid, cudaStream_t s) {
  switch (g_gb_threads) {
    case 1024: return launch_groupby_cfg<W, 1024>(a, grid, s);
    case 512: return launch_groupby_cfg<W, 512>(a, grid, s);
    default: return launch_groupby_cfg<W, 256>(a, grid, s);
  }
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

  // Two attempts. First LARGE bins (~120k rows): typical group-bys have far fewer groups than rows, a bin's groups then fit
  // the table (1M groups over 125M rows: 1024 bins of ~1000 groups each) and the per-bin costs -- barriers, table reset,
  // output reservation -- are paid 1024 times instead of 65536 (measured: aggregate 3.8 ms against 5.9 ms). If any bin's
  // groups overflow the table the attempt is abandoned (kernels exit early on the flag) and repeated with bins of
  // <= 0.47 * slots rows, which cannot hold more groups than that.
  for (int attempt = 0; attempt < 2; attempt++) {
  int64_t nbins = 1;
  const int64_t rows_per_bin = attempt == 0 ? 30 * (int64_t)g_gb_slots : (int64_t)(g_gb_slots * 47 / 100);
  while (nbins * rows_per_bin < n && nbins < (1ll << 22)) nbins *= 2;
  if (g_gb_bins_log2 > 0) nbins = 1ll << g_gb_bins_log2;
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
  a.slots = g_gb_slots;
  CYL_TRY(counters.alloc(64, s));
  CYL_CUDA(cudaMemsetAsync(counters.p, 0, 64, s));
  a.out_cursor = counters.as<unsigned long long>();
  a.work_counter = reinterpret_cast<unsigned int *>(counters.as<char>() + 16);
  a.overflow = reinterpret_cast<int *>(counters.as<char>() + 32);
  int64_t per_sm = (200 * 1024) / (int64_t)(gb_smem_bytes(naggs, g_gb_slots) + 1024);
  if (per_sm > 1024 / g_gb_threads) per_sm = 1024 / g_gb_threads;
  if (per_sm < 1) per_sm = 1;
  int64_t grid = (int64_t)num_sms() * per_sm;
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
  if (reinterpret_cast<int *>(h)[8]) {
    if (attempt == 0 && g_gb_bins_log2 == 0) continue;  // too many groups for large bins: small bins next
    return fail(CYL_CapacityError, "cyl_groupby_fused: a bin held more groups than the shared-memory table has slots");
  }
  *num_groups = (int64_t)h[0];
  return CYL_OK;
  }
  return fail(CYL_CapacityError, "cyl_groupby_fused: no bin size fitted");
}

}  // extern "C"
