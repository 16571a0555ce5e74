// join_fused.cu -- radix-partitioned, shared-memory hash join that materialises the result directly.
//
// The fast path of cylon::Join / DistributedJoin (INNER, one fixed-width key of <= 8 bytes, fixed-width payload
// columns without nulls). It replaces the same reference code as join_hash.cu + gather.cu
// (join/hash_join.cpp:188-307 ArrayIndexHashJoin + join/join_utils.cpp:126-181 build_final_table +
// util/copy_arrray.cpp:24-50) but is organised around the B200 memory system instead of around index vectors:
//
//   1. both tables are radix-partitioned on the top bits of mix64(key) into K bins of ~3k build rows
//      (histogram by global reductions -> device scan -> scatter of every column with an atomic cursor per bin;
//      126 MB of L2 merges the K write frontiers into full lines before they reach HBM);
//   2. one CTA per bin loads the bin's build rows (key AND payload columns) into shared memory, links them into a
//      chained hash table there (32-bit heads, 16-bit next links), then streams the bin's probe rows, counts matches,
//      reserves output rows with one global atomicAdd per 512-row tile and writes every output column straight from
//      shared memory / the probe stream -- no row-index vectors, no gather pass, no hash table in HBM.
//
// HBM traffic per join ~= keys once (histograms) + 2 x tables (scatter) + tables + result: about 2.25x the algorithmic
// bytes, all of it sequential, versus ~16x random-sector traffic for the global table of join_hash.cu.
// The result size is not known before the probe, so the caller passes a capacity (optimistic materialisation): the
// kernel always returns the exact row count and the caller repeats the call with a larger capacity if it did not fit.
// Bins larger than the shared-memory budget (key skew) are processed in several build chunks, which is exact for an
// inner join. Output order is unspecified (row multiset == reference).
#include <memory>

#include "common.cuh"
#include "join_fused.cuh"
#include "scan.cuh"

namespace cylb {

constexpr int kFusedMaxCols = 8;
constexpr int kFusedThreads = 512;
constexpr uint32_t kFusedNil = 0xFFFFu;

struct FusedSide {
  const uint8_t *src[kFusedMaxCols];  // partition-ordered columns
  uint8_t *out[kFusedMaxCols];        // output columns (capacity rows each)
  int32_t width[kFusedMaxCols];
  int32_t smem_off[kFusedMaxCols];    // build side only: byte offset of the column's staging array in shared memory
  int32_t n;
  int32_t key;       // index of the key column
  int32_t key_type;  // cyl_type_t of the key
};

struct FusedArgs {
  FusedSide b, p;             // build / probe side
  const int64_t *b_off;       // [K + 1] bin boundaries in the partition-ordered build table
  const int64_t *p_off;       // [K + 1]
  int64_t nbins;
  int32_t cap_rows;           // build rows staged per chunk
  int32_t nheads;             // power of two
  int32_t heads_off, next_off;
  int64_t capacity;           // output rows available
  unsigned long long *out_cursor;
  unsigned int *work_counter;
};

// key bits as the join compares them: zero-extended raw bits (floating-point identity = bit pattern, see common.cuh)
__device__ __forceinline__ uint64_t canon_key(const uint8_t *base, int64_t i, int32_t width, int32_t type) {
  uint64_t v;
  switch (width) {
    case 1: v = base[i]; break;
    case 2: v = reinterpret_cast<const uint16_t *>(base)[i]; break;
    case 4: v = reinterpret_cast<const uint32_t *>(base)[i]; break;
    default: v = reinterpret_cast<const uint64_t *>(base)[i]; break;
  }
  (void)type;
  return v;
}

__device__ __forceinline__ uint32_t bin_of(uint64_t key, int shift) { return shift >= 64 ? 0u : (uint32_t)(mix64(key) >> shift); }

__device__ __forceinline__ void copy_value(uint8_t *dst, int64_t d, const uint8_t *src, int64_t s, int32_t w) {
  switch (w) {
    case 8: reinterpret_cast<uint64_t *>(dst)[d] = reinterpret_cast<const uint64_t *>(src)[s]; break;
    case 4: reinterpret_cast<uint32_t *>(dst)[d] = reinterpret_cast<const uint32_t *>(src)[s]; break;
    case 2: reinterpret_cast<uint16_t *>(dst)[d] = reinterpret_cast<const uint16_t *>(src)[s]; break;
    default: dst[d] = src[s]; break;
  }
}

// ---- partition pass ------------------------------------------------------------------------------------------
struct FusedPartArgs {
  const uint8_t *key;
  int32_t key_width, key_type;
  int64_t n;
  int32_t shift;
  unsigned int *hist;             // [K]        (histogram pass)
  unsigned long long *cursor;     // [K]        (scatter pass; starts at the bin offsets)
  const uint8_t *src[kFusedMaxCols];
  uint8_t *dst[kFusedMaxCols];
  int32_t width[kFusedMaxCols];
  int32_t ncols;
};

constexpr int kPartRowsPerThread = 4;

__global__ void __launch_bounds__(256) fused_hist_kernel(const __grid_constant__ FusedPartArgs a) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < a.n; i0 += stride * kPartRowsPerThread) {
    uint32_t bin[kPartRowsPerThread];
#pragma unroll
    for (int u = 0; u < kPartRowsPerThread; u++) {
      const int64_t i = i0 + u * stride;
      bin[u] = i < a.n ? bin_of(canon_key(a.key, i, a.key_width, a.key_type), a.shift) : 0xFFFFFFFFu;
    }
#pragma unroll
    for (int u = 0; u < kPartRowsPerThread; u++)
      if (bin[u] != 0xFFFFFFFFu) atomicAdd(&a.hist[bin[u]], 1u);
  }
}

__global__ void __launch_bounds__(256) fused_scatter_kernel(const __grid_constant__ FusedPartArgs a) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < a.n; i0 += stride * kPartRowsPerThread) {
    int64_t pos[kPartRowsPerThread];
#pragma unroll
    for (int u = 0; u < kPartRowsPerThread; u++) {
      const int64_t i = i0 + u * stride;
      pos[u] = -1;
      if (i < a.n) {
        const uint32_t bin = bin_of(canon_key(a.key, i, a.key_width, a.key_type), a.shift);
        pos[u] = (int64_t)atomicAdd(&a.cursor[bin], 1ull);
      }
    }
    for (int c = 0; c < a.ncols; c++) {
#pragma unroll
      for (int u = 0; u < kPartRowsPerThread; u++) {
        const int64_t i = i0 + u * stride;
        if (pos[u] >= 0) copy_value(a.dst[c], pos[u], a.src[c], i, a.width[c]);
      }
    }
  }
}

// ---- join pass ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kFusedThreads) fused_join_kernel(const __grid_constant__ FusedArgs a) {
  extern __shared__ __align__(16) unsigned char smem[];
  __shared__ unsigned int s_bin;
  __shared__ unsigned int s_warp_tot[kFusedThreads / 32];
  __shared__ unsigned long long s_base;
  uint32_t *heads = reinterpret_cast<uint32_t *>(smem + a.heads_off);
  uint16_t *next = reinterpret_cast<uint16_t *>(smem + a.next_off);
  const uint8_t *bkey_s = smem + a.b.smem_off[a.b.key];
  const int bkw = a.b.width[a.b.key], pkw = a.p.width[a.p.key];
  const uint32_t hmask = (uint32_t)a.nheads - 1;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  while (true) {
    __syncthreads();  // previous bin fully done before s_bin / shared tables are reused
    if (tid == 0) s_bin = atomicAdd(a.work_counter, 1u);
    __syncthreads();
    const int64_t bin = s_bin;
    if (bin >= a.nbins) break;
    const int64_t b0 = a.b_off[bin], b1 = a.b_off[bin + 1];
    const int64_t p0 = a.p_off[bin], p1 = a.p_off[bin + 1];
    if (b1 == b0 || p1 == p0) continue;

    for (int64_t c0 = b0; c0 < b1; c0 += a.cap_rows) {  // build chunks (more than one only under key skew)
      const int nb = (int)((b1 - c0) < a.cap_rows ? (b1 - c0) : a.cap_rows);
      __syncthreads();
      for (int h = tid; h < a.nheads; h += kFusedThreads) heads[h] = kFusedNil;
      for (int c = 0; c < a.b.n; c++) {  // stage every build column of the chunk in shared memory
        uint8_t *dst = smem + a.b.smem_off[c];
        const int w = a.b.width[c];
        for (int i = tid; i < nb; i += kFusedThreads) copy_value(dst, i, a.b.src[c], c0 + i, w);
      }
      __syncthreads();
      for (int i = tid; i < nb; i += kFusedThreads) {  // link the chunk's rows into the chained table
        const uint64_t k = canon_key(bkey_s, i, bkw, a.b.key_type);
        const uint32_t h = (uint32_t)mix64(k) & hmask;
        next[i] = (uint16_t)atomicExch(&heads[h], (uint32_t)i);
      }
      __syncthreads();

      for (int64_t t0 = p0; t0 < p1; t0 += kFusedThreads) {  // probe tiles
        const int64_t r = t0 + tid;
        uint64_t k = 0;
        uint32_t first = kFusedNil;
        unsigned int cnt = 0;
        if (r < p1) {
          k = canon_key(a.p.src[a.p.key], r, pkw, a.p.key_type);
          // a NaN key joins nothing (operator== in the reference's single-key comparator); NaN build rows are linked but
          // can only be reached by an identical NaN probe key, which stops here
          uint32_t j = key_bits_is_nan(a.p.key_type, k) ? kFusedNil : heads[(uint32_t)mix64(k) & hmask];
          for (; j != kFusedNil; j = next[j]) {
            if (canon_key(bkey_s, j, bkw, a.b.key_type) == k) {
              if (cnt == 0) first = j;
              cnt++;
            }
          }
        }
        // block-wide exclusive scan of cnt -> this thread's first output slot inside the tile
        unsigned int inc = cnt;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
          const unsigned int n = __shfl_up_sync(0xffffffffu, inc, d);
          if (lane >= d) inc += n;
        }
        if (lane == 31) s_warp_tot[warp] = inc;
        __syncthreads();
        unsigned int wbase = 0, total = 0;
#pragma unroll
        for (int w = 0; w < kFusedThreads / 32; w++) {
          const unsigned int t = s_warp_tot[w];
          if (w < warp) wbase += t;
          total += t;
        }
        if (tid == 0 && total) s_base = atomicAdd(a.out_cursor, (unsigned long long)total);
        __syncthreads();
        if (total && cnt) {
          int64_t o = (int64_t)s_base + wbase + inc - cnt;
          if ((int64_t)s_base + total <= a.capacity) {
            for (uint32_t j = first; j != kFusedNil; j = next[j]) {
              if (j != first && canon_key(bkey_s, j, bkw, a.b.key_type) != k) continue;
              for (int c = 0; c < a.b.n; c++) copy_value(a.b.out[c], o, smem + a.b.smem_off[c], j, a.b.width[c]);
              for (int c = 0; c < a.p.n; c++) copy_value(a.p.out[c], o, a.p.src[c], r, a.p.width[c]);
              o++;
            }
          }
        }
      }
    }
  }
}

static inline int grid1d(int64_t n, int per_block) {
  int64_t b = div_up(n, per_block);
  const int64_t cap = (int64_t)num_sms() * 8;
  return (int)(b < 1 ? 1 : (b < cap ? b : cap));
}

// Validity bitmaps do not disqualify a column: the join's result carries no source validity (copy_arrray.cpp:39-47
// appends the value slot of every matched row) and keys are hashed and compared by their value slot, so
// an INNER join never reads a validity bit.
static bool fused_type_ok(const cyl_column_t &c) {
  switch (c.type) {
    case CYL_UINT8: case CYL_INT8: return c.byte_width == 1;
    case CYL_UINT16: case CYL_INT16: return c.byte_width == 2;
    case CYL_UINT32: case CYL_INT32: case CYL_FLOAT: case CYL_DATE32: case CYL_TIME32: return c.byte_width == 4;
    case CYL_UINT64: case CYL_INT64: case CYL_DOUBLE: case CYL_DATE64: case CYL_TIMESTAMP: case CYL_TIME64: return c.byte_width == 8;
    default: return false;
  }
}

struct FusedPlan {
  int cap_rows, nheads, heads_off, next_off, smem_bytes;
  int smem_off[kFusedMaxCols];
};

// shared-memory layout for a build side with the given column widths
static bool plan_smem(const cyl_column_t *bcols, int nb_cols, FusedPlan *pl) {
  int row_bytes = 0;
  for (int c = 0; c < nb_cols; c++) row_bytes += bcols[c].byte_width;
  const int budget = 100 * 1024;  // two CTAs per SM
  int cap = budget / (row_bytes + 2 + 16);  // 16 = worst-case bytes of heads per row (nheads < 4 * cap)
  cap = cap / 256 * 256;
  if (cap > 16384) cap = 16384;  // next links are 16-bit, 0xFFFF is the end marker
  if (cap < 256) return false;
  int nheads = 256;
  while (nheads < cap * 2) nheads *= 2;
  int off = 0;
  for (int c = 0; c < nb_cols; c++) {
    pl->smem_off[c] = off;
    off += (cap * bcols[c].byte_width + 15) / 16 * 16;
  }
  pl->next_off = off;
  off += (cap * 2 + 15) / 16 * 16;
  pl->heads_off = off;
  off += nheads * 4;
  pl->cap_rows = cap;
  pl->nheads = nheads;
  pl->smem_bytes = off;
  return off <= 110 * 1024;
}

static int partition_side(const cyl_column_t *cols, int ncols, int key, int64_t n, int64_t nbins, int shift, uint8_t *const *dst,
                          int64_t *bin_off, cudaStream_t s) {
  FusedPartArgs a;
  memset(&a, 0, sizeof(a));
  a.key = static_cast<const uint8_t *>(cols[key].data);
  a.key_width = cols[key].byte_width;
  a.key_type = cols[key].type;
  a.n = n;
  a.shift = shift;
  a.ncols = ncols;
  for (int c = 0; c < ncols; c++) {
    a.src[c] = static_cast<const uint8_t *>(cols[c].data);
    a.dst[c] = dst[c];
    a.width[c] = cols[c].byte_width;
  }
  Scratch hist, cursor;
  CYL_TRY(hist.alloc((size_t)nbins * sizeof(unsigned int), s));
  CYL_TRY(cursor.alloc((size_t)(nbins + 1) * sizeof(unsigned long long), s));
  CYL_CUDA(cudaMemsetAsync(hist.p, 0, (size_t)nbins * sizeof(unsigned int), s));
  a.hist = hist.as<unsigned int>();
  a.cursor = cursor.as<unsigned long long>();
  if (n > 0) {
    ProfScope ps("fused_join_hist", s);
    fused_hist_kernel<<<grid1d(n, 256 * kPartRowsPerThread), 256, 0, s>>>(a);
    CYL_LAUNCH_CHECK();
  }
  CYL_TRY(exclusive_scan<uint32_t>(hist.as<unsigned int>(), nbins, bin_off, s));
  CYL_CUDA(cudaMemcpyAsync(cursor.p, bin_off, (size_t)nbins * sizeof(int64_t), cudaMemcpyDeviceToDevice, s));
  if (n > 0) {
    ProfScope ps("fused_join_scatter", s);
    fused_scatter_kernel<<<grid1d(n, 256 * kPartRowsPerThread), 256, 0, s>>>(a);
    CYL_LAUNCH_CHECK();
  }
  return CYL_OK;
}

}  // namespace cylb

using namespace cylb;

extern "C" {

int cyl_hash_join_fused_supported(const cyl_column_t *lcols, int32_t nlcols, int32_t lkey, const cyl_column_t *rcols, int32_t nrcols,
                                  int32_t rkey, int32_t join_type) {
  if (join_type != CYL_JOIN_INNER) return 0;
  if (nlcols < 1 || nrcols < 1 || nlcols > kFusedMaxCols || nrcols > kFusedMaxCols) return 0;
  if (lkey < 0 || lkey >= nlcols || rkey < 0 || rkey >= nrcols) return 0;
  for (int c = 0; c < nlcols; c++)
    if (!fused_type_ok(lcols[c])) return 0;
  for (int c = 0; c < nrcols; c++)
    if (!fused_type_ok(rcols[c])) return 0;
  if (lcols[lkey].type != rcols[rkey].type) return 0;
  FusedPlan pl;
  return plan_smem(lcols, nlcols, &pl) && plan_smem(rcols, nrcols, &pl) ? 1 : 0;
}

// ---- the all-8-byte fused join in two steps (see include/cylon_b200.h) ----
struct cyl_join_side {
  F8Partitioned *p = nullptr;
  int ncols = 0;
};

int cyl_hash_join_fused_partition(const cyl_column_t *cols, int32_t ncols, int32_t key, int64_t nrows, const cyl_join_side_t *like, int64_t build_rows_hint,
                                  cyl_join_side_t **side, void *stream) {
  if (!side) return fail(CYL_Invalid, "cyl_hash_join_fused_partition: null output");
  *side = nullptr;
  if (ncols < 1 || ncols > kFused8MaxCols || key < 0 || key >= ncols) return fail(CYL_Invalid, "cyl_hash_join_fused_partition: bad column / key count");
  for (int c = 0; c < ncols; c++)
    if (cols[c].byte_width != 8 || cols[c].length != nrows) return fail(CYL_Invalid, "cyl_hash_join_fused_partition: column %d is not an 8-byte column of %lld rows", c, (long long)nrows);
  if (nrows < 0 || nrows > 0xFFFFFFFEll) return fail(CYL_CapacityError, "cyl_hash_join_fused_partition: row count out of range");
  int64_t nbins;
  int shift;
  if (like) {
    nbins = fused8_side_bins(like->p);
    shift = fused8_side_shift(like->p);
  } else {
    fused8_bins_for(ncols, build_rows_hint > nrows ? build_rows_hint : nrows, &nbins, &shift);
  }
  std::unique_ptr<cyl_join_side> out(new cyl_join_side());
  out->ncols = ncols;
  CYL_TRY(fused8_partition_side(cols, ncols, key, nrows, nbins, shift, &out->p, (cudaStream_t)stream));
  *side = out.release();
  return CYL_OK;
}

int cyl_hash_join_fused_probe(const cyl_join_side_t *left, const cyl_join_side_t *right, int64_t capacity, void *const *out_cols, int64_t *num_rows,
                              void *stream) {
  if (!left || !right || !num_rows) return fail(CYL_Invalid, "cyl_hash_join_fused_probe: null argument");
  // build side = the table that was partitioned first decides nothing here: either works; take the left (ties -> left, hash_join.cpp:226-231)
  return fused8_probe(left->p, right->p, out_cols, out_cols + left->ncols, capacity, num_rows, (cudaStream_t)stream);
}

void cyl_hash_join_fused_side_free(cyl_join_side_t *side) {
  if (!side) return;
  fused8_side_free(side->p);
  delete side;
}

int cyl_tune_set(int32_t key, int32_t value) {
  fused8_tune(key, value);
  return CYL_OK;
}

int cyl_hash_join_fused_set_bins_log2(int32_t bits) {
  if (bits < 0 || bits > 18) return fail(CYL_Invalid, "cyl_hash_join_fused_set_bins_log2: bits must be in [0, 18], got %d", bits);
  fused8_force_bins_log2(bits);
  return CYL_OK;
}

int cyl_hash_join_fused_inner(const cyl_column_t *lcols, int32_t nlcols, int32_t lkey, const cyl_column_t *rcols, int32_t nrcols, int32_t rkey,
                              int64_t nl, int64_t nr, int64_t capacity, void *const *out_cols, int64_t *num_rows, void *stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (!num_rows) return fail(CYL_Invalid, "cyl_hash_join_fused_inner: null output");
  if (!cyl_hash_join_fused_supported(lcols, nlcols, lkey, rcols, nrcols, rkey, CYL_JOIN_INNER))
    return fail(CYL_NotImplemented, "cyl_hash_join_fused_inner: needs an INNER join on one fixed-width key, fixed-width columns without nulls, "
                                    "at most %d columns per side", kFusedMaxCols);
  if (nl < 0 || nr < 0 || nl > 0xFFFFFFFEll || nr > 0xFFFFFFFEll) return fail(CYL_CapacityError, "cyl_hash_join_fused_inner: row count out of range");
  for (int c = 0; c < nlcols; c++)
    if (lcols[c].length != nl) return fail(CYL_Invalid, "cyl_hash_join_fused_inner: left column %d length mismatch", c);
  for (int c = 0; c < nrcols; c++)
    if (rcols[c].length != nr) return fail(CYL_Invalid, "cyl_hash_join_fused_inner: right column %d length mismatch", c);
  *num_rows = 0;
  if (nl == 0 || nr == 0) return CYL_OK;

  // build side = the smaller table, ties left (join/hash_join.cpp:226-231)
  const bool build_left = nl <= nr;
  const cyl_column_t *bcols = build_left ? lcols : rcols, *pcols = build_left ? rcols : lcols;
  const int nbc = build_left ? nlcols : nrcols, npc = build_left ? nrcols : nlcols;
  const int bkey = build_left ? lkey : rkey, pkey = build_left ? rkey : lkey;
  const int64_t nb = build_left ? nl : nr, np = build_left ? nr : nl;
  void *const *bout = build_left ? out_cols : out_cols + nlcols;
  void *const *pout = build_left ? out_cols + nlcols : out_cols;

  if (fused8_eligible(lcols, nlcols, rcols, nrcols))  // all columns 8 bytes wide: packed-record kernels (join_fused8.cu)
    return fused8_join(bcols, nbc, bkey, nb, pcols, npc, pkey, np, bout, pout, capacity, num_rows, s);

  FusedPlan pl;
  plan_smem(bcols, nbc, &pl);
  int64_t nbins = 1;
  while (nbins * (int64_t)(pl.cap_rows * 8 / 10) < nb && nbins < (1ll << 22)) nbins *= 2;
  int bits = 0;
  while ((1ll << bits) < nbins) bits++;
  const int shift = 64 - bits;

  // partition-ordered copies of both tables
  Scratch bpart[kFusedMaxCols], ppart[kFusedMaxCols], boff, poff, counters;
  uint8_t *bdst[kFusedMaxCols], *pdst[kFusedMaxCols];
  for (int c = 0; c < nbc; c++) {
    CYL_TRY(bpart[c].alloc((size_t)nb * bcols[c].byte_width, s));
    bdst[c] = bpart[c].as<uint8_t>();
  }
  for (int c = 0; c < npc; c++) {
    CYL_TRY(ppart[c].alloc((size_t)np * pcols[c].byte_width, s));
    pdst[c] = ppart[c].as<uint8_t>();
  }
  CYL_TRY(boff.alloc((size_t)(nbins + 1) * sizeof(int64_t), s));
  CYL_TRY(poff.alloc((size_t)(nbins + 1) * sizeof(int64_t), s));
  CYL_TRY(partition_side(bcols, nbc, bkey, nb, nbins, shift, bdst, boff.as<int64_t>(), s));
  CYL_TRY(partition_side(pcols, npc, pkey, np, nbins, shift, pdst, poff.as<int64_t>(), s));

  FusedArgs a;
  memset(&a, 0, sizeof(a));
  a.b.n = nbc;
  a.b.key = bkey;
  a.b.key_type = bcols[bkey].type;
  for (int c = 0; c < nbc; c++) {
    a.b.src[c] = bdst[c];
    a.b.out[c] = static_cast<uint8_t *>(bout[c]);
    a.b.width[c] = bcols[c].byte_width;
    a.b.smem_off[c] = pl.smem_off[c];
  }
  a.p.n = npc;
  a.p.key = pkey;
  a.p.key_type = pcols[pkey].type;
  for (int c = 0; c < npc; c++) {
    a.p.src[c] = pdst[c];
    a.p.out[c] = static_cast<uint8_t *>(pout[c]);
    a.p.width[c] = pcols[c].byte_width;
  }
  a.b_off = boff.as<int64_t>();
  a.p_off = poff.as<int64_t>();
  a.nbins = nbins;
  a.cap_rows = pl.cap_rows;
  a.nheads = pl.nheads;
  a.heads_off = pl.heads_off;
  a.next_off = pl.next_off;
  a.capacity = capacity;
  CYL_TRY(counters.alloc(64, s));
  CYL_CUDA(cudaMemsetAsync(counters.p, 0, 64, s));
  a.out_cursor = counters.as<unsigned long long>();
  a.work_counter = reinterpret_cast<unsigned int *>(counters.as<char>() + 16);

  static PerDeviceOnce attr_once;
  if (auto once_guard = attr_once.first()) {
    CYL_CUDA(cudaFuncSetAttribute(fused_join_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024));
  }
  int64_t grid = (int64_t)num_sms() * 2;
  if (grid > nbins) grid = nbins;
  {
    ProfScope ps("fused_join_probe", s);
    fused_join_kernel<<<(unsigned)grid, kFusedThreads, pl.smem_bytes, s>>>(a);
    CYL_LAUNCH_CHECK();
  }
  unsigned long long m = 0;
  CYL_CUDA(cudaMemcpyAsync(&m, a.out_cursor, sizeof(m), cudaMemcpyDeviceToHost, s));
  CYL_CUDA(cudaStreamSynchronize(s));
  *num_rows = (int64_t)m;
  return CYL_OK;
}

}  // extern "C"
