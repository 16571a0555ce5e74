// join_fused8x.cu -- the fused radix-partitioned join for everything the INNER / one-key kernel of join_fused8.cu leaves
// out: LEFT, RIGHT and FULL_OUTER joins, keys of two columns, and tables that mix 4- and 8-byte columns (BASELINE config
// 3: an (int64, int32) key with a double payload, Zipf-skewed left side, left / right / full-outer).
//
// Replaces (reference, relative to cpp/src/cylon/): join/hash_join.cpp:21-90 (the three probe functions), :92-186
// (multi_index_hash_join), :188-307 (ArrayIndexHashJoin), join/sort_join.cpp:396-556 (do_multi_index_sorted_join; its
// result is the same row multiset) and the build_final_table / copy_array_by_indices pair that materialises them
// (join/join_utils.cpp:126-181, util/copy_arrray.cpp:24-50) -- index -1 becomes a null, source validity is not carried.
//
// Same engine as join_fused8.cu: both tables are radix-partitioned on a hash of the key into bins whose build side fits
// shared memory (f8_partition: histogram + two staged passes), then one CTA joins a bin and writes result rows directly.
// What is added around it:
//   * a pre-pass turns the table into all-8-byte "word columns": 4-byte columns are widened, and a two-column key gets
//     one extra word h = mix(k0, k1) in front, which is what the partition hashes and the chains bucket on; equality
//     always compares the real key words, so h only has to spread rows;
//   * the probe side is the preserved side of a LEFT / RIGHT join (as in the reference, hash_join.cpp:207-232): a probe
//     row without a partner emits (nulls, probe row); for FULL_OUTER every build row remembers in shared memory whether
//     it was matched and the unmatched ones are emitted after the bin's probe rows have gone through;
//   * a bin whose build side does not fit (one hot key) is processed in build chunks; the "did this probe row match in
//     ANY chunk" bit then lives in a global byte per probe row and the unmatched probe rows are emitted after the last
//     chunk;
//   * output rows carry one validity byte per row and nullable side, packed into Arrow validity bitmaps by the caller.
#include "common.cuh"
#include "join_fused.cuh"
#include "scan.cuh"

namespace cylb {

constexpr int kXThreads = 512;
constexpr int kXProbeRows = 4;
constexpr uint32_t kXNil = 0xFFFFu;
constexpr int kXMaxOut = kFused8MaxCols;

struct F8XSide {
  int32_t nout;                 // output columns of this side
  int32_t word[kXMaxOut];       // record word each output column takes
  int32_t narrow[kXMaxOut];     // 1: the column is 4 bytes wide (the word's low half)
  uint8_t *out[kXMaxOut];
  uint8_t *null_flag;           // [capacity] validity byte of this side per output row: 1 = present, 0 = null (nullptr: never null)
  int32_t kw[2];                // record words holding the key columns
};

struct F8XArgs {
  const uint64_t *brec, *prec;
  const int64_t *b_off, *p_off;  // [nbins + 1]
  int64_t nbins;
  F8XSide b, p;
  int32_t nkw;                   // key words to compare (1 or 2)
  int32_t bucket_word_b, bucket_word_p;  // word the chains bucket on (the hash word for two-column keys, else the key)
  int32_t nan_type;              // single FLOAT / DOUBLE key: its type id (a NaN key matches nothing), else 0
  int32_t keep_probe, keep_build;
  int32_t cap_rows, nheads;
  int64_t capacity;
  unsigned long long *out_cursor;
  unsigned int *work_counter;
  uint8_t *pflag;                // [np] probe row matched in some build chunk (multi-chunk bins of outer joins)
  const int64_t *slice_off;      // [nbins + 1] work items before each bin: a bin with many probe rows is cut into slices
  uint8_t *bflag;                // [nb] FULL_OUTER, bins cut into slices: build row matched by some slice
  unsigned int *bin_ticket;      // [nbins] FULL_OUTER: slices of the bin that have finished
};

// Probe rows one work item covers. A bin normally holds a few thousand probe rows (one item); a hot key on the probe side
// puts millions into one bin (config 3: 12 % of the left table), which one CTA would chew through alone -- measured 27.7 ms
// for 50M x 50M rows -- so such a bin becomes many items that each rebuild the bin's small table.
constexpr int64_t kXSliceRows = 16 * kXThreads * kXProbeRows;

// fullest bin of each side -> out[0] (left), out[1] (right)
__global__ void x_maxbin_kernel(const int64_t *__restrict__ l_off, const int64_t *__restrict__ r_off, int64_t nbins, unsigned long long *__restrict__ out) {
  unsigned long long ml = 0, mr = 0;
  for (int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; b < nbins; b += (int64_t)gridDim.x * blockDim.x) {
    const unsigned long long l = (unsigned long long)(l_off[b + 1] - l_off[b]), r = (unsigned long long)(r_off[b + 1] - r_off[b]);
    ml = l > ml ? l : ml;
    mr = r > mr ? r : mr;
  }
  if (ml) atomicMax(&out[0], ml);
  if (mr) atomicMax(&out[1], mr);
}

__global__ void x_slices_kernel(const int64_t *__restrict__ p_off, int64_t nbins, uint32_t *__restrict__ nslices) {
  for (int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; b < nbins; b += (int64_t)gridDim.x * blockDim.x) {
    const int64_t pn = p_off[b + 1] - p_off[b];
    nslices[b] = pn <= kXSliceRows ? 1u : (uint32_t)((pn + kXSliceRows - 1) / kXSliceRows);
  }
}

__device__ __forceinline__ uint32_t x_bucket(uint64_t k) { return (uint32_t)((k * 0x9E3779B97F4A7C15ULL) >> 38); }

template <int W>
__device__ __forceinline__ uint64_t x_pick(const uint64_t (&v)[W], int idx) {
  uint64_t r = v[0];
#pragma unroll
  for (int w = 1; w < W; w++)
    if (w == idx) r = v[w];
  return r;
}

__device__ __forceinline__ void x_store(const F8XSide &s, int c, int64_t o, uint64_t word) {
  if (s.narrow[c]) reinterpret_cast<uint32_t *>(s.out[c])[o] = (uint32_t)word;
  else reinterpret_cast<uint64_t *>(s.out[c])[o] = word;
}

// block-wide exclusive scan of one count per thread; returns this thread's offset, *total = sum (same for every thread)
__device__ __forceinline__ unsigned int x_block_scan(unsigned int c, unsigned int *warp_tot, unsigned int *total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  unsigned int inc = c;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const unsigned int n = __shfl_up_sync(0xffffffffu, inc, d);
    if (lane >= d) inc += n;
  }
  __syncthreads();  // warp_tot may still be read from the previous use
  if (lane == 31) warp_tot[warp] = inc;
  __syncthreads();
  unsigned int before = 0, all = 0;
#pragma unroll
  for (int w = 0; w < kXThreads / 32; w++) {
    const unsigned int t = warp_tot[w];
    before += w < warp ? t : 0u;
    all += t;
  }
  *total = all;
  return before + inc - c;
}

template <int BW, int PW>
__global__ void __launch_bounds__(kXThreads, 2) fused8x_join_kernel(const __grid_constant__ F8XArgs a) {
  extern __shared__ __align__(16) unsigned char smem[];
  __shared__ unsigned int s_bin, s_last;
  __shared__ unsigned int s_warp_tot[kXThreads / 32];
  __shared__ unsigned long long s_base;
  uint64_t *brec = reinterpret_cast<uint64_t *>(smem);                                  // [cap_rows][BW]
  uint32_t *heads = reinterpret_cast<uint32_t *>(smem + (size_t)a.cap_rows * BW * 8);   // [nheads]
  uint16_t *next = reinterpret_cast<uint16_t *>(heads + a.nheads);                      // [cap_rows]
  uint8_t *bmatched = reinterpret_cast<uint8_t *>(next + a.cap_rows);                   // [cap_rows] (FULL_OUTER)
  const uint32_t hmask = (uint32_t)a.nheads - 1;
  const int tid = threadIdx.x;

  auto keys_equal = [&](const uint64_t *b, const uint64_t (&pv)[PW]) -> bool {
    bool eq = b[a.b.kw[0]] == x_pick<PW>(pv, a.p.kw[0]);
    if (a.nkw > 1) eq = eq && b[a.b.kw[1]] == x_pick<PW>(pv, a.p.kw[1]);
    return eq;
  };
  // one result row: build words (or nulls) + probe words (or nulls). The probe record lives in registers: no dynamic index.
  auto emit = [&](int64_t o, const uint64_t *b, const uint64_t (&pv)[PW], bool has_p) {
#pragma unroll
    for (int c = 0; c < kXMaxOut; c++)
      if (c < a.b.nout) x_store(a.b, c, o, b ? b[a.b.word[c]] : 0ull);
#pragma unroll
    for (int c = 0; c < kXMaxOut; c++)
      if (c < a.p.nout) x_store(a.p, c, o, has_p ? x_pick<PW>(pv, a.p.word[c]) : 0ull);
    if (a.b.null_flag) a.b.null_flag[o] = b ? 1 : 0;
    if (a.p.null_flag) a.p.null_flag[o] = has_p ? 1 : 0;
  };

  const int64_t total_items = a.slice_off[a.nbins];
  while (true) {
    __syncthreads();
    if (tid == 0) s_bin = atomicAdd(a.work_counter, 1u);
    __syncthreads();
    const int64_t item = s_bin;
    if (item >= total_items) break;
    int64_t bin;
    {  // last bin whose first item is <= item
      int64_t lo = 0, hi = a.nbins;
      while (hi - lo > 1) {
        const int64_t mid = (lo + hi) >> 1;
        if (a.slice_off[mid] <= item) lo = mid; else hi = mid;
      }
      bin = lo;
    }
    const int64_t slice = item - a.slice_off[bin], nslices = a.slice_off[bin + 1] - a.slice_off[bin];
    const int64_t b0 = a.b_off[bin], b1 = a.b_off[bin + 1];
    int64_t p0 = a.p_off[bin], p1 = a.p_off[bin + 1];
    if (p1 == p0 && !(a.keep_build && b1 > b0)) continue;
    if (b1 == b0 && !(a.keep_probe && p1 > p0)) continue;
    if (nslices > 1) {  // this item's share of the bin's probe rows
      p0 += slice * kXSliceRows;
      if (p1 > p0 + kXSliceRows) p1 = p0 + kXSliceRows;
    }
    const bool multi_chunk = (b1 - b0) > a.cap_rows;
    const bool shared_bin = nslices > 1;

    int64_t c0 = b0;
    do {  // build chunks: more than one only when a hot key overfills the bin; runs once with nb == 0 for an empty build side
      const int nb = (int)((b1 - c0) < a.cap_rows ? (b1 - c0) : a.cap_rows);
      __syncthreads();
      for (int h = tid; h < a.nheads; h += kXThreads) heads[h] = kXNil;
      {
        const uint64_t *src = a.brec + c0 * BW;
        const int nwords = nb * BW;
        if constexpr (BW % 2 == 0) {
          const ulonglong2 *s2 = reinterpret_cast<const ulonglong2 *>(src);
          ulonglong2 *d2 = reinterpret_cast<ulonglong2 *>(brec);
          for (int i = tid; i < nwords / 2; i += kXThreads) d2[i] = s2[i];
        } else {
          for (int i = tid; i < nwords; i += kXThreads) brec[i] = src[i];
        }
      }
      if (a.keep_build)
        for (int i = tid; i < nb; i += kXThreads) bmatched[i] = 0;
      __syncthreads();
      for (int i = tid; i < nb; i += kXThreads) {
        const uint64_t k = brec[i * BW + a.bucket_word_b];
        next[i] = (uint16_t)atomicExch(&heads[x_bucket(k) & hmask], (uint32_t)i);
      }
      __syncthreads();

      for (int64_t t0 = p0; t0 < p1; t0 += kXThreads * kXProbeRows) {
        uint64_t pv[kXProbeRows][PW];
        unsigned int cnt[kXProbeRows], e[kXProbeRows];
#pragma unroll
        for (int u = 0; u < kXProbeRows; u++) {
          const int64_t r = t0 + (int64_t)u * kXThreads + tid;  // coalesced: consecutive lanes read consecutive records
#pragma unroll
          for (int w = 0; w < PW; w++) pv[u][w] = r < p1 ? a.prec[r * PW + w] : 0ull;
        }
        unsigned int mine = 0;
#pragma unroll
        for (int u = 0; u < kXProbeRows; u++) {
          const int64_t r = t0 + (int64_t)u * kXThreads + tid;
          cnt[u] = 0;
          e[u] = 0;
          if (r < p1) {
            const uint64_t k0 = x_pick<PW>(pv[u], a.p.kw[0]);
            const bool nan = a.nan_type && key_bits_is_nan(a.nan_type, k0);  // operator== of the single-key comparator
            uint32_t j = nan ? kXNil : heads[x_bucket(x_pick<PW>(pv[u], a.bucket_word_p)) & hmask];
            for (; j != kXNil; j = next[j]) {
              if (keys_equal(brec + (size_t)j * BW, pv[u])) {
                cnt[u]++;
                if (a.keep_build) bmatched[j] = 1;  // every writer stores 1
              }
            }
            if (cnt[u] && multi_chunk && a.keep_probe) a.pflag[r] = 1;
            e[u] = cnt[u] ? cnt[u] : ((a.keep_probe && !multi_chunk) ? 1u : 0u);
          }
          mine += e[u];
        }
        unsigned int total;
        const unsigned int my_off = x_block_scan(mine, s_warp_tot, &total);
        if (tid == 0 && total) s_base = atomicAdd(a.out_cursor, (unsigned long long)total);  // one reservation per tile
        __syncthreads();
        if (mine && (int64_t)(s_base + total) <= a.capacity) {
          int64_t o = (int64_t)s_base + my_off;
#pragma unroll
          for (int u = 0; u < kXProbeRows; u++) {
            if (e[u] == 0) continue;
            if (cnt[u] == 0) {
              emit(o++, nullptr, pv[u], true);
              continue;
            }
            unsigned int left = cnt[u];
            for (uint32_t j = heads[x_bucket(x_pick<PW>(pv[u], a.bucket_word_p)) & hmask]; left; j = next[j]) {
              if (!keys_equal(brec + (size_t)j * BW, pv[u])) continue;
              emit(o++, brec + (size_t)j * BW, pv[u], true);
              left--;
            }
          }
        }
      }
      if (a.keep_build && nb > 0 && shared_bin) {  // other slices probe this chunk too: publish, the last slice settles it
        __syncthreads();
        for (int i = tid; i < nb; i += kXThreads)
          if (bmatched[i]) a.bflag[c0 + i] = 1;
      } else if (a.keep_build && nb > 0) {  // FULL_OUTER: build rows of this chunk that no probe row matched
        __syncthreads();
        unsigned int mine = 0;
        for (int i = tid; i < nb; i += kXThreads) mine += bmatched[i] ? 0u : 1u;
        unsigned int total;
        const unsigned int my_off = x_block_scan(mine, s_warp_tot, &total);
        if (tid == 0 && total) s_base = atomicAdd(a.out_cursor, (unsigned long long)total);
        __syncthreads();
        if (mine && (int64_t)(s_base + total) <= a.capacity) {
          int64_t o = (int64_t)s_base + my_off;
          const uint64_t none[PW] = {};
          for (int i = tid; i < nb; i += kXThreads)
            if (!bmatched[i]) emit(o++, brec + (size_t)i * BW, none, false);
        }
      }
      c0 += a.cap_rows;
    } while (c0 < b1);

    if (a.keep_build && shared_bin) {  // FULL_OUTER, bin cut into slices: whoever finishes last emits the unmatched build rows
      __threadfence();
      __syncthreads();
      if (tid == 0) s_last = atomicAdd(&a.bin_ticket[bin], 1u) == (unsigned int)(nslices - 1) ? 1u : 0u;
      __syncthreads();
      if (s_last) {
        __threadfence();
        const uint64_t none[PW] = {};
        for (int64_t t0 = b0; t0 < b1; t0 += kXThreads) {
          const int64_t i = t0 + tid;
          const unsigned int mine = (i < b1 && !*reinterpret_cast<volatile uint8_t *>(&a.bflag[i])) ? 1u : 0u;
          unsigned int total;
          const unsigned int my_off = x_block_scan(mine, s_warp_tot, &total);
          if (tid == 0 && total) s_base = atomicAdd(a.out_cursor, (unsigned long long)total);
          __syncthreads();
          if (mine && (int64_t)(s_base + total) <= a.capacity) emit((int64_t)s_base + my_off, a.brec + (size_t)i * BW, none, false);
        }
      }
    }
    if (multi_chunk && a.keep_probe) {  // probe rows that matched in no chunk
      __threadfence();
      __syncthreads();
      for (int64_t t0 = p0; t0 < p1; t0 += kXThreads * kXProbeRows) {
        unsigned int mine = 0;
#pragma unroll
        for (int u = 0; u < kXProbeRows; u++) {
          const int64_t r = t0 + (int64_t)u * kXThreads + tid;
          if (r < p1 && !a.pflag[r]) mine++;
        }
        unsigned int total;
        const unsigned int my_off = x_block_scan(mine, s_warp_tot, &total);
        if (tid == 0 && total) s_base = atomicAdd(a.out_cursor, (unsigned long long)total);
        __syncthreads();
        if (mine && (int64_t)(s_base + total) <= a.capacity) {
          int64_t o = (int64_t)s_base + my_off;
#pragma unroll
          for (int u = 0; u < kXProbeRows; u++) {
            const int64_t r = t0 + (int64_t)u * kXThreads + tid;
            if (r < p1 && !a.pflag[r]) {
              uint64_t pv[PW];
#pragma unroll
              for (int w = 0; w < PW; w++) pv[w] = a.prec[r * PW + w];
              emit(o++, nullptr, pv, true);
            }
          }
        }
      }
    }
  }
}


// ---- pre-pass: 4-byte columns widened to words, two-column keys hashed into one leading word -------------------------
__global__ void __launch_bounds__(256) x_widen32_kernel(const uint32_t *__restrict__ src, int64_t n, uint64_t *__restrict__ dst) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) dst[i] = src[i];
}

__global__ void __launch_bounds__(256) x_hash2_kernel(const uint8_t *__restrict__ k0, int w0, const uint8_t *__restrict__ k1, int w1, int64_t n,
                                                      uint64_t *__restrict__ dst) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const uint64_t a = w0 == 8 ? reinterpret_cast<const uint64_t *>(k0)[i] : reinterpret_cast<const uint32_t *>(k0)[i];
    const uint64_t b = w1 == 8 ? reinterpret_cast<const uint64_t *>(k1)[i] : reinterpret_cast<const uint32_t *>(k1)[i];
    dst[i] = mix64(a) ^ (mix64(b + 0x9E3779B97F4A7C15ULL) * 0xD6E8FEB86659FD93ULL);
  }
}

static inline int x_grid(int64_t n) {
  int64_t b = div_up(n, 256);
  const int64_t cap = (int64_t)num_sms() * 16;
  return (int)(b < 1 ? 1 : (b < cap ? b : cap));
}

static bool x_type_ok(const cyl_column_t &c) {
  switch (c.type) {
    case CYL_UINT32: case CYL_INT32: case CYL_FLOAT: case CYL_DATE32: case CYL_TIME32: return c.byte_width == 4;
    case CYL_UINT64: case CYL_INT64: case CYL_DOUBLE: case CYL_DATE64: case CYL_TIMESTAMP: case CYL_TIME64: return c.byte_width == 8;
    default: return false;
  }
}

// One side of the join as all-8-byte word columns: [hash word (two-column keys)] ++ every table column (widened if needed)
struct XWords {
  cyl_column_t cols[kFused8MaxCols];
  int nwords = 0;
  int col_word[kFused8MaxCols];  // word of table column c
  int key_word[2];
  int hash_word = -1;
  Scratch tmp[kFused8MaxCols];
};

static int x_build_words(const cyl_column_t *cols, int ncols, const int32_t *keys, int nkeys, int64_t n, XWords *out, cudaStream_t s) {
  int w = 0;
  if (nkeys == 2) {
    CYL_TRY(out->tmp[w].alloc((size_t)(n + 2) * 8, s));
    if (n > 0) {
      x_hash2_kernel<<<x_grid(n), 256, 0, s>>>(static_cast<const uint8_t *>(cols[keys[0]].data), cols[keys[0]].byte_width,
                                                static_cast<const uint8_t *>(cols[keys[1]].data), cols[keys[1]].byte_width, n,
                                                out->tmp[w].as<uint64_t>());
      CYL_LAUNCH_CHECK();
    }
    out->cols[w] = cyl_column_t{};
    out->cols[w].type = CYL_UINT64;
    out->cols[w].byte_width = 8;
    out->cols[w].length = n;
    out->cols[w].data = out->tmp[w].p;
    out->hash_word = w;
    w++;
  }
  for (int c = 0; c < ncols; c++, w++) {
    out->col_word[c] = w;
    out->cols[w] = cols[c];
    out->cols[w].validity = nullptr;
    if (cols[c].byte_width == 4) {
      CYL_TRY(out->tmp[w].alloc((size_t)(n + 2) * 8, s));
      if (n > 0) {
        x_widen32_kernel<<<x_grid(n), 256, 0, s>>>(static_cast<const uint32_t *>(cols[c].data), n, out->tmp[w].as<uint64_t>());
        CYL_LAUNCH_CHECK();
      }
      out->cols[w].type = CYL_UINT64;
      out->cols[w].byte_width = 8;
      out->cols[w].data = out->tmp[w].p;
    }
  }
  out->nwords = w;
  for (int k = 0; k < nkeys; k++) out->key_word[k] = out->col_word[keys[k]];
  return CYL_OK;
}

template <int BW, int PW>
static int x_launch(const F8XArgs &a, int grid, size_t smem_bytes, cudaStream_t s) {
  static PerDeviceOnce attr_once;
  if (auto once_guard = attr_once.first()) CYL_CUDA(cudaFuncSetAttribute(fused8x_join_kernel<BW, PW>, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024));
  fused8x_join_kernel<BW, PW><<<grid, kXThreads, smem_bytes, s>>>(a);
  CYL_LAUNCH_CHECK();
  return CYL_OK;
}

template <int BW>
static int x_dispatch_pw(int pw, const F8XArgs &a, int grid, size_t smem_bytes, cudaStream_t s) {
  switch (pw) {
    case 1: return x_launch<BW, 1>(a, grid, smem_bytes, s);
    case 2: return x_launch<BW, 2>(a, grid, smem_bytes, s);
    case 3: return x_launch<BW, 3>(a, grid, smem_bytes, s);
    default: return x_launch<BW, 4>(a, grid, smem_bytes, s);
  }
}

}  // namespace cylb

using namespace cylb;

extern "C" {

int cyl_join_fused_supported(const cyl_column_t *lcols, int32_t nlcols, const int32_t *lkeys, const cyl_column_t *rcols, int32_t nrcols,
                             const int32_t *rkeys, int32_t nkeys, int32_t join_type) {
  if (nkeys < 1 || nkeys > 2 || nlcols < 1 || nrcols < 1) return 0;
  if (join_type < CYL_JOIN_INNER || join_type > CYL_JOIN_FULL_OUTER) return 0;
  const int extra = nkeys == 2 ? 1 : 0;
  if (nlcols + extra > kFused8MaxCols || nrcols + extra > kFused8MaxCols) return 0;
  for (int c = 0; c < nlcols; c++)
    if (!x_type_ok(lcols[c])) return 0;
  for (int c = 0; c < nrcols; c++)
    if (!x_type_ok(rcols[c])) return 0;
  for (int k = 0; k < nkeys; k++) {
    if (lkeys[k] < 0 || lkeys[k] >= nlcols || rkeys[k] < 0 || rkeys[k] >= nrcols) return 0;
    if (lcols[lkeys[k]].type != rcols[rkeys[k]].type) return 0;
  }
  return 1;
}

int cyl_join_fused(const cyl_column_t *lcols, int32_t nlcols, const int32_t *lkeys, const cyl_column_t *rcols, int32_t nrcols, const int32_t *rkeys,
                   int32_t nkeys, int32_t join_type, int64_t capacity, void *const *out_cols, uint8_t *left_valid, uint8_t *right_valid,
                   int64_t *num_rows, void *stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (!num_rows) return fail(CYL_Invalid, "cyl_join_fused: null output");
  *num_rows = 0;
  if (!cyl_join_fused_supported(lcols, nlcols, lkeys, rcols, nrcols, rkeys, nkeys, join_type))
    return fail(CYL_Invalid, "cyl_join_fused: unsupported shape (1-2 key columns, 4/8-byte numeric columns, <= %d words per side)", kFused8MaxCols);
  const int64_t nl = lcols[0].length, nr = rcols[0].length;
  if (nl > 0xFFFFFFFEll || nr > 0xFFFFFFFEll) return fail(CYL_CapacityError, "cyl_join_fused: more than 2^32-2 rows per side");
  const bool left_can_be_null = join_type == CYL_JOIN_RIGHT || join_type == CYL_JOIN_FULL_OUTER;
  const bool right_can_be_null = join_type == CYL_JOIN_LEFT || join_type == CYL_JOIN_FULL_OUTER;
  if ((left_can_be_null && !left_valid) || (right_can_be_null && !right_valid)) return fail(CYL_Invalid, "cyl_join_fused: validity byte buffer missing");
  // Both tables are partitioned first; then the side whose fullest bin is smaller becomes the build side. The reference
  // builds from the non-preserved side of a LEFT / RIGHT join and from the smaller table otherwise (hash_join.cpp:207-232),
  // which only changes the order of its output. Here it decides the cost: a key with millions of rows on the BUILD side
  // means thousands of build chunks and single probe rows that emit millions of matches from one thread (measured:
  // FULL_OUTER of config 3 with the Zipf side as build took 2.8 s instead of 7 ms). Either side can be preserved in
  // either role -- an unmatched preserved row is "probe row without partner" or "build row nobody matched".
  XWords lw, rw;
  CYL_TRY(x_build_words(lcols, nlcols, lkeys, nkeys, nl, &lw, s));
  CYL_TRY(x_build_words(rcols, nrcols, rkeys, nkeys, nr, &rw, s));
  const size_t budget = 104 * 1024;
  auto plan_cap = [&](int nwords, int *nheads_out, size_t *smem_out) {
    int cap = 16384 - 256;
    for (;; cap -= 256) {
      int nheads = 256;
      while (nheads < cap * 2) nheads *= 2;
      const size_t bytes = (size_t)cap * nwords * 8 + (size_t)nheads * 4 + (size_t)cap * 3 + 32;
      if (bytes <= budget || cap <= 256) {
        *nheads_out = nheads;
        *smem_out = bytes;
        return cap;
      }
    }
  };
  int lheads = 0, rheads = 0;
  size_t lsmem = 0, rsmem = 0;
  const int lcap = plan_cap(lw.nwords, &lheads, &lsmem), rcap = plan_cap(rw.nwords, &rheads, &rsmem);
  int64_t nbins = 1;
  while ((nbins * (int64_t)(lcap * 8 / 10) < nl || nbins * (int64_t)(rcap * 8 / 10) < nr) && nbins < (1ll << 22)) nbins *= 2;
  int bits = 0;
  while ((1ll << bits) < nbins) bits++;
  const int shift = 64 - bits;

  Scratch lrec, rrec, loff, roff, counters, pflag, maxbin;
  CYL_TRY(lrec.alloc((size_t)nl * lw.nwords * 8 + 16, s));
  CYL_TRY(rrec.alloc((size_t)nr * rw.nwords * 8 + 16, s));
  CYL_TRY(loff.alloc((size_t)(nbins + 1) * sizeof(int64_t), s));
  CYL_TRY(roff.alloc((size_t)(nbins + 1) * sizeof(int64_t), s));
  const int lpart = nkeys == 2 ? lw.hash_word : lw.key_word[0], rpart = nkeys == 2 ? rw.hash_word : rw.key_word[0];
  CYL_TRY(f8_partition(lw.cols, lw.nwords, lpart, nl, nbins, shift, lrec.as<uint64_t>(), loff.as<int64_t>(), s));
  CYL_TRY(f8_partition(rw.cols, rw.nwords, rpart, nr, nbins, shift, rrec.as<uint64_t>(), roff.as<int64_t>(), s));
  CYL_TRY(maxbin.alloc(64, s));
  CYL_CUDA(cudaMemsetAsync(maxbin.p, 0, 64, s));
  x_maxbin_kernel<<<x_grid(nbins), 256, 0, s>>>(loff.as<int64_t>(), roff.as<int64_t>(), nbins, maxbin.as<unsigned long long>());
  CYL_LAUNCH_CHECK();
  unsigned long long fullest[2] = {0, 0};
  CYL_CUDA(cudaMemcpyAsync(fullest, maxbin.p, sizeof(fullest), cudaMemcpyDeviceToHost, s));
  CYL_CUDA(cudaStreamSynchronize(s));
  bool build_right;
  if (fullest[0] > (unsigned long long)lcap || fullest[1] > (unsigned long long)rcap) {
    build_right = fullest[1] < fullest[0];  // some bin overflows shared memory on one side: build from the other
  } else {
    switch (join_type) {  // no skew to speak of: the reference's choice
      case CYL_JOIN_LEFT: build_right = true; break;
      case CYL_JOIN_RIGHT: build_right = false; break;
      default: build_right = !(nl <= nr); break;
    }
  }
  const cyl_column_t *bcols = build_right ? rcols : lcols, *pcols = build_right ? lcols : rcols;
  const int nbc = build_right ? nrcols : nlcols, npc = build_right ? nlcols : nrcols;
  const int32_t *bkeys = build_right ? rkeys : lkeys;
  const int64_t nb = build_right ? nr : nl, np = build_right ? nl : nr;
  XWords &bw = build_right ? rw : lw, &pw = build_right ? lw : rw;
  Scratch &brec = build_right ? rrec : lrec, &prec = build_right ? lrec : rrec;
  Scratch &boff = build_right ? roff : loff, &poff = build_right ? loff : roff;
  const int bpart = build_right ? rpart : lpart, ppart = build_right ? lpart : rpart;
  const int cap = build_right ? rcap : lcap, nheads = build_right ? rheads : lheads;
  const size_t smem_bytes = build_right ? rsmem : lsmem;

  F8XArgs a;
  memset(&a, 0, sizeof(a));
  a.brec = brec.as<uint64_t>();
  a.prec = prec.as<uint64_t>();
  a.b_off = boff.as<int64_t>();
  a.p_off = poff.as<int64_t>();
  a.nbins = nbins;
  a.nkw = nkeys;
  for (int k = 0; k < nkeys; k++) {
    a.b.kw[k] = bw.key_word[k];
    a.p.kw[k] = pw.key_word[k];
  }
  a.bucket_word_b = bpart;
  a.bucket_word_p = ppart;
  const int32_t ktype = bcols[bkeys[0]].type;
  a.nan_type = (nkeys == 1 && (ktype == CYL_FLOAT || ktype == CYL_DOUBLE)) ? ktype : 0;
  const bool left_kept = join_type == CYL_JOIN_LEFT || join_type == CYL_JOIN_FULL_OUTER;
  const bool right_kept = join_type == CYL_JOIN_RIGHT || join_type == CYL_JOIN_FULL_OUTER;
  a.keep_probe = build_right ? left_kept : right_kept;
  a.keep_build = build_right ? right_kept : left_kept;
  // output columns: all left columns then all right columns (build_final_table, join_utils.cpp:126-181)
  void *const *bout = build_right ? out_cols + nlcols : out_cols;
  void *const *pout = build_right ? out_cols : out_cols + nlcols;
  a.b.nout = nbc;
  for (int c = 0; c < nbc; c++) {
    a.b.word[c] = bw.col_word[c];
    a.b.narrow[c] = bcols[c].byte_width == 4;
    a.b.out[c] = static_cast<uint8_t *>(bout[c]);
  }
  a.p.nout = npc;
  for (int c = 0; c < npc; c++) {
    a.p.word[c] = pw.col_word[c];
    a.p.narrow[c] = pcols[c].byte_width == 4;
    a.p.out[c] = static_cast<uint8_t *>(pout[c]);
  }
  a.b.null_flag = build_right ? right_valid : left_valid;
  a.p.null_flag = build_right ? left_valid : right_valid;
  if (!(build_right ? right_can_be_null : left_can_be_null)) a.b.null_flag = nullptr;
  if (!(build_right ? left_can_be_null : right_can_be_null)) a.p.null_flag = nullptr;
  a.cap_rows = cap;
  a.nheads = nheads;
  a.capacity = capacity;
  CYL_TRY(counters.alloc(64, s));
  CYL_CUDA(cudaMemsetAsync(counters.p, 0, 64, s));
  a.out_cursor = counters.as<unsigned long long>();
  a.work_counter = reinterpret_cast<unsigned int *>(counters.as<char>() + 16);
  Scratch nsl, slice_off, bflag, tickets;
  CYL_TRY(nsl.alloc((size_t)(nbins + 1) * sizeof(uint32_t), s));
  CYL_TRY(slice_off.alloc((size_t)(nbins + 2) * sizeof(int64_t), s));
  x_slices_kernel<<<x_grid(nbins), 256, 0, s>>>(poff.as<int64_t>(), nbins, nsl.as<uint32_t>());
  CYL_LAUNCH_CHECK();
  CYL_TRY(exclusive_scan<uint32_t>(nsl.as<uint32_t>(), nbins, slice_off.as<int64_t>(), s));
  a.slice_off = slice_off.as<int64_t>();
  if (a.keep_build) {
    CYL_TRY(bflag.alloc((size_t)nb + 16, s));
    CYL_CUDA(cudaMemsetAsync(bflag.p, 0, (size_t)nb + 16, s));
    CYL_TRY(tickets.alloc((size_t)nbins * sizeof(unsigned int), s));
    CYL_CUDA(cudaMemsetAsync(tickets.p, 0, (size_t)nbins * sizeof(unsigned int), s));
    a.bflag = bflag.as<uint8_t>();
    a.bin_ticket = tickets.as<unsigned int>();
  }
  if (a.keep_probe) {
    CYL_TRY(pflag.alloc((size_t)np + 16, s));
    CYL_CUDA(cudaMemsetAsync(pflag.p, 0, (size_t)np + 16, s));
    a.pflag = pflag.as<uint8_t>();
  }
  const int64_t grid = (int64_t)num_sms() * 2;  // work items are handed out dynamically; their number is only known on the device
  {
    ProfScope ps("fused_join_probe_x", s);
    int rc;
    switch (bw.nwords) {
      case 1: rc = x_dispatch_pw<1>(pw.nwords, a, (int)grid, smem_bytes, s); break;
      case 2: rc = x_dispatch_pw<2>(pw.nwords, a, (int)grid, smem_bytes, s); break;
      case 3: rc = x_dispatch_pw<3>(pw.nwords, a, (int)grid, smem_bytes, s); break;
      default: rc = x_dispatch_pw<4>(pw.nwords, a, (int)grid, smem_bytes, s); break;
    }
    CYL_TRY(rc);
  }
  unsigned long long m = 0;
  CYL_CUDA(cudaMemcpyAsync(&m, a.out_cursor, sizeof(m), cudaMemcpyDeviceToHost, s));
  CYL_CUDA(cudaStreamSynchronize(s));
  *num_rows = (int64_t)m;
  return CYL_OK;
}

}  // extern "C"
