// join_fused8.cu -- the fused radix-partitioned join specialised for tables whose columns are all 8 bytes wide
// (int64 / uint64 / double / timestamp ...: the shape of BASELINE config 2, int64 key + double payload).
//
// Same algorithm as join_fused.cu, tuned for the two limits ncu showed on the generic version
// (profiles/r01_ncu_fused_v2.md): the L2 sector-operation rate in the scatter and the instruction-issue rate in the
// probe. Rows travel as packed RECORDS of W 8-byte words (W = number of columns, <= 4):
//   * the scatter writes one record per row with 128-bit stores (one L2 sector operation for a 16-byte row instead of
//     one per column) after reserving its slot with a 32-bit atomic cursor;
//   * a bin is copied into shared memory as one flat, fully coalesced block of words (the records of a bin are
//     contiguous), probed 4 rows per thread per tile so the block scan, the barrier pair and the single global
//     atomicAdd that reserves output rows are paid once per 2048 probe rows, and everything is compile-time sized
//     (no per-value width switches).
#include <memory>

#include "common.cuh"
#include "join_fused.cuh"
#include "scan.cuh"

namespace cylb {

constexpr int kF8Threads = 512;
constexpr int kF8ProbeRows = 4;
constexpr uint32_t kF8Nil = 0xFFFFu;
constexpr int kF8PartRows = 4;

struct F8PartArgs {
  const uint64_t *cols[kFused8MaxCols];
  int32_t ncols, key, key_is_f64;
  int64_t n;
  int32_t shift;
  unsigned int *hist;
  unsigned int *cursor;
  uint64_t *rec;
};

// floating-point identity = bit pattern (common.cuh); kept as a named step so every use of a key reads the same
__device__ __forceinline__ uint64_t f8_canon(uint64_t v, int key_is_f64) {
  (void)key_is_f64;
  return v;
}
__device__ __forceinline__ uint32_t f8_bin(uint64_t k, int shift) { return shift >= 64 ? 0u : (uint32_t)(mix64(k) >> shift); }
__device__ __forceinline__ uint32_t f8_bucket(uint64_t k) { return (uint32_t)((k * 0x9E3779B97F4A7C15ULL) >> 38); }
// v[idx] for a register array without a dynamic index (which would force the array into local memory)
template <int W>
__device__ __forceinline__ uint64_t f8_pick(const uint64_t (&v)[W], int idx) {
  uint64_t r = v[0];
#pragma unroll
  for (int w = 1; w < W; w++)
    if (w == idx) r = v[w];
  return r;
}

__global__ void __launch_bounds__(256) fused8_hist_kernel(const __grid_constant__ F8PartArgs a) {
  const uint64_t *__restrict__ key = a.cols[a.key];
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < a.n; i0 += stride * kF8PartRows) {
    uint64_t k[kF8PartRows];
#pragma unroll
    for (int u = 0; u < kF8PartRows; u++) {
      const int64_t i = i0 + u * stride;
      k[u] = i < a.n ? ld_stream_u64(key + i) : 0;
    }
#pragma unroll
    for (int u = 0; u < kF8PartRows; u++)
      if (i0 + u * stride < a.n) atomicAdd(&a.hist[f8_bin(f8_canon(k[u], a.key_is_f64), a.shift)], 1u);
  }
}

__global__ void narrow_cursor_kernel(const int64_t *__restrict__ off, int64_t n, unsigned int *__restrict__ cursor) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) cursor[i] = (unsigned int)off[i];
}

template <int W>
__device__ __forceinline__ void store_record(uint64_t *rec, int64_t pos, const uint64_t (&v)[W]) {
  uint64_t *p = rec + pos * W;
  if constexpr (W == 2) {
    *reinterpret_cast<ulonglong2 *>(p) = make_ulonglong2(v[0], v[1]);
  } else if constexpr (W == 4) {
    reinterpret_cast<ulonglong2 *>(p)[0] = make_ulonglong2(v[0], v[1]);
    reinterpret_cast<ulonglong2 *>(p)[1] = make_ulonglong2(v[2], v[3]);
  } else {
#pragma unroll
    for (int w = 0; w < W; w++) p[w] = v[w];
  }
}

template <int W>
__global__ void __launch_bounds__(256) fused8_scatter_kernel(const __grid_constant__ F8PartArgs a) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < a.n; i0 += stride * kF8PartRows) {
    uint64_t v[kF8PartRows][W];
#pragma unroll
    for (int u = 0; u < kF8PartRows; u++) {
      const int64_t i = i0 + u * stride;
#pragma unroll
      for (int w = 0; w < W; w++) v[u][w] = i < a.n ? ld_stream_u64(a.cols[w] + i) : 0;
    }
    unsigned int pos[kF8PartRows];
#pragma unroll
    for (int u = 0; u < kF8PartRows; u++) {
      pos[u] = 0;
      if (i0 + u * stride < a.n) pos[u] = atomicAdd(&a.cursor[f8_bin(f8_canon(f8_pick<W>(v[u], a.key), a.key_is_f64), a.shift)], 1u);
    }
#pragma unroll
    for (int u = 0; u < kF8PartRows; u++)
      if (i0 + u * stride < a.n) store_record<W>(a.rec, pos[u], v[u]);
  }
}

struct F8JoinArgs {
  const uint64_t *brec, *prec;  // partition-ordered records
  const int64_t *b_off, *p_off; // [K + 1]
  uint64_t *bout[kFused8MaxCols], *pout[kFused8MaxCols];
  int64_t nbins;
  int32_t bkey, pkey, key_is_f64;
  int32_t cap_rows, nheads;
  int64_t capacity;
  unsigned long long *out_cursor;
  unsigned int *work_counter;
};

template <int BW, int PW>
__global__ void __launch_bounds__(kF8Threads) fused8_join_kernel(const __grid_constant__ F8JoinArgs a) {
  extern __shared__ __align__(16) unsigned char smem[];
  __shared__ unsigned int s_bin;
  __shared__ unsigned int s_warp_tot[kF8ProbeRows][kF8Threads / 32];
  __shared__ unsigned long long s_base;
  __shared__ unsigned int s_total;
  uint64_t *brec = reinterpret_cast<uint64_t *>(smem);                                  // [cap_rows][BW]
  uint32_t *heads = reinterpret_cast<uint32_t *>(smem + (size_t)a.cap_rows * BW * 8);   // [nheads]
  uint16_t *next = reinterpret_cast<uint16_t *>(heads + a.nheads);                      // [cap_rows]
  const uint32_t hmask = (uint32_t)a.nheads - 1;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  while (true) {
    __syncthreads();
    if (tid == 0) s_bin = atomicAdd(a.work_counter, 1u);
    __syncthreads();
    const int64_t bin = s_bin;
    if (bin >= a.nbins) break;
    const int64_t b0 = a.b_off[bin], b1 = a.b_off[bin + 1];
    const int64_t p0 = a.p_off[bin], p1 = a.p_off[bin + 1];
    if (b1 == b0 || p1 == p0) continue;

    for (int64_t c0 = b0; c0 < b1; c0 += a.cap_rows) {  // more than one chunk only under key skew
      const int nb = (int)((b1 - c0) < a.cap_rows ? (b1 - c0) : a.cap_rows);
      __syncthreads();
      for (int h = tid; h < a.nheads; h += kF8Threads) heads[h] = kF8Nil;
      {  // the bin's records are contiguous: one flat coalesced copy into shared memory
        const uint64_t *src = a.brec + c0 * BW;
        const int nwords = nb * BW;
        if constexpr (BW % 2 == 0) {
          const ulonglong2 *s2 = reinterpret_cast<const ulonglong2 *>(src);
          ulonglong2 *d2 = reinterpret_cast<ulonglong2 *>(brec);
          for (int i = tid; i < nwords / 2; i += kF8Threads) d2[i] = s2[i];
        } else {
          for (int i = tid; i < nwords; i += kF8Threads) brec[i] = src[i];
        }
      }
      __syncthreads();
      for (int i = tid; i < nb; i += kF8Threads) {
        const uint64_t k = f8_canon(brec[i * BW + a.bkey], a.key_is_f64);
        next[i] = (uint16_t)atomicExch(&heads[f8_bucket(k) & hmask], (uint32_t)i);
      }
      __syncthreads();

      for (int64_t t0 = p0; t0 < p1; t0 += kF8Threads * kF8ProbeRows) {
        uint64_t pv[kF8ProbeRows][PW];
        uint32_t first[kF8ProbeRows];
        unsigned int cnt[kF8ProbeRows];
#pragma unroll
        for (int u = 0; u < kF8ProbeRows; u++) {
          const int64_t r = t0 + (int64_t)u * kF8Threads + tid;
          if (r < p1) {
            if constexpr (PW == 2) {
              const ulonglong2 t = reinterpret_cast<const ulonglong2 *>(a.prec)[r];
              pv[u][0] = t.x;
              pv[u][1] = t.y;
            } else {
#pragma unroll
              for (int w = 0; w < PW; w++) pv[u][w] = a.prec[r * PW + w];
            }
          } else {
#pragma unroll
            for (int w = 0; w < PW; w++) pv[u][w] = 0;
          }
        }
        unsigned int mine = 0;
#pragma unroll
        for (int u = 0; u < kF8ProbeRows; u++) {
          const int64_t r = t0 + (int64_t)u * kF8Threads + tid;
          first[u] = kF8Nil;
          cnt[u] = 0;
          if (r < p1) {
            const uint64_t k = f8_canon(f8_pick<PW>(pv[u], a.pkey), a.key_is_f64);
            // a NaN key joins nothing (operator== in the reference, arrow_comparator.cpp:240-243)
            uint32_t j = (a.key_is_f64 && key_bits_is_nan(CYL_DOUBLE, k)) ? kF8Nil : heads[f8_bucket(k) & hmask];
            for (; j != kF8Nil; j = next[j]) {
              if (f8_canon(brec[j * BW + a.bkey], a.key_is_f64) == k) {
                if (cnt[u] == 0) first[u] = j;
                cnt[u]++;
              }
            }
          }
          mine += cnt[u];
        }
        // Output slots are handed out u-major (every thread's probe row 0, then every thread's row 1, ...): inside one
        // store instruction the lanes then write consecutive output rows, i.e. whole sectors. (Thread-major slots cost
        // 3.6 L2 sector writes per sector of output: ncu, profiles/r01_ncu_final.md.)
        unsigned int inc[kF8ProbeRows];
#pragma unroll
        for (int u = 0; u < kF8ProbeRows; u++) {
          inc[u] = cnt[u];
#pragma unroll
          for (int d = 1; d < 32; d <<= 1) {
            const unsigned int n = __shfl_up_sync(0xffffffffu, inc[u], d);
            if (lane >= d) inc[u] += n;
          }
          if (lane == 31) s_warp_tot[u][warp] = inc[u];
        }
        __syncthreads();
        if (warp == 0) {  // one warp turns the per-warp totals into exclusive bases (u-major) for everybody
          unsigned int run = 0;
#pragma unroll
          for (int u = 0; u < kF8ProbeRows; u++) {
            const unsigned int t = lane < kF8Threads / 32 ? s_warp_tot[u][lane] : 0u;
            unsigned int sc = t;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
              const unsigned int n = __shfl_up_sync(0xffffffffu, sc, d);
              if (lane >= d) sc += n;
            }
            if (lane < kF8Threads / 32) s_warp_tot[u][lane] = run + sc - t;
            run += __shfl_sync(0xffffffffu, sc, 31);
          }
          if (lane == 0) {
            s_total = run;
            if (run) s_base = atomicAdd(a.out_cursor, (unsigned long long)run);  // one reservation per tile
          }
        }
        __syncthreads();
        const unsigned int total = s_total;
        unsigned int slot0[kF8ProbeRows];
#pragma unroll
        for (int u = 0; u < kF8ProbeRows; u++) slot0[u] = s_warp_tot[u][warp] + inc[u] - cnt[u];
        if (mine && (int64_t)s_base + total <= a.capacity) {
#pragma unroll
          for (int u = 0; u < kF8ProbeRows; u++) {
            if (cnt[u] == 0) continue;
            int64_t o = (int64_t)s_base + slot0[u];
            unsigned int left = cnt[u];
            const uint64_t k = f8_canon(f8_pick<PW>(pv[u], a.pkey), a.key_is_f64);
            for (uint32_t j = first[u]; left; j = next[j]) {
              if (j != first[u] && f8_canon(brec[j * BW + a.bkey], a.key_is_f64) != k) continue;
#pragma unroll
              for (int w = 0; w < BW; w++) a.bout[w][o] = brec[j * BW + w];
#pragma unroll
              for (int w = 0; w < PW; w++) a.pout[w][o] = pv[u][w];
              o++;
              left--;
            }
          }
        }
      }
    }
  }
}

static inline int f8_grid(int64_t n, int per_block) {
  int64_t b = div_up(n, per_block);
  const int64_t cap = (int64_t)num_sms() * 8;
  return (int)(b < 1 ? 1 : (b < cap ? b : cap));
}

template <int W>
static void launch_scatter(const F8PartArgs &a, cudaStream_t s) {
  fused8_scatter_kernel<W><<<f8_grid(a.n, 256 * kF8PartRows), 256, 0, s>>>(a);
}

// ---- block-tile partitioner: ranks rows with SHARED-memory atomics, no global atomics ----------------------------
// ncu on the global-cursor scatter above showed ~34 G returning atomics/s chip-wide (the L2 atomic units), i.e. 2.9 ms
// per 100M rows whatever the cursor width. Here one CTA owns a tile of consecutive rows: pass 1 histograms the tile
// into shared memory and stores the K counters as one coalesced row of counts[block][bin]; a column scan over blocks
// plus the scan of the bin totals turns them into absolute cursors; pass 2 loads its row of cursors into shared
// memory and reserves every row's slot with a shared-memory atomicAdd.
constexpr int kBlkThreads = 1024;
constexpr int64_t kBlkMaxBins = 32768;  // 128 KB of 32-bit cursors per CTA

struct F8BlockArgs {
  const uint64_t *cols[kFused8MaxCols];
  int32_t key, key_is_f64, shift;
  int64_t n, tile_rows, nbins, nblocks;
  int64_t bin_lo;                // histogram mode: only bins [bin_lo, bin_lo + nbins) are counted
  unsigned int *counts;          // [nblocks][nbins]
  const int64_t *bin_off;        // [nbins + 1]
  unsigned int *bin_total;       // [nbins]
  uint64_t *rec;
};

__global__ void __launch_bounds__(kBlkThreads) fused8_block_hist_kernel(const __grid_constant__ F8BlockArgs a) {
  extern __shared__ unsigned int s_cnt[];
  for (int64_t b = threadIdx.x; b < a.nbins; b += kBlkThreads) s_cnt[b] = 0;
  __syncthreads();
  const uint64_t *__restrict__ key = a.cols[a.key];
  const int64_t r0 = (int64_t)blockIdx.x * a.tile_rows;
  const int64_t r1 = r0 + a.tile_rows < a.n ? r0 + a.tile_rows : a.n;
  for (int64_t i0 = r0 + threadIdx.x; i0 < r1; i0 += (int64_t)kBlkThreads * kF8PartRows) {
    uint64_t k[kF8PartRows];
#pragma unroll
    for (int u = 0; u < kF8PartRows; u++) {
      const int64_t i = i0 + (int64_t)u * kBlkThreads;
      k[u] = i < r1 ? ld_stream_u64(key + i) : 0;
    }
#pragma unroll
    for (int u = 0; u < kF8PartRows; u++)
      if (i0 + (int64_t)u * kBlkThreads < r1) {
        const uint32_t b = f8_bin(f8_canon(k[u], a.key_is_f64), a.shift) - (uint32_t)a.bin_lo;
        if (b < (uint32_t)a.nbins) atomicAdd(&s_cnt[b], 1u);
      }
  }
  __syncthreads();
  if (a.counts) {  // block-private cursors: keep the tile's counters as one coalesced row
    unsigned int *row = a.counts + (int64_t)blockIdx.x * a.nbins;
    for (int64_t b = threadIdx.x; b < a.nbins; b += kBlkThreads) row[b] = s_cnt[b];
  } else {  // shared global cursors: only the bin totals are needed
    for (int64_t b = threadIdx.x; b < a.nbins; b += kBlkThreads)
      if (s_cnt[b]) atomicAdd(&a.bin_total[b], s_cnt[b]);
  }
}

// counts[block][bin] -> exclusive prefix over blocks (in place); bin_total[bin] = column sum. One thread per bin,
// consecutive threads read consecutive bins of a row: coalesced.
__global__ void __launch_bounds__(256) fused8_column_scan_kernel(const __grid_constant__ F8BlockArgs a) {
  const int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= a.nbins) return;
  unsigned int run = 0;
#pragma unroll 8
  for (int64_t blk = 0; blk < a.nblocks; blk++) {
    unsigned int *p = a.counts + blk * a.nbins + b;
    const unsigned int c = *p;
    *p = run;
    run += c;
  }
  a.bin_total[b] = run;
}

template <int W>
__global__ void __launch_bounds__(kBlkThreads) fused8_block_scatter_kernel(const __grid_constant__ F8BlockArgs a) {
  extern __shared__ unsigned int s_cur[];
  const unsigned int *row = a.counts + (int64_t)blockIdx.x * a.nbins;
  for (int64_t b = threadIdx.x; b < a.nbins; b += kBlkThreads) s_cur[b] = (unsigned int)a.bin_off[b] + row[b];
  __syncthreads();
  const int64_t r0 = (int64_t)blockIdx.x * a.tile_rows;
  const int64_t r1 = r0 + a.tile_rows < a.n ? r0 + a.tile_rows : a.n;
  for (int64_t i0 = r0 + threadIdx.x; i0 < r1; i0 += (int64_t)kBlkThreads * kF8PartRows) {
    uint64_t v[kF8PartRows][W];
#pragma unroll
    for (int u = 0; u < kF8PartRows; u++) {
      const int64_t i = i0 + (int64_t)u * kBlkThreads;
#pragma unroll
      for (int w = 0; w < W; w++) v[u][w] = i < r1 ? ld_stream_u64(a.cols[w] + i) : 0;
    }
    unsigned int pos[kF8PartRows];
#pragma unroll
    for (int u = 0; u < kF8PartRows; u++)  // all reservations first, then all stores: no store waits behind an atomic
      pos[u] = i0 + (int64_t)u * kBlkThreads < r1 ? atomicAdd(&s_cur[f8_bin(f8_canon(f8_pick<W>(v[u], a.key), a.key_is_f64), a.shift)], 1u) : 0u;
#pragma unroll
    for (int u = 0; u < kF8PartRows; u++)
      if (i0 + (int64_t)u * kBlkThreads < r1) store_record<W>(a.rec, pos[u], v[u]);
  }
}

template <int W>
static int launch_block_scatter(const F8BlockArgs &a, size_t smem, cudaStream_t s) {
  static PerDeviceOnce attr_once;
  if (auto once_guard = attr_once.first()) {
    CYL_CUDA(cudaFuncSetAttribute(fused8_block_scatter_kernel<W>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(kBlkMaxBins * 4)));
  }
  fused8_block_scatter_kernel<W><<<(unsigned)a.nblocks, kBlkThreads, smem, s>>>(a);
  CYL_LAUNCH_CHECK();
  return CYL_OK;
}

static int f8_partition_blocks(const cyl_column_t *cols, int ncols, int key, int64_t n, int64_t nbins, int shift, uint64_t *rec,
                               int64_t *bin_off, cudaStream_t s) {
  F8BlockArgs a;
  memset(&a, 0, sizeof(a));
  for (int c = 0; c < ncols; c++) a.cols[c] = static_cast<const uint64_t *>(cols[c].data);
  a.key = key;
  a.key_is_f64 = cols[key].type == CYL_DOUBLE;
  a.shift = shift;
  a.n = n;
  a.nbins = nbins;
  a.rec = rec;
  // tile: big enough that the K counters per block stay a small fraction of the data, small enough to fill the chip
  int64_t tile = n / ((int64_t)num_sms() * 4);
  if (tile < nbins * 2) tile = nbins * 2;
  if (tile < 8192) tile = 8192;
  tile = div_up(tile, (int64_t)kBlkThreads * kF8PartRows) * kBlkThreads * kF8PartRows;
  a.tile_rows = tile;
  a.nblocks = div_up(n, tile);
  Scratch counts, totals;
  CYL_TRY(counts.alloc((size_t)a.nblocks * nbins * sizeof(unsigned int), s));
  CYL_TRY(totals.alloc((size_t)nbins * sizeof(unsigned int), s));
  a.counts = counts.as<unsigned int>();
  a.bin_total = totals.as<unsigned int>();
  const size_t smem = (size_t)nbins * sizeof(unsigned int);
  static PerDeviceOnce attr_once;
  if (auto once_guard = attr_once.first()) {
    CYL_CUDA(cudaFuncSetAttribute(fused8_block_hist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(kBlkMaxBins * 4)));
  }
  {
    ProfScope ps("fused_join_hist", s);
    fused8_block_hist_kernel<<<(unsigned)a.nblocks, kBlkThreads, smem, s>>>(a);
    CYL_LAUNCH_CHECK();
  }
  {
    ProfScope ps("fused_join_offsets", s);
    fused8_column_scan_kernel<<<(unsigned)div_up(nbins, 256), 256, 0, s>>>(a);
    CYL_LAUNCH_CHECK();
    CYL_TRY(exclusive_scan<uint32_t>(totals.as<unsigned int>(), nbins, bin_off, s));
  }
  a.bin_off = bin_off;
  {
    ProfScope ps("fused_join_scatter", s);
    switch (ncols) {
      case 1: CYL_TRY(launch_block_scatter<1>(a, smem, s)); break;
      case 2: CYL_TRY(launch_block_scatter<2>(a, smem, s)); break;
      case 3: CYL_TRY(launch_block_scatter<3>(a, smem, s)); break;
      default: CYL_TRY(launch_block_scatter<4>(a, smem, s)); break;
    }
  }
  return CYL_OK;
}

// ---- two-pass staged radix partitioner -----------------------------------------------------------------------------
// What limits every single-pass variant above is one scattered 16-byte store per row (~35 G/s chip-wide). Splitting the
// bin number into a coarse digit (<= 8 bits) and a fine digit (<= 8 bits) lets a CTA sort a 4096-row tile by ONE digit
// inside shared memory and write each (tile, digit) run as consecutive records: 16-32 rows, 256-512 contiguous bytes, so
// the stores coalesce and the pass runs near copy speed. Two such passes move 4x the table's bytes and still beat one
// pass of scattered stores. Slots inside a run come from shared-memory atomics, ranges inside a bin from one global
// atomicAdd per (tile, digit) on cursors preset from the full K-bin histogram -- row order inside a bin is arbitrary,
// as it already was. All of a tile's loads are issued before the first use.
// Measured and rejected in round 2 (profiles/r02_join_experiments.md; cyl_tune_set switched between them on the same
// 100M x 100M input): ranking a tile's rows with ballots over the digit bits or with match.any instead of the
// shared-memory atomicAdd below (passes 0.70 -> 1.01-1.13 ms), linking the join's chains in rounds of plain stores
// instead of atomicExch (probe 2.49 -> 2.75 ms), and join CTAs of 128 / 256 threads over 2^17 / 2^16 smaller bins (probe
// 2.19 / 2.30 ms but histogram 0.66 / 0.42 ms: no gain for the step).
void groupby_fused8_tune(int key, int value);
void fused8_tune(int key, int value) { groupby_fused8_tune(key, value); }
constexpr int kRsThreads = 512;  // == kRsMaxDigits: one digit per thread in the per-tile scan
static int g_force_bins_log2 = 0;
void fused8_force_bins_log2(int bits) { g_force_bins_log2 = bits; }
static int64_t g_two_pass_min_rows = 1 << 21;  // below this the launch count matters more than the store pattern
constexpr int kRsMaxDigits = 512;
static_assert(kRsThreads == kRsMaxDigits, "the per-tile scan maps one digit to one thread");

struct F8RadixArgs {
  const uint64_t *cols[kFused8MaxCols];  // pass A input (columns)
  const uint64_t *src;                   // pass B input (records ordered by coarse digit)
  uint64_t *dst;
  int32_t key, key_is_f64, shift, lo_bits, ndigits, cols_aligned16;
  int64_t n;
  unsigned int *cursor;                  // pass A: [coarse]; pass B: [K]
  const int64_t *bin_off;                // [K + 1]
  const unsigned int *tile_prefix;       // pass B: [ncoarse + 1] tiles before each coarse partition
  int32_t ncoarse;
};

template <int W>
struct RsCfg {
  static constexpr int R = W <= 2 ? 8 : 4;   // rows per thread
  static constexpr int T = R * kRsThreads;   // rows per tile: 64 KB of records for W = 2 and W = 4
};

__global__ void fused8_tile_prefix_kernel(const int64_t *__restrict__ bin_off, int lo_bits, int ncoarse, int tile_rows, unsigned int *__restrict__ prefix) {
  __shared__ unsigned int s[kRsMaxDigits + 1];
  const int h = threadIdx.x;
  if (h < ncoarse) {
    const int64_t rows = bin_off[(int64_t)(h + 1) << lo_bits] - bin_off[(int64_t)h << lo_bits];
    s[h + 1] = (unsigned int)((rows + tile_rows - 1) / tile_rows);
  }
  if (h == 0) s[0] = 0;
  __syncthreads();
  if (h == 0)
    for (int i = 1; i <= ncoarse; i++) s[i] += s[i - 1];
  __syncthreads();
  if (h <= ncoarse) prefix[h] = s[h];
}

__global__ void fused8_coarse_cursor_kernel(const int64_t *__restrict__ bin_off, int lo_bits, int ncoarse, unsigned int *__restrict__ cursor) {
  const int h = blockIdx.x * blockDim.x + threadIdx.x;
  if (h < ncoarse) cursor[h] = (unsigned int)bin_off[(int64_t)h << lo_bits];
}

template <int W, bool PASS_B>
__global__ void __launch_bounds__(kRsThreads, 2) fused8_radix_pass_kernel(const __grid_constant__ F8RadixArgs a) {
  constexpr int R = RsCfg<W>::R;
  constexpr int T = RsCfg<W>::T;
  extern __shared__ __align__(16) unsigned char rs_smem[];
  uint64_t *stage = reinterpret_cast<uint64_t *>(rs_smem);                     // [T][W]
  unsigned int *cnt = reinterpret_cast<unsigned int *>(stage + (size_t)T * W);  // [512] rows per digit, then run starts
  unsigned int *delta = cnt + kRsMaxDigits;                                     // [512] destination - staging slot
  unsigned int *warp_tot = delta + kRsMaxDigits;                                // [16]
  unsigned int *prefix = warp_tot + kRsThreads / 32;                                          // [513] (pass B)
  const int tid = threadIdx.x;
  const unsigned lane = tid & 31, warp = tid >> 5;
  const uint32_t lo_mask = (1u << a.lo_bits) - 1;

  int64_t ntiles;
  if (PASS_B) {
    for (int i = tid; i <= a.ncoarse; i += kRsThreads) prefix[i] = a.tile_prefix[i];
    __syncthreads();
    ntiles = prefix[a.ncoarse];
  } else {
    ntiles = (a.n + T - 1) / T;
  }
  // rows [r0, r1) of tile g and (pass B) the first fine cursor of its coarse partition
  auto tile_range = [&](int64_t g, int64_t &r0, int64_t &r1, uint32_t &cursor_base) {
    cursor_base = 0;
    if (PASS_B) {
      int lo = 0, hi = a.ncoarse;  // last h with prefix[h] <= g
      while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (prefix[mid] <= (unsigned int)g) lo = mid; else hi = mid;
      }
      const int64_t p0 = a.bin_off[(int64_t)lo << a.lo_bits], p1 = a.bin_off[(int64_t)(lo + 1) << a.lo_bits];
      r0 = p0 + (g - prefix[lo]) * T;
      r1 = r0 + T < p1 ? r0 + T : p1;
      cursor_base = (uint32_t)lo << a.lo_bits;
    } else {
      r0 = g * T;
      r1 = r0 + T < a.n ? r0 + T : a.n;
    }
  };
  // Thread `tid` owns rows r0 + (u/2)*2*256 + 2*tid + (u&1): pairs of neighbouring rows, so that column input can be read
  // with 128-bit loads when the tile start is 16-byte aligned (r0 is even and the column base is).
  uint64_t v[R][W];
  auto row_of = [&](int64_t r0, int u) { return r0 + (int64_t)(u >> 1) * (2 * kRsThreads) + 2 * tid + (u & 1); };
  auto load_tile = [&](int64_t r0, int64_t r1) {
    if (PASS_B) {
#pragma unroll
      for (int u = 0; u < R; u++) {
        const int64_t i = row_of(r0, u);
        if constexpr (W == 2) {
          ulonglong2 t = make_ulonglong2(0, 0);
          if (i < r1) t = __ldcs(reinterpret_cast<const ulonglong2 *>(a.src + i * 2));
          v[u][0] = t.x;
          v[u][1] = t.y;
        } else if constexpr (W == 4) {
          ulonglong2 t0 = make_ulonglong2(0, 0), t1 = t0;
          if (i < r1) {
            t0 = __ldcs(reinterpret_cast<const ulonglong2 *>(a.src + i * 4));
            t1 = __ldcs(reinterpret_cast<const ulonglong2 *>(a.src + i * 4 + 2));
          }
          v[u][0] = t0.x; v[u][1] = t0.y; v[u][2] = t1.x; v[u][3] = t1.y;
        } else {
#pragma unroll
          for (int w = 0; w < W; w++) v[u][w] = i < r1 ? ld_stream_u64(a.src + i * W + w) : 0;
        }
      }
    } else if (a.cols_aligned16 && !(r0 & 1)) {
#pragma unroll
      for (int u = 0; u < R; u += 2) {
        const int64_t i = row_of(r0, u);
#pragma unroll
        for (int w = 0; w < W; w++) {
          ulonglong2 t = make_ulonglong2(0, 0);
          if (i + 1 < r1) t = __ldcs(reinterpret_cast<const ulonglong2 *>(a.cols[w] + i));
          else if (i < r1) t.x = ld_stream_u64(a.cols[w] + i);
          v[u][w] = t.x;
          v[u + 1][w] = t.y;
        }
      }
    } else {
#pragma unroll
      for (int u = 0; u < R; u++) {
        const int64_t i = row_of(r0, u);
#pragma unroll
        for (int w = 0; w < W; w++) v[u][w] = i < r1 ? ld_stream_u64(a.cols[w] + i) : 0;
      }
    }
  };

  int64_t r0 = 0, r1 = 0;
  uint32_t cursor_base = 0;
  if ((int64_t)blockIdx.x < ntiles) {
    tile_range(blockIdx.x, r0, r1, cursor_base);
    load_tile(r0, r1);
  }
  for (int64_t g = blockIdx.x; g < ntiles; g += gridDim.x) {
    cnt[tid] = 0;
    __syncthreads();  // counters are zero
    uint32_t dr[R];   // digit << 16 | rank inside the (tile, digit) run
#pragma unroll
    for (int u = 0; u < R; u++) {
      const int64_t i = row_of(r0, u);
      const uint32_t bin = f8_bin(f8_canon(f8_pick<W>(v[u], a.key), a.key_is_f64), a.shift);
      const uint32_t d = PASS_B ? (bin & lo_mask) : (bin >> a.lo_bits);
      dr[u] = i < r1 ? ((d << 16) | atomicAdd(&cnt[d], 1u)) : 0xFFFFFFFFu;
    }
    __syncthreads();
    {  // exclusive scan of the 256 counters; digit `tid` reserves its run's destination range
      const unsigned int c = cnt[tid];
      unsigned int inc = c;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const unsigned int t = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= (unsigned)d) inc += t;
      }
      if (lane == 31) warp_tot[warp] = inc;
      unsigned int g0 = 0;
      if (c) g0 = atomicAdd(&a.cursor[cursor_base + tid], c);
      __syncthreads();
      unsigned int before = 0;
#pragma unroll
      for (int w = 0; w < kRsThreads / 32; w++) before += w < (int)warp ? warp_tot[w] : 0u;
      const unsigned int start = before + inc - c;
      cnt[tid] = start;
      delta[tid] = g0 - start;
    }
    __syncthreads();
#pragma unroll
    for (int u = 0; u < R; u++) {
      if (dr[u] != 0xFFFFFFFFu) {
        const uint32_t d = dr[u] >> 16, slot = cnt[d] + (dr[u] & 0xFFFFu);
        if constexpr (W == 2) {
          *reinterpret_cast<ulonglong2 *>(stage + (size_t)slot * 2) = make_ulonglong2(v[u][0], v[u][1]);
        } else {
#pragma unroll
          for (int w = 0; w < W; w++) stage[(size_t)slot * W + w] = v[u][w];
        }
      }
    }
    const int nvalid = (int)(r1 - r0);
    if (g + gridDim.x < ntiles) {  // the registers are free again: the next tile's loads fly while this one is written out
      tile_range(g + gridDim.x, r0, r1, cursor_base);
      load_tile(r0, r1);
    }
    __syncthreads();
    for (int j = tid; j < nvalid; j += kRsThreads) {  // consecutive threads -> consecutive records of a run
      uint64_t rec[W];
      if constexpr (W == 2) {
        const ulonglong2 t = *reinterpret_cast<const ulonglong2 *>(stage + (size_t)j * 2);
        rec[0] = t.x;
        rec[1] = t.y;
      } else {
#pragma unroll
        for (int w = 0; w < W; w++) rec[w] = stage[(size_t)j * W + w];
      }
      // the run a slot belongs to = the digit of its key, recomputed: a few ALU instructions instead of a per-slot
      // digit array in shared memory (measured the same; shared memory is the busy resource of this kernel)
      const uint32_t bin = f8_bin(f8_canon(f8_pick<W>(rec, a.key), a.key_is_f64), a.shift);
      const uint32_t d = PASS_B ? (bin & lo_mask) : (bin >> a.lo_bits);
      store_record<W>(a.dst, delta[d] + (uint32_t)j, rec);
    }
    __syncthreads();
  }
}

template <int W>
static size_t rs_smem_bytes() {
  return (size_t)RsCfg<W>::T * W * 8 + (2 * kRsMaxDigits + kRsThreads / 32 + kRsMaxDigits + 1) * sizeof(unsigned int) + 16;
}

template <int W, bool PASS_B>
static int launch_radix_pass(const F8RadixArgs &a, int64_t tiles_bound, cudaStream_t s) {
  static PerDeviceOnce attr_once;
  if (auto once_guard = attr_once.first()) {
    CYL_CUDA(cudaFuncSetAttribute(fused8_radix_pass_kernel<W, PASS_B>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rs_smem_bytes<W>()));
  }
  int64_t grid = (int64_t)num_sms() * 2;
  if (grid > tiles_bound) grid = tiles_bound < 1 ? 1 : tiles_bound;
  fused8_radix_pass_kernel<W, PASS_B><<<(unsigned)grid, kRsThreads, rs_smem_bytes<W>(), s>>>(a);
  CYL_LAUNCH_CHECK();
  return CYL_OK;
}

template <int W>
static int radix_two_pass(F8RadixArgs a, uint64_t *tmp, uint64_t *rec, unsigned int *coarse_cursor, unsigned int *fine_cursor, cudaStream_t s) {
  constexpr int T = RsCfg<W>::T;
  {
    ProfScope ps("fused_join_scatter_coarse", s);
    a.dst = tmp;
    a.cursor = coarse_cursor;
    a.ndigits = a.ncoarse;
    CYL_TRY((launch_radix_pass<W, false>(a, div_up(a.n, (int64_t)T), s)));
  }
  {
    ProfScope ps("fused_join_scatter_fine", s);
    a.src = tmp;
    a.dst = rec;
    a.cursor = fine_cursor;
    a.ndigits = 1 << a.lo_bits;
    CYL_TRY((launch_radix_pass<W, true>(a, div_up(a.n, (int64_t)T) + a.ncoarse, s)));
  }
  return CYL_OK;
}

// a.hist[bin] += rows of every bin (a.hist zeroed by the caller)
static int f8_bin_histogram(const F8PartArgs &a, int64_t nbins, int64_t tile_guess, cudaStream_t s) {
  if (nbins <= kBlkMaxBins) {  // histogram in shared memory, totals merged with one reduction per (tile, bin)
    F8BlockArgs h;
    memset(&h, 0, sizeof(h));
    for (int c = 0; c < a.ncols; c++) h.cols[c] = a.cols[c];
    h.key = a.key;
    h.key_is_f64 = a.key_is_f64;
    h.shift = a.shift;
    h.n = a.n;
    h.nbins = nbins;
    int64_t tile = div_up(tile_guess < 8192 ? 8192 : tile_guess, (int64_t)kBlkThreads * kF8PartRows) * kBlkThreads * kF8PartRows;
    h.tile_rows = tile;
    h.nblocks = div_up(a.n, tile);
    h.bin_total = a.hist;
    static PerDeviceOnce attr_once;
    if (auto once_guard = attr_once.first()) {
      CYL_CUDA(cudaFuncSetAttribute(fused8_block_hist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(kBlkMaxBins * 4)));
    }
    ProfScope ps("fused_join_hist", s);
    fused8_block_hist_kernel<<<(unsigned)h.nblocks, kBlkThreads, (size_t)nbins * sizeof(unsigned int), s>>>(h);
    CYL_LAUNCH_CHECK();
  } else if (nbins == 2 * kBlkMaxBins) {
    // 65536 bins do not fit one CTA's shared memory: two launches, each counting one half of the bin range and reading
    // all keys (2 x 0.2 ms per 100 M rows; reductions straight into HBM cost 0.74 ms)
    F8BlockArgs h;
    memset(&h, 0, sizeof(h));
    for (int c = 0; c < a.ncols; c++) h.cols[c] = a.cols[c];
    h.key = a.key;
    h.key_is_f64 = a.key_is_f64;
    h.shift = a.shift;
    h.n = a.n;
    h.nbins = kBlkMaxBins;
    int64_t tile = div_up(tile_guess < 8192 ? 8192 : tile_guess, (int64_t)kBlkThreads * kF8PartRows) * kBlkThreads * kF8PartRows;
    h.tile_rows = tile;
    h.nblocks = div_up(a.n, tile);
    static PerDeviceOnce attr_once2;
    if (auto once_guard2 = attr_once2.first()) {
      CYL_CUDA(cudaFuncSetAttribute(fused8_block_hist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(kBlkMaxBins * 4)));
    }
    ProfScope ps("fused_join_hist", s);
    for (int half = 0; half < 2; half++) {
      h.bin_lo = half * kBlkMaxBins;
      h.bin_total = a.hist + half * kBlkMaxBins;
      fused8_block_hist_kernel<<<(unsigned)h.nblocks, kBlkThreads, (size_t)kBlkMaxBins * sizeof(unsigned int), s>>>(h);
      CYL_LAUNCH_CHECK();
    }
  } else {
    ProfScope ps("fused_join_hist", s);
    fused8_hist_kernel<<<f8_grid(a.n, 256 * kF8PartRows), 256, 0, s>>>(a);
    CYL_LAUNCH_CHECK();
  }
  return CYL_OK;
}

static int f8_partition_two_pass(const cyl_column_t *cols, int ncols, int key, int64_t n, int64_t nbins, int shift, uint64_t *rec,
                                 int64_t *bin_off, cudaStream_t s) {
  const int kb = 64 - shift;
  const int hi_bits = (kb + 1) / 2, lo_bits = kb - hi_bits;  // each <= 9 = log2(kRsMaxDigits)
  F8PartArgs h;
  memset(&h, 0, sizeof(h));
  for (int c = 0; c < ncols; c++) h.cols[c] = static_cast<const uint64_t *>(cols[c].data);
  h.ncols = ncols;
  h.key = key;
  h.key_is_f64 = cols[key].type == CYL_DOUBLE;
  h.n = n;
  h.shift = shift;
  Scratch hist, fine, coarse, prefix, tmp;
  CYL_TRY(hist.alloc((size_t)nbins * sizeof(unsigned int), s));
  CYL_TRY(fine.alloc((size_t)nbins * sizeof(unsigned int), s));
  CYL_TRY(coarse.alloc((size_t)kRsMaxDigits * sizeof(unsigned int), s));
  CYL_TRY(prefix.alloc((size_t)(kRsMaxDigits + 1) * sizeof(unsigned int), s));
  CYL_TRY(tmp.alloc((size_t)n * ncols * 8 + 16, s));
  CYL_CUDA(cudaMemsetAsync(hist.p, 0, (size_t)nbins * sizeof(unsigned int), s));
  h.hist = hist.as<unsigned int>();
  CYL_TRY(f8_bin_histogram(h, nbins, n / ((int64_t)num_sms() * 4) + 1, s));
  F8RadixArgs a;
  memset(&a, 0, sizeof(a));
  for (int c = 0; c < ncols; c++) a.cols[c] = h.cols[c];
  a.cols_aligned16 = 1;
  for (int c = 0; c < ncols; c++) a.cols_aligned16 &= (reinterpret_cast<uintptr_t>(a.cols[c]) & 15) == 0;
  a.key = key;
  a.key_is_f64 = h.key_is_f64;
  a.shift = shift;
  a.lo_bits = lo_bits;
  a.n = n;
  a.bin_off = bin_off;
  a.ncoarse = 1 << hi_bits;
  a.tile_prefix = prefix.as<unsigned int>();
  const int tile_rows = ncols <= 2 ? RsCfg<2>::T : RsCfg<4>::T;
  {
    ProfScope ps("fused_join_offsets", s);
    CYL_TRY(exclusive_scan<uint32_t>(hist.as<unsigned int>(), nbins, bin_off, s));
    narrow_cursor_kernel<<<f8_grid(nbins, 256), 256, 0, s>>>(bin_off, nbins, fine.as<unsigned int>());
    CYL_LAUNCH_CHECK();
    fused8_coarse_cursor_kernel<<<1, kRsMaxDigits, 0, s>>>(bin_off, lo_bits, a.ncoarse, coarse.as<unsigned int>());
    CYL_LAUNCH_CHECK();
    fused8_tile_prefix_kernel<<<1, kRsMaxDigits + 32, 0, s>>>(bin_off, lo_bits, a.ncoarse, tile_rows, prefix.as<unsigned int>());
    CYL_LAUNCH_CHECK();
  }
  switch (ncols) {
    case 1: return radix_two_pass<1>(a, tmp.as<uint64_t>(), rec, coarse.as<unsigned int>(), fine.as<unsigned int>(), s);
    case 2: return radix_two_pass<2>(a, tmp.as<uint64_t>(), rec, coarse.as<unsigned int>(), fine.as<unsigned int>(), s);
    case 3: return radix_two_pass<3>(a, tmp.as<uint64_t>(), rec, coarse.as<unsigned int>(), fine.as<unsigned int>(), s);
    default: return radix_two_pass<4>(a, tmp.as<uint64_t>(), rec, coarse.as<unsigned int>(), fine.as<unsigned int>(), s);
  }
}

int f8_partition(const cyl_column_t *cols, int ncols, int key, int64_t n, int64_t nbins, int shift, uint64_t *rec, int64_t *bin_off,
                 cudaStream_t s) {
  // Measured on B200 at 100M rows x 32768 bins (profiles/r01_fused_join_notes.md): block-private cursors make the
  // scatter SLOWER (4.4 ms vs 2.9 ms) because 148 CTAs x 32768 bins leave ~5M partially written lines in flight, far
  // more than L2 can hold until they fill; shared global cursors keep only K write frontiers alive. A two-level split
  // (2048 coarse bins with block-private cursors, then one CTA per coarse bin) was also measured and rejected: its first
  // level alone took 2.8 ms (shared-memory atomic and store-queue stalls), 4.3 ms per table in total against 3.1 ms.
  // Forcing block-private cursors at 1024-4096 bins gave 2.7-3.5 ms as well: what costs is the ONE scattered 16-byte
  // store per row (~37 G scattered stores/s chip-wide whatever the table size, so not TLB reach either), not the atomic.
  // So: block-private ranges only while a (tile, bin) run is long enough to fill lines, else shared cursors -- but
  // always the shared-memory histogram, which is 4x faster than global reductions.
  if (n == 0) {  // nothing to move: every bin is empty
    CYL_CUDA(cudaMemsetAsync(bin_off, 0, (size_t)(nbins + 1) * sizeof(int64_t), s));
    return CYL_OK;
  }
  const int64_t tile_guess = n / ((int64_t)num_sms() * 4) + 1;
  {  // large tables: two staged passes with coalesced stores instead of one pass of scattered stores
    const int kb = 64 - shift;
    if (kb >= 9 && kb <= 18 && n >= (int64_t)g_two_pass_min_rows && n < 0xFFFFFFFFll) return f8_partition_two_pass(cols, ncols, key, n, nbins, shift, rec, bin_off, s);
  }
  if (nbins <= kBlkMaxBins && tile_guess / nbins >= 32) return f8_partition_blocks(cols, ncols, key, n, nbins, shift, rec, bin_off, s);
  F8PartArgs a;
  memset(&a, 0, sizeof(a));
  for (int c = 0; c < ncols; c++) a.cols[c] = static_cast<const uint64_t *>(cols[c].data);
  a.ncols = ncols;
  a.key = key;
  a.key_is_f64 = cols[key].type == CYL_DOUBLE;
  a.n = n;
  a.shift = shift;
  a.rec = rec;
  Scratch hist, cursor;
  CYL_TRY(hist.alloc((size_t)nbins * sizeof(unsigned int), s));
  CYL_TRY(cursor.alloc((size_t)nbins * sizeof(unsigned int), s));
  CYL_CUDA(cudaMemsetAsync(hist.p, 0, (size_t)nbins * sizeof(unsigned int), s));
  a.hist = hist.as<unsigned int>();
  a.cursor = cursor.as<unsigned int>();
  CYL_TRY(f8_bin_histogram(a, nbins, tile_guess, s));
  CYL_TRY(exclusive_scan<uint32_t>(hist.as<unsigned int>(), nbins, bin_off, s));
  narrow_cursor_kernel<<<f8_grid(nbins, 256), 256, 0, s>>>(bin_off, nbins, a.cursor);
  CYL_LAUNCH_CHECK();
  {
    ProfScope ps("fused_join_scatter", s);
    switch (ncols) {
      case 1: launch_scatter<1>(a, s); break;
      case 2: launch_scatter<2>(a, s); break;
      case 3: launch_scatter<3>(a, s); break;
      default: launch_scatter<4>(a, s); break;
    }
    CYL_LAUNCH_CHECK();
  }
  return CYL_OK;
}

template <int BW, int PW>
static int launch_join(const F8JoinArgs &a, int grid, size_t smem_bytes, cudaStream_t s) {
  static PerDeviceOnce attr_once;
  if (auto once_guard = attr_once.first()) {
    CYL_CUDA(cudaFuncSetAttribute(fused8_join_kernel<BW, PW>, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024));
  }
  fused8_join_kernel<BW, PW><<<grid, kF8Threads, smem_bytes, s>>>(a);
  CYL_LAUNCH_CHECK();
  return CYL_OK;
}

template <int BW>
static int dispatch_pw(int pw, const F8JoinArgs &a, int grid, size_t smem_bytes, cudaStream_t s) {
  switch (pw) {
    case 1: return launch_join<BW, 1>(a, grid, smem_bytes, s);
    case 2: return launch_join<BW, 2>(a, grid, smem_bytes, s);
    case 3: return launch_join<BW, 3>(a, grid, smem_bytes, s);
    default: return launch_join<BW, 4>(a, grid, smem_bytes, s);
  }
}

bool fused8_eligible(const cyl_column_t *lcols, int nlcols, const cyl_column_t *rcols, int nrcols) {
  if (nlcols < 1 || nrcols < 1 || nlcols > kFused8MaxCols || nrcols > kFused8MaxCols) return false;
  for (int c = 0; c < nlcols; c++)
    if (lcols[c].byte_width != 8) return false;
  for (int c = 0; c < nrcols; c++)
    if (rcols[c].byte_width != 8) return false;
  return true;
}

// Shared-memory plan of the join kernel for a build side of `nbc` words: rows per bin, chain heads, bytes; and the bin count
// for `nb` build rows.
static void f8_plan(int nbc, int64_t nb, int *cap_out, int *nheads_out, size_t *smem_out, int64_t *nbins_out, int *shift_out) {
  // shared memory per CTA: records + 16-bit next links + 32-bit heads (2 CTAs per SM)
  const size_t budget = 104 * 1024;
  int cap = 16384, nheads = 0;  // next links are 16-bit and 0xFFFF ends a chain
  size_t smem_bytes = 0;
  for (;; cap -= 256) {  // largest multiple of 256 rows whose records + links + heads fit the budget
    nheads = 256;
    while (nheads < cap * 2) nheads *= 2;
    smem_bytes = (size_t)cap * nbc * 8 + (size_t)nheads * 4 + (size_t)cap * 2 + 16;
    if (smem_bytes <= budget || cap <= 256) break;
  }
  int64_t nbins = 1;
  while (nbins * (int64_t)(cap * 8 / 10) < nb && nbins < (1ll << 22)) nbins *= 2;
  if (g_force_bins_log2 > 0) nbins = 1ll << g_force_bins_log2;  // test hook: more (smaller) bins than the size asks for
  int bits = 0;
  while ((1ll << bits) < nbins) bits++;
  *cap_out = cap;
  *nheads_out = nheads;
  *smem_out = smem_bytes;
  *nbins_out = nbins;
  *shift_out = 64 - bits;
}

// One table radix-partitioned into the join's bins: the unit the distributed join prepares for the first table while the
// second one is still being shuffled (host/shuffle.cpp).
struct F8Partitioned {
  Scratch rec, off;
  int64_t n = 0, nbins = 0;
  int ncols = 0, key = 0, shift = 0, key_is_f64 = 0;
};

int fused8_partition_side(const cyl_column_t *cols, int ncols, int key, int64_t n, int64_t nbins, int shift, F8Partitioned **out, cudaStream_t s) {
  std::unique_ptr<F8Partitioned> p(new F8Partitioned());
  p->n = n;
  p->nbins = nbins;
  p->ncols = ncols;
  p->key = key;
  p->shift = shift;
  p->key_is_f64 = cols[key].type == CYL_DOUBLE;
  CYL_TRY(p->rec.alloc((size_t)n * ncols * 8 + 16, s));
  CYL_TRY(p->off.alloc((size_t)(nbins + 1) * sizeof(int64_t), s));
  CYL_TRY(f8_partition(cols, ncols, key, n, nbins, shift, p->rec.as<uint64_t>(), p->off.as<int64_t>(), s));
  *out = p.release();
  return CYL_OK;
}

void fused8_side_free(F8Partitioned *p) { delete p; }
int64_t fused8_side_bins(const F8Partitioned *p) { return p->nbins; }
int fused8_side_shift(const F8Partitioned *p) { return p->shift; }
void fused8_bins_for(int build_words, int64_t build_rows, int64_t *nbins, int *shift) {
  int cap, nheads;
  size_t smem;
  f8_plan(build_words, build_rows, &cap, &nheads, &smem, nbins, shift);
}

int fused8_probe(const F8Partitioned *b, const F8Partitioned *p, void *const *bout, void *const *pout, int64_t capacity, int64_t *num_rows,
                 cudaStream_t s) {
  if (b->nbins != p->nbins || b->shift != p->shift) return fail(CYL_Invalid, "fused join: the two sides were partitioned into different bins");
  int cap, nheads, shift;
  size_t smem_bytes;
  int64_t nbins_unused;
  f8_plan(b->ncols, b->n, &cap, &nheads, &smem_bytes, &nbins_unused, &shift);  // cap depends on the record width only
  const int nbc = b->ncols, npc = p->ncols;
  Scratch counters;
  F8JoinArgs a;
  memset(&a, 0, sizeof(a));
  a.brec = b->rec.as<uint64_t>();
  a.prec = p->rec.as<uint64_t>();
  a.b_off = b->off.as<int64_t>();
  a.p_off = p->off.as<int64_t>();
  for (int c = 0; c < nbc; c++) a.bout[c] = static_cast<uint64_t *>(bout[c]);
  for (int c = 0; c < npc; c++) a.pout[c] = static_cast<uint64_t *>(pout[c]);
  a.nbins = b->nbins;
  a.bkey = b->key;
  a.pkey = p->key;
  a.key_is_f64 = b->key_is_f64;
  a.cap_rows = cap;
  a.nheads = nheads;
  a.capacity = capacity;
  CYL_TRY(counters.alloc(64, s));
  CYL_CUDA(cudaMemsetAsync(counters.p, 0, 64, s));
  a.out_cursor = counters.as<unsigned long long>();
  a.work_counter = reinterpret_cast<unsigned int *>(counters.as<char>() + 16);
  int64_t grid = (int64_t)num_sms() * 2;
  if (grid > a.nbins) grid = a.nbins;
  {
    ProfScope ps("fused_join_probe", s);
    int rc;
    switch (nbc) {
      case 1: rc = dispatch_pw<1>(npc, a, (int)grid, smem_bytes, s); break;
      case 2: rc = dispatch_pw<2>(npc, a, (int)grid, smem_bytes, s); break;
      case 3: rc = dispatch_pw<3>(npc, a, (int)grid, smem_bytes, s); break;
      default: rc = dispatch_pw<4>(npc, a, (int)grid, smem_bytes, s); break;
    }
    CYL_TRY(rc);
  }
  unsigned long long m = 0;
  CYL_CUDA(cudaMemcpyAsync(&m, a.out_cursor, sizeof(m), cudaMemcpyDeviceToHost, s));
  CYL_CUDA(cudaStreamSynchronize(s));
  *num_rows = (int64_t)m;
  return CYL_OK;
}

int fused8_join(const cyl_column_t *bcols, int nbc, int bkey, int64_t nb, const cyl_column_t *pcols, int npc, int pkey, int64_t np,
                void *const *bout, void *const *pout, int64_t capacity, int64_t *num_rows, cudaStream_t s) {
  int64_t nbins;
  int shift;
  fused8_bins_for(nbc, nb, &nbins, &shift);
  F8Partitioned *b = nullptr, *p = nullptr;
  CYL_TRY(fused8_partition_side(bcols, nbc, bkey, nb, nbins, shift, &b, s));
  std::unique_ptr<F8Partitioned> bguard(b);
  CYL_TRY(fused8_partition_side(pcols, npc, pkey, np, nbins, shift, &p, s));
  std::unique_ptr<F8Partitioned> pguard(p);
  return fused8_probe(b, p, bout, pout, capacity, num_rows, s);
}

}  // namespace cylb
