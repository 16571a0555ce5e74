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

__device__ __forceinline__ uint64_t f8_canon(uint64_t v, int key_is_f64) { return (key_is_f64 && v == 0x8000000000000000ULL) ? 0ull : v; }
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
            for (uint32_t j = heads[f8_bucket(k) & hmask]; j != kF8Nil; j = next[j]) {
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
      if (i0 + (int64_t)u * kBlkThreads < r1) atomicAdd(&s_cnt[f8_bin(f8_canon(k[u], a.key_is_f64), a.shift)], 1u);
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
  static bool attr_set = false;
  if (!attr_set) {
    CYL_CUDA(cudaFuncSetAttribute(fused8_block_scatter_kernel<W>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(kBlkMaxBins * 4)));
    attr_set = true;
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
  static bool attr_set = false;
  if (!attr_set) {
    CYL_CUDA(cudaFuncSetAttribute(fused8_block_hist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(kBlkMaxBins * 4)));
    attr_set = true;
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

static int f8_partition(const cyl_column_t *cols, int ncols, int key, int64_t n, int64_t nbins, int shift, uint64_t *rec, int64_t *bin_off,
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
  const int64_t tile_guess = n / ((int64_t)num_sms() * 4) + 1;
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
  if (nbins <= kBlkMaxBins) {  // histogram in shared memory, totals merged with one reduction per (tile, bin)
    F8BlockArgs h;
    memset(&h, 0, sizeof(h));
    for (int c = 0; c < ncols; c++) h.cols[c] = a.cols[c];
    h.key = key;
    h.key_is_f64 = a.key_is_f64;
    h.shift = shift;
    h.n = n;
    h.nbins = nbins;
    int64_t tile = div_up(tile_guess < 8192 ? 8192 : tile_guess, (int64_t)kBlkThreads * kF8PartRows) * kBlkThreads * kF8PartRows;
    h.tile_rows = tile;
    h.nblocks = div_up(n, tile);
    h.bin_total = a.hist;
    static bool attr_set = false;
    if (!attr_set) {
      CYL_CUDA(cudaFuncSetAttribute(fused8_block_hist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(kBlkMaxBins * 4)));
      attr_set = true;
    }
    ProfScope ps("fused_join_hist", s);
    fused8_block_hist_kernel<<<(unsigned)h.nblocks, kBlkThreads, (size_t)nbins * sizeof(unsigned int), s>>>(h);
    CYL_LAUNCH_CHECK();
  } else {
    ProfScope ps("fused_join_hist", s);
    fused8_hist_kernel<<<f8_grid(n, 256 * kF8PartRows), 256, 0, s>>>(a);
    CYL_LAUNCH_CHECK();
  }
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
  static bool attr_set = false;
  if (!attr_set) {
    CYL_CUDA(cudaFuncSetAttribute(fused8_join_kernel<BW, PW>, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024));
    attr_set = true;
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
    if (lcols[c].byte_width != 8 || lcols[c].validity) return false;
  for (int c = 0; c < nrcols; c++)
    if (rcols[c].byte_width != 8 || rcols[c].validity) return false;
  return true;
}

int fused8_join(const cyl_column_t *bcols, int nbc, int bkey, int64_t nb, const cyl_column_t *pcols, int npc, int pkey, int64_t np,
                void *const *bout, void *const *pout, int64_t capacity, int64_t *num_rows, cudaStream_t s) {
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
  int bits = 0;
  while ((1ll << bits) < nbins) bits++;
  const int shift = 64 - bits;

  Scratch brec, prec, boff, poff, counters;
  CYL_TRY(brec.alloc((size_t)nb * nbc * 8 + 16, s));
  CYL_TRY(prec.alloc((size_t)np * npc * 8 + 16, s));
  CYL_TRY(boff.alloc((size_t)(nbins + 1) * sizeof(int64_t), s));
  CYL_TRY(poff.alloc((size_t)(nbins + 1) * sizeof(int64_t), s));
  CYL_TRY(f8_partition(bcols, nbc, bkey, nb, nbins, shift, brec.as<uint64_t>(), boff.as<int64_t>(), s));
  CYL_TRY(f8_partition(pcols, npc, pkey, np, nbins, shift, prec.as<uint64_t>(), poff.as<int64_t>(), s));

  F8JoinArgs a;
  memset(&a, 0, sizeof(a));
  a.brec = brec.as<uint64_t>();
  a.prec = prec.as<uint64_t>();
  a.b_off = boff.as<int64_t>();
  a.p_off = poff.as<int64_t>();
  for (int c = 0; c < nbc; c++) a.bout[c] = static_cast<uint64_t *>(bout[c]);
  for (int c = 0; c < npc; c++) a.pout[c] = static_cast<uint64_t *>(pout[c]);
  a.nbins = nbins;
  a.bkey = bkey;
  a.pkey = pkey;
  a.key_is_f64 = bcols[bkey].type == CYL_DOUBLE;
  a.cap_rows = cap;
  a.nheads = nheads;
  a.capacity = capacity;
  CYL_TRY(counters.alloc(64, s));
  CYL_CUDA(cudaMemsetAsync(counters.p, 0, 64, s));
  a.out_cursor = counters.as<unsigned long long>();
  a.work_counter = reinterpret_cast<unsigned int *>(counters.as<char>() + 16);
  int64_t grid = (int64_t)num_sms() * 2;
  if (grid > nbins) grid = nbins;
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

}  // namespace cylb
