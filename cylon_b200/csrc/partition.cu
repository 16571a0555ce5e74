// partition.cu -- row hashing, hash partitioning and the stable scatter ("Split") kernels.
//
// Replaces (reference, relative to cpp/src/cylon/): arrow/arrow_partition_kernels.cpp:67-330 (hash kernels),
// partition/partition.cpp:27-166 (MapToHashPartitions, Split) and arrow/arrow_kernels.cpp:33-195 (splitters).
//
// Design (B200): HBM-bound byte shuffling. One WARP owns a tile of consecutive rows and walks it 32 rows at a
// time (256 B coalesced key loads, several rounds in flight). Inside a round `match.any` groups the lanes by
// partition, so a row's stable rank is  cursor[pid] + popc(peers below me)  and one leader lane per group
// advances the warp-private shared-memory cursor -- no atomics, no inter-warp synchronisation, and the result
// is deterministic (rows keep their input order inside a partition, like the reference's builders).
// Pass 1 writes per-tile histograms (partition-major), a device scan turns them into per-tile write cursors,
// pass 2 recomputes the partition ids (cheaper than storing 4 B/row) and either emits dest[] or scatters every
// fixed-width column directly. The scattered stores of one warp form <= P short runs that the 126 MB L2 merges
// into full lines before they reach HBM.
#include <memory>
#include <mutex>
#include <vector>

#include "common.cuh"
#include "scan.cuh"

namespace cylb {

constexpr int kPartThreads = 256;
constexpr int kPartWarps = kPartThreads / 32;
constexpr int kUnroll = 4;  // rounds whose loads are issued together

enum PartMode : int { PM_PID_HIST = 0, PM_TILE_HIST = 1, PM_DEST = 2 };

struct ScatterCols {
  const uint8_t *src[kMaxCols];
  uint8_t *dst[kMaxCols];
  int32_t width[kMaxCols];
  int32_t n;
};

struct PartArgs {
  DCols keys;
  const uint32_t *pid_in;  // when non-null, partition ids are read instead of hashed
  const uint64_t *digit_keys;  // when non-null, pid = 8-bit digit of these keys (radix sort pass)
  int32_t digit_shift;
  uint32_t *pid_out;       // optional
  int64_t n;
  uint32_t P;
  int32_t pow2;
  int32_t rounds;  // 32-row rounds per warp tile (multiple of kUnroll)
  int64_t ntiles;
  uint32_t *tile_hist;                // [P][ntiles]  (PM_TILE_HIST)
  const int64_t *tile_base;           // [P][ntiles]  (PM_DEST)
  int64_t *dest;                      // optional     (PM_DEST)
  unsigned long long *global_hist;    // [P]          (PM_PID_HIST)
  ScatterCols sc;                     // optional fused scatter (PM_DEST)
  const struct MultiDst *multi;       // optional: one destination set per partition (PM_DEST, with sc.src)
};

// Per-partition destinations for the scatter that writes straight into other GPUs' receive buffers: partition p's row
// with stable in-partition rank k lands at dst[c] + (row_base + k) * width. `delta` = row_base - part_offsets[p].
struct MultiDst {
  uint8_t *dst[kMaxCols];
  int64_t delta;
};

__device__ __forceinline__ uint32_t row_pid(const PartArgs &a, int64_t i) {
  if (a.digit_keys) return (uint32_t)(a.digit_keys[i] >> a.digit_shift) & 0xFFu;
  if (a.pid_in) {  // caller-supplied ids index shared-memory counters: an id outside [0, P) marks the row as skipped
    const uint32_t p = a.pid_in[i];
    return p < a.P ? p : 0xFFFFFFFFu;
  }
  return ref_partition_of(ref_row_hash(a.keys, i), a.P, a.pow2 != 0);
}

__device__ __forceinline__ void scatter_one(const ScatterCols &sc, int c, int64_t i, int64_t d) {
  switch (sc.width[c]) {
    case 8: reinterpret_cast<uint64_t *>(sc.dst[c])[d] = reinterpret_cast<const uint64_t *>(sc.src[c])[i]; break;
    case 4: reinterpret_cast<uint32_t *>(sc.dst[c])[d] = reinterpret_cast<const uint32_t *>(sc.src[c])[i]; break;
    case 16: reinterpret_cast<uint4 *>(sc.dst[c])[d] = reinterpret_cast<const uint4 *>(sc.src[c])[i]; break;
    case 2: reinterpret_cast<uint16_t *>(sc.dst[c])[d] = reinterpret_cast<const uint16_t *>(sc.src[c])[i]; break;
    case 1: sc.dst[c][d] = sc.src[c][i]; break;
    default: {
      const int w = sc.width[c];
      const uint8_t *s = sc.src[c] + i * (int64_t)w;
      uint8_t *t = sc.dst[c] + d * (int64_t)w;
      for (int b = 0; b < w; b++) t[b] = s[b];
    }
  }
}

__device__ __forceinline__ void scatter_to(uint8_t *dst, int64_t d, const uint8_t *src, int64_t i, int w) {
  switch (w) {
    case 8: reinterpret_cast<uint64_t *>(dst)[d] = reinterpret_cast<const uint64_t *>(src)[i]; break;
    case 4: reinterpret_cast<uint32_t *>(dst)[d] = reinterpret_cast<const uint32_t *>(src)[i]; break;
    case 16: reinterpret_cast<uint4 *>(dst)[d] = reinterpret_cast<const uint4 *>(src)[i]; break;
    case 2: reinterpret_cast<uint16_t *>(dst)[d] = reinterpret_cast<const uint16_t *>(src)[i]; break;
    default: dst[d] = src[i]; break;
  }
}

template <int MODE>
__global__ void __launch_bounds__(kPartThreads) partition_tiles_kernel(const __grid_constant__ PartArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int warp = threadIdx.x >> 5;
  const unsigned lane = lane_id();
  const unsigned lt = lanemask_lt();
  // warp-private cursors: u32 counts (hist modes) or i64 destinations (dest mode)
  uint32_t *cnt32 = reinterpret_cast<uint32_t *>(smem_raw) + (size_t)warp * a.P;
  int64_t *cur64 = reinterpret_cast<int64_t *>(smem_raw) + (size_t)warp * a.P;
  const int64_t tile_rows = (int64_t)a.rounds * 32;
  const int64_t warps_total = (int64_t)gridDim.x * kPartWarps;

  if (MODE == PM_PID_HIST) {
    for (uint32_t p = lane; p < a.P; p += 32) cnt32[p] = 0;
    __syncwarp();
  }

  for (int64_t tile = (int64_t)blockIdx.x * kPartWarps + warp; tile < a.ntiles; tile += warps_total) {
    if (MODE == PM_TILE_HIST) {
      for (uint32_t p = lane; p < a.P; p += 32) cnt32[p] = 0;
    } else if (MODE == PM_DEST) {
      for (uint32_t p = lane; p < a.P; p += 32) cur64[p] = a.tile_base[(int64_t)p * a.ntiles + tile];
    }
    __syncwarp();
    const int64_t start = tile * tile_rows;
    for (int r0 = 0; r0 < a.rounds; r0 += kUnroll) {
      uint32_t pid[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; u++) {
        const int64_t i = start + (int64_t)(r0 + u) * 32 + lane;
        pid[u] = i < a.n ? row_pid(a, i) : 0xFFFFFFFFu;
      }
      int64_t d[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; u++) {
        const int64_t i = start + (int64_t)(r0 + u) * 32 + lane;
        const bool valid = pid[u] != 0xFFFFFFFFu;  // past the end, or a caller-supplied id >= P
        const unsigned peers = __match_any_sync(0xffffffffu, pid[u]);
        const unsigned rank = __popc(peers & lt);
        if (MODE == PM_DEST) {
          d[u] = valid ? cur64[pid[u]] + rank : -1;
          __syncwarp();
          if (valid && rank == 0) cur64[pid[u]] += __popc(peers);
          __syncwarp();
        } else {
          if (valid && rank == 0) cnt32[pid[u]] += __popc(peers);
          __syncwarp();
          if (MODE == PM_PID_HIST && valid && a.pid_out) a.pid_out[i] = pid[u];
        }
      }
      if (MODE == PM_DEST) {
        if (a.dest) {
#pragma unroll
          for (int u = 0; u < kUnroll; u++) {
            const int64_t i = start + (int64_t)(r0 + u) * 32 + lane;
            if (d[u] >= 0) a.dest[i] = d[u];
          }
        }
        if (a.pid_out) {
#pragma unroll
          for (int u = 0; u < kUnroll; u++) {
            const int64_t i = start + (int64_t)(r0 + u) * 32 + lane;
            if (d[u] >= 0) a.pid_out[i] = pid[u];
          }
        }
        if (a.multi) {  // destinations differ per partition (local buffer or a peer GPU's buffer over NVLink)
          for (int c = 0; c < a.sc.n; c++) {
#pragma unroll
            for (int u = 0; u < kUnroll; u++) {
              const int64_t i = start + (int64_t)(r0 + u) * 32 + lane;
              if (d[u] >= 0) {
                const MultiDst &m = a.multi[pid[u]];
                scatter_to(m.dst[c], d[u] + m.delta, a.sc.src[c], i, a.sc.width[c]);
              }
            }
          }
        } else {
          for (int c = 0; c < a.sc.n; c++) {
#pragma unroll
            for (int u = 0; u < kUnroll; u++) {
              const int64_t i = start + (int64_t)(r0 + u) * 32 + lane;
              if (d[u] >= 0) scatter_one(a.sc, c, i, d[u]);
            }
          }
        }
      }
    }
    if (MODE == PM_TILE_HIST) {
      __syncwarp();
      for (uint32_t p = lane; p < a.P; p += 32) a.tile_hist[(int64_t)p * a.ntiles + tile] = cnt32[p];
      __syncwarp();
    }
  }
  if (MODE == PM_PID_HIST) {
    __syncwarp();
    for (uint32_t p = lane; p < a.P; p += 32)
      if (cnt32[p]) atomicAdd(&a.global_hist[p], (unsigned long long)cnt32[p]);
  }
}

__global__ void extract_part_offsets(const int64_t *__restrict__ base, int64_t ntiles, uint32_t P, int64_t *__restrict__ part_offsets) {
  const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p <= P) part_offsets[p] = base[(int64_t)p * ntiles];
}

__global__ void row_hash_kernel(const __grid_constant__ DCols keys, int64_t n, uint32_t *__restrict__ h) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) h[i] = ref_row_hash(keys, i);
}

// ---- staged scatter for the peer-memory shuffle ------------------------------------------------------------------
// Measured with direct stores (partition_tiles_kernel<PM_DEST> + MultiDst): 394 GB/s over NVLink at 2 GPUs, 310 GB/s at
// 4 -- every warp instruction sends 32/P rows (128 B, then 64 B, then 32 B at 8 GPUs) to each peer, and remote stores
// are not merged by the local L2. Here a warp first sorts its whole tile by partition inside shared memory (same
// stable ranks as above), then streams each partition's run to its destination with full 256-byte warp stores.
constexpr int kStageMaxP = 32;

struct StageArgs {
  PartArgs a;              // keys, n, P, rounds, ntiles, tile_hist (pass-1 counts), tile_base, multi, sc
  int32_t col_off[kMaxCols];  // byte offset of column c inside a warp's staging area
  int32_t warp_bytes;         // staging bytes per warp (all columns) + the three small per-warp tables
  int32_t bulk;               // 1: all columns are 8 or 16 bytes wide -> runs leave through cp.async.bulk (TMA) stores
  int32_t key_col;            // the one 8-byte key column's index among the scattered columns, or -1 (generic row hash)
  int32_t key_kind;           // HashKind of that column
};

// Reference hash of a single 8-byte key from its bits (ref_value_hash without the column descriptor).
__device__ __forceinline__ uint32_t key8_hash(uint64_t v, int kind) { return kind == HK_MODULO ? (uint32_t)v : murmur3_32_u64(v, 0); }

// Host side: keys are exactly one 8-byte MODULO / MURMUR_FIXED column.
static bool single_key8(const DCols &k) { return k.n == 1 && k.c[0].width == 8 && (k.c[0].kind == HK_MODULO || k.c[0].kind == HK_MURMUR_FIXED); }

// One partition's run of one column, shared memory -> (peer) global memory, as a single asynchronous bulk store.
// Both addresses are 16-byte aligned and `bytes` is a positive multiple of 16.
__device__ __forceinline__ void bulk_store_run(void *dst, const void *src_smem, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"((uint32_t)__cvta_generic_to_shared(src_smem)), "r"(bytes)
               : "memory");
}

__global__ void __launch_bounds__(kPartThreads) partition_stage_scatter_kernel(const __grid_constant__ StageArgs s) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const PartArgs &a = s.a;
  const int warp = threadIdx.x >> 5;
  const unsigned lane = lane_id();
  const unsigned lt = lanemask_lt();
  unsigned char *wbase = smem_raw + (size_t)warp * s.warp_bytes;
  uint32_t *loff = reinterpret_cast<uint32_t *>(wbase);      // [P + 1] first staging slot of each partition
  uint32_t *run = loff + (kStageMaxP + 1);                    // [P] rows of the partition staged so far
  unsigned char *stage = wbase + (2 * kStageMaxP + 4) * sizeof(uint32_t);
  const int64_t tile_rows = (int64_t)a.rounds * 32;
  const int64_t warps_total = (int64_t)gridDim.x * kPartWarps;

  for (int64_t tile = (int64_t)blockIdx.x * kPartWarps + warp; tile < a.ntiles; tile += warps_total) {
    {  // partition counts of this tile (pass 1) -> staging offsets
      uint32_t c = lane < a.P ? a.tile_hist[(int64_t)lane * a.ntiles + tile] : 0u;
      uint32_t odd = 0;
      if (s.bulk && lane < a.P) {
        // A run's staging slot gets the parity of its destination row, so that source and destination of the bulk store
        // are 16-byte aligned together: every region starts on an even slot and has an even size.
        odd = (uint32_t)((a.tile_base[(int64_t)lane * a.ntiles + tile] + a.multi[lane].delta) & 1);
        c = (c + odd + 1) & ~1u;
      }
      uint32_t inc = c;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= (unsigned)d) inc += t;
      }
      if (lane < a.P) {
        loff[lane] = inc - c + odd;
        run[lane] = 0;
      }
      if (lane == 31) loff[a.P] = inc;
    }
    __syncwarp();
    const int64_t start = tile * tile_rows;
    for (int r0 = 0; r0 < a.rounds; r0 += kUnroll) {
      uint32_t pid[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; u++) {
        const int64_t i = start + (int64_t)(r0 + u) * 32 + lane;
        pid[u] = i < a.n ? row_pid(a, i) : 0xFFFFFFFFu;
      }
      uint32_t slot[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; u++) {
        const bool valid = pid[u] != 0xFFFFFFFFu;
        const unsigned peers = __match_any_sync(0xffffffffu, pid[u]);
        const unsigned rank = __popc(peers & lt);
        slot[u] = valid ? loff[pid[u]] + run[pid[u]] + rank : 0xFFFFFFFFu;
        __syncwarp();
        if (valid && rank == 0) run[pid[u]] += __popc(peers);
        __syncwarp();
      }
      for (int c = 0; c < a.sc.n; c++) {
        unsigned char *col = stage + (size_t)s.col_off[c];
#pragma unroll
        for (int u = 0; u < kUnroll; u++) {
          const int64_t i = start + (int64_t)(r0 + u) * 32 + lane;
          if (slot[u] != 0xFFFFFFFFu) scatter_to(col, slot[u], a.sc.src[c], i, a.sc.width[c]);
        }
      }
    }
    __syncwarp();
    if (s.bulk) {  // lane p hands partition p's runs to the TMA engine; odd first/last 8-byte elements go as plain stores
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      if (lane < a.P && run[lane] != 0) {
        const uint32_t first = loff[lane], cnt = run[lane];
        const MultiDst &m = a.multi[lane];
        const int64_t gbase = a.tile_base[(int64_t)lane * a.ntiles + tile] + m.delta;
        const uint32_t head = (uint32_t)(gbase & 1);
        const uint32_t mid = (cnt - head) & ~1u;
        for (int c = 0; c < a.sc.n; c++) {
          const unsigned char *col = stage + (size_t)s.col_off[c];
          if (a.sc.width[c] == 16) {
            bulk_store_run(m.dst[c] + (size_t)gbase * 16, col + (size_t)first * 16, cnt * 16);
          } else {
            const uint64_t *src64 = reinterpret_cast<const uint64_t *>(col) + first;
            uint64_t *dst64 = reinterpret_cast<uint64_t *>(m.dst[c]) + gbase;
            if (head) dst64[0] = src64[0];
            if (mid) bulk_store_run(dst64 + head, src64 + head, mid * 8);
            if (head + mid < cnt) dst64[cnt - 1] = src64[cnt - 1];
          }
        }
      }
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");  // the staging area is rewritten by the next tile
      __syncwarp();
      continue;
    }
    for (uint32_t p = 0; p < a.P; p++) {  // stream every partition's run to its (possibly remote) destination
      const uint32_t first = loff[p], cnt = loff[p + 1] - first;
      if (cnt == 0) continue;
      const MultiDst &m = a.multi[p];
      const int64_t gbase = a.tile_base[(int64_t)p * a.ntiles + tile] + m.delta;
      for (int c = 0; c < a.sc.n; c++) {
        const unsigned char *col = stage + (size_t)s.col_off[c];
        if (a.sc.width[c] == 8) {  // 8-byte columns: 128-bit stores (512 B per warp instruction) once the destination is 16-byte aligned
          const uint64_t *src64 = reinterpret_cast<const uint64_t *>(col) + first;
          uint64_t *dst64 = reinterpret_cast<uint64_t *>(m.dst[c]) + gbase;
          const uint32_t head = (uint32_t)(gbase & 1);  // element offsets are 8 bytes: odd gbase = misaligned by 8
          if (head && lane == 0) dst64[0] = src64[0];
          const uint32_t npairs = (cnt - head) >> 1;
          for (uint32_t q = lane; q < npairs; q += 32) {
            const uint32_t k = head + 2 * q;
            *reinterpret_cast<ulonglong2 *>(dst64 + k) = make_ulonglong2(src64[k], src64[k + 1]);
          }
          if (((cnt - head) & 1) && lane == 0) dst64[cnt - 1] = src64[cnt - 1];
        } else {
          for (uint32_t k = lane; k < cnt; k += 32) scatter_to(m.dst[c], gbase + k, col, first + k, a.sc.width[c]);
        }
      }
    }
    __syncwarp();
  }
}

static int pick_rounds(int64_t n, uint32_t P) {
  // keep #counters (ntiles * P) well below #rows, and tiles small enough to spread over 148 SMs x 64 warps
  int rounds = 32;
  while (rounds < 4096 && (int64_t)rounds * 32 < (int64_t)P * 8) rounds *= 2;
  (void)n;
  return rounds;
}

// All-8-byte tables (the join benchmark's int64 key + double payload, up to three columns). The generic kernel above walks a
// tile in groups of kUnroll rounds and waits for DRAM once per group and once more per column: ~10 dependent round trips per
// 512-row tile, and ncu shows the warps parked on the long scoreboard with 17% issue utilisation at 16 warps/SM. Here
// every load a tile needs -- pass-1 counts, destination bases, all keys, all payloads -- is issued before anything is
// consumed, so a tile costs one round trip; the values wait in registers, are sorted by partition into the warp's
// staging area, and leave as one bulk (TMA) store per partition and column.
template <int NC, int KC>  // KC: index of the single 8-byte key among the columns, -1 = generic row hash
__global__ void __launch_bounds__(kPartThreads, NC == 3 ? 1 : 2) partition_stage_scatter8_kernel(const __grid_constant__ StageArgs s) {
  constexpr int R = 16;  // 32-row rounds per tile, == PartArgs::rounds for staged plans
  constexpr bool FASTKEY = KC >= 0;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const PartArgs &a = s.a;
  const int warp = threadIdx.x >> 5;
  const unsigned lane = lane_id();
  const unsigned lt = lanemask_lt();
  unsigned char *wbase = smem_raw + (size_t)warp * s.warp_bytes;
  uint32_t *loff = reinterpret_cast<uint32_t *>(wbase);
  uint32_t *run = loff + (kStageMaxP + 1);
  unsigned char *stage = wbase + (2 * kStageMaxP + 4) * sizeof(uint32_t);
  const int64_t warps_total = (int64_t)gridDim.x * kPartWarps;

  for (int64_t tile = (int64_t)blockIdx.x * kPartWarps + warp; tile < a.ntiles; tile += warps_total) {
    const int64_t start = tile * (int64_t)(R * 32);
    uint32_t cnt = 0;
    int64_t gbase = 0;
    uint8_t *dstp[NC];
    if (lane < a.P) {
      cnt = a.tile_hist[(int64_t)lane * a.ntiles + tile];
      const MultiDst &m = a.multi[lane];
      gbase = a.tile_base[(int64_t)lane * a.ntiles + tile] + m.delta;
#pragma unroll
      for (int c = 0; c < NC; c++) dstp[c] = m.dst[c];
    }
    uint64_t v[NC][R];
#pragma unroll
    for (int u = 0; u < R; u++) {
      const int64_t i = start + u * 32 + lane;
#pragma unroll
      for (int c = 0; c < NC; c++) v[c][u] = i < a.n ? __ldg(reinterpret_cast<const uint64_t *>(a.sc.src[c]) + i) : 0ull;
    }
    uint32_t pid[FASTKEY ? 1 : R];
    if (!FASTKEY) {
#pragma unroll
      for (int u = 0; u < R; u++) {
        const int64_t i = start + u * 32 + lane;
        pid[u] = i < a.n ? row_pid(a, i) : 0xFFFFFFFFu;
      }
    }
    const uint32_t odd = (uint32_t)(gbase & 1);  // staging slot parity follows the destination row: see the generic kernel
    {
      const uint32_t padded = lane < a.P ? ((cnt + odd + 1) & ~1u) : 0u;
      uint32_t inc = padded;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= (unsigned)d) inc += t;
      }
      if (lane < a.P) {
        loff[lane] = inc - padded + odd;
        run[lane] = 0;
      }
    }
    __syncwarp();
#pragma unroll
    for (int u = 0; u < R; u++) {
      uint32_t q;
      if (FASTKEY) {  // the key is one of the values already in registers: no second load, no descriptor decoding
        const uint64_t kv = v[FASTKEY ? KC : 0][u];
        q = start + u * 32 + lane < a.n ? ref_partition_of(key8_hash(kv, s.key_kind), a.P, a.pow2 != 0) : 0xFFFFFFFFu;
      } else {
        q = pid[u];
      }
      const bool valid = q != 0xFFFFFFFFu;
      const unsigned peers = __match_any_sync(0xffffffffu, q);
      const unsigned rank = __popc(peers & lt);
      const uint32_t slot = valid ? loff[q] + run[q] + rank : 0u;
      __syncwarp();
      if (valid && rank == 0) run[q] += __popc(peers);
      if (valid) {
#pragma unroll
        for (int c = 0; c < NC; c++) reinterpret_cast<uint64_t *>(stage + (size_t)s.col_off[c])[slot] = v[c][u];
      }
      __syncwarp();
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (lane < a.P && cnt != 0) {
      const uint32_t first = loff[lane];
      const uint32_t mid = (cnt - odd) & ~1u;
#pragma unroll
      for (int c = 0; c < NC; c++) {
        const uint64_t *src64 = reinterpret_cast<const uint64_t *>(stage + (size_t)s.col_off[c]) + first;
        uint64_t *dst64 = reinterpret_cast<uint64_t *>(dstp[c]) + gbase;
        if (odd) dst64[0] = src64[0];
        if (mid) bulk_store_run(dst64 + odd, src64 + odd, mid * 8);
        if (odd + mid < cnt) dst64[cnt - 1] = src64[cnt - 1];
      }
    }
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    __syncwarp();
  }
}

// Pass 1 for staged plans with one 8-byte key and at most 8 partitions. The generic histogram spends ~80 instructions
// per 32 rows (descriptor decoding, match.any, shared-memory counters: 42% issue utilisation, 1.4 TB/s). Here a tile's 16
// key loads go out together and every lane counts in registers, 16 bits per partition packed in two 64-bit words
// (a 512-row tile cannot overflow 16 bits); five butterfly steps add the lanes up.
__global__ void __launch_bounds__(kPartThreads, 3) partition_tile_hist_k8_kernel(const __grid_constant__ PartArgs a, int key_kind) {
  constexpr int R = 16;
  const int warp = threadIdx.x >> 5;
  const unsigned lane = lane_id();
  const int64_t warps_total = (int64_t)gridDim.x * kPartWarps;
  const uint64_t *__restrict__ key = reinterpret_cast<const uint64_t *>(a.keys.c[0].data);
  for (int64_t tile = (int64_t)blockIdx.x * kPartWarps + warp; tile < a.ntiles; tile += warps_total) {
    const int64_t start = tile * (int64_t)(R * 32);
    uint64_t kv[R];
#pragma unroll
    for (int u = 0; u < R; u++) {
      const int64_t i = start + u * 32 + lane;
      kv[u] = i < a.n ? __ldg(key + i) : 0ull;
    }
    uint64_t lo = 0, hi = 0;  // partitions 0-3 and 4-7
#pragma unroll
    for (int u = 0; u < R; u++) {
      const int64_t i = start + u * 32 + lane;
      const uint32_t pid = ref_partition_of(key8_hash(kv[u], key_kind), a.P, a.pow2 != 0);
      const uint64_t one = (i < a.n ? 1ull : 0ull) << ((pid & 3) * 16);
      lo += pid < 4 ? one : 0ull;
      hi += pid < 4 ? 0ull : one;
    }
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) {
      lo += __shfl_xor_sync(0xffffffffu, lo, d);
      hi += __shfl_xor_sync(0xffffffffu, hi, d);
    }
    if (lane < a.P) a.tile_hist[(int64_t)lane * a.ntiles + tile] = (uint32_t)(((lane < 4 ? lo : hi) >> ((lane & 3) * 16)) & 0xFFFFu);
  }
}

static int grid_for(int64_t ntiles, int blocks_per_sm) {
  int64_t want = div_up(ntiles, kPartWarps);
  int64_t cap = (int64_t)num_sms() * blocks_per_sm;
  return (int)(want < cap ? (want < 1 ? 1 : want) : cap);
}

template <int MODE>
static int launch_part(const PartArgs &a, size_t smem, cudaStream_t stream) {
  static PerDeviceOnce attr_once;  // the attribute belongs to the (function, device) pair
  if (smem > 48 * 1024) {
    if (auto once_guard = attr_once.first()) CYL_CUDA(cudaFuncSetAttribute(partition_tiles_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
  }
  const int bps = smem > 16 * 1024 ? 2 : 8;
  partition_tiles_kernel<MODE><<<grid_for(a.ntiles, bps), kPartThreads, smem, stream>>>(a);
  CYL_LAUNCH_CHECK();
  return CYL_OK;
}

// shared driver: tile histogram -> scan -> dest / fused scatter
static int partition_map_impl(const DCols *keys, const uint32_t *pid_in, int64_t n, uint32_t P, uint32_t *pid_out, int64_t *dest,
                              const ScatterCols *sc, int64_t *part_offsets, cudaStream_t stream, const char *who,
                              const uint64_t *digit_keys = nullptr, int digit_shift = 0) {
  if (P == 0 || P > 1024) return fail(CYL_Invalid, "%s: num_partitions must be in [1, 1024], got %u", who, P);
  if (n < 0 || n > 0xFFFFFFFEll) return fail(CYL_CapacityError, "%s: %lld rows (max 2^32-2 per GPU)", who, (long long)n);
  PartArgs a;
  memset(&a, 0, sizeof(a));
  if (keys) a.keys = *keys;
  a.pid_in = pid_in;
  a.digit_keys = digit_keys;
  a.digit_shift = digit_shift;
  a.pid_out = pid_out;
  a.n = n;
  a.P = P;
  a.pow2 = (P & (P - 1)) == 0;
  a.rounds = pick_rounds(n, P);
  a.ntiles = n == 0 ? 0 : div_up(n, (int64_t)a.rounds * 32);
  if (n == 0) {
    if (part_offsets) CYL_CUDA(cudaMemsetAsync(part_offsets, 0, (P + 1) * sizeof(int64_t), stream));
    return CYL_OK;
  }
  const int64_t ncounters = a.ntiles * P;
  Scratch hist, base;
  CYL_TRY(hist.alloc((size_t)ncounters * sizeof(uint32_t), stream));
  CYL_TRY(base.alloc((size_t)(ncounters + 1) * sizeof(int64_t), stream));
  a.tile_hist = hist.as<uint32_t>();
  {
    ProfScope ps(digit_keys ? "radix_tile_hist" : "partition_tile_hist", stream);
    CYL_TRY(launch_part<PM_TILE_HIST>(a, (size_t)kPartWarps * P * sizeof(uint32_t), stream));
  }
  {
    ProfScope ps(digit_keys ? "radix_scan" : "partition_scan", stream);
    CYL_TRY(exclusive_scan<uint32_t>(hist.as<uint32_t>(), ncounters, base.as<int64_t>(), stream));
  }
  if (part_offsets) {
    extract_part_offsets<<<(P + 1 + 255) / 256, 256, 0, stream>>>(base.as<int64_t>(), a.ntiles, P, part_offsets);
    CYL_LAUNCH_CHECK();
  }
  a.tile_base = base.as<int64_t>();
  a.dest = dest;
  if (sc) a.sc = *sc;
  {
    ProfScope ps(digit_keys ? "radix_scatter" : "partition_scatter", stream);
    CYL_TRY(launch_part<PM_DEST>(a, (size_t)kPartWarps * P * sizeof(int64_t), stream));
  }
  return CYL_OK;
}

// ---- element-wise scatter helpers --------------------------------------------------------------------------
template <typename T>
__global__ void scatter_fixed_kernel(const T *__restrict__ src, const int64_t *__restrict__ dest, int64_t n, T *__restrict__ dst) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) dst[dest[i]] = src[i];
}

__global__ void scatter_bytes_kernel(const uint8_t *__restrict__ src, int32_t w, const int64_t *__restrict__ dest, int64_t n,
                                     uint8_t *__restrict__ dst) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const uint8_t *s = src + i * (int64_t)w;
    uint8_t *t = dst + dest[i] * (int64_t)w;
    for (int b = 0; b < w; b++) t[b] = s[b];
  }
}

__global__ void scatter_lengths_kernel(const int32_t *__restrict__ off, const int64_t *__restrict__ dest, int64_t n,
                                       int32_t *__restrict__ len_out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    len_out[dest[i]] = off[i + 1] - off[i];
}

__global__ void narrow_offsets_kernel(const int64_t *__restrict__ in, int64_t n1, int32_t *__restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n1; i += (int64_t)gridDim.x * blockDim.x) out[i] = (int32_t)in[i];
}

// one warp copies one row's bytes (rows are short strings; lanes stride over the bytes)
__global__ void scatter_var_bytes_kernel(const int32_t *__restrict__ soff, const uint8_t *__restrict__ sdata,
                                         const int64_t *__restrict__ dest, int64_t n, const int32_t *__restrict__ doff,
                                         uint8_t *__restrict__ ddata) {
  const int64_t gthread = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;
  // 4 lanes per row
  for (int64_t i = gthread >> 2; i < n; i += nthreads >> 2) {
    const int32_t b = soff[i], len = soff[i + 1] - b;
    const int32_t o = doff[dest ? dest[i] : i];
    for (int32_t t = (int32_t)(gthread & 3); t < len; t += 4) ddata[o + t] = sdata[b + t];
  }
}

__global__ void scatter_validity_kernel(const uint8_t *__restrict__ vsrc, const int64_t *__restrict__ dest, int64_t n,
                                        uint32_t *__restrict__ vdst) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    if (bit_is_set(vsrc, i)) {
      const int64_t d = dest[i];
      atomicOr(&vdst[d >> 5], 1u << (d & 31));
    }
  }
}

static inline int ew_grid(int64_t n, int threads = 256) {
  int64_t b = div_up(n, threads);
  int64_t cap = (int64_t)num_sms() * 16;
  return (int)(b < 1 ? 1 : (b < cap ? b : cap));
}

int scatter_fixed(const void *src, int32_t w, const int64_t *dest, int64_t n, void *dst, cudaStream_t stream) {
  if (n == 0) return CYL_OK;
  const int g = ew_grid(n);
  switch (w) {
    case 1: scatter_fixed_kernel<uint8_t><<<g, 256, 0, stream>>>((const uint8_t *)src, dest, n, (uint8_t *)dst); break;
    case 2: scatter_fixed_kernel<uint16_t><<<g, 256, 0, stream>>>((const uint16_t *)src, dest, n, (uint16_t *)dst); break;
    case 4: scatter_fixed_kernel<uint32_t><<<g, 256, 0, stream>>>((const uint32_t *)src, dest, n, (uint32_t *)dst); break;
    case 8: scatter_fixed_kernel<uint64_t><<<g, 256, 0, stream>>>((const uint64_t *)src, dest, n, (uint64_t *)dst); break;
    case 16: scatter_fixed_kernel<uint4><<<g, 256, 0, stream>>>((const uint4 *)src, dest, n, (uint4 *)dst); break;
    default:
      if (w <= 0) return fail(CYL_Invalid, "scatter_fixed: byte_width %d", w);
      scatter_bytes_kernel<<<g, 256, 0, stream>>>((const uint8_t *)src, w, dest, n, (uint8_t *)dst);
  }
  CYL_LAUNCH_CHECK();
  return CYL_OK;
}

// ---- stable scatter map for callers that cannot allocate ---------------------------------------------------------
// The three scratch arrays of the scatter map (per-tile histograms, their scanned cursors, the scan's tile sums) live
// in ONE file-scope device array instead of coming from cudaMallocAsync. Its size bounds the problem:
// tiles x partitions <= kStaticMaxCounters.
constexpr int64_t kStaticMaxCounters = 262144;
struct StaticScatterScratch {
  int64_t base[kStaticMaxCounters + 1];                       // scanned cursors, [P][ntiles] (+ total)
  int64_t scan_tiles[kStaticMaxCounters / kScanTile + 2];     // tile sums of the device scan
  uint32_t hist[kStaticMaxCounters];                          // per-tile histograms, [P][ntiles]
};
static __device__ __align__(256) unsigned char g_static_scatter_scratch[sizeof(StaticScatterScratch)];

// In host code the name above is only a handle: the address a kernel can use has to be asked for, once per device and
// with that device current. Cached per device ordinal.
static int static_scatter_scratch(StaticScatterScratch **out) {
  static std::mutex mu;
  static StaticScatterScratch *per_device[64] = {nullptr};
  int dev = -1;
  CYL_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return fail(CYL_Invalid, "cyl_scatter_map_static: device ordinal %d not supported", dev);
  std::lock_guard<std::mutex> lk(mu);
  if (per_device[dev] == nullptr) {
    void *p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_static_scatter_scratch);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetSymbolAddress(g_static_scatter_scratch)", __FILE__, __LINE__);
    per_device[dev] = static_cast<StaticScatterScratch *>(p);
  }
  *out = per_device[dev];
  return CYL_OK;
}

// one stable 8-bit LSD radix pass over (u64 key, u32 value) pairs; shares the warp-tile ranking above
int radix_pass_pairs(const uint64_t *keys_in, const uint32_t *vals_in, uint64_t *keys_out, uint32_t *vals_out, int64_t n, int shift,
                     cudaStream_t stream) {
  ScatterCols sc;
  memset(&sc, 0, sizeof(sc));
  sc.n = 2;
  sc.src[0] = reinterpret_cast<const uint8_t *>(keys_in);
  sc.dst[0] = reinterpret_cast<uint8_t *>(keys_out);
  sc.width[0] = 8;
  sc.src[1] = reinterpret_cast<const uint8_t *>(vals_in);
  sc.dst[1] = reinterpret_cast<uint8_t *>(vals_out);
  sc.width[1] = 4;
  return partition_map_impl(nullptr, nullptr, n, 256, nullptr, nullptr, &sc, nullptr, stream, "radix_pass", keys_in, shift);
}

}  // namespace cylb

using namespace cylb;

extern "C" {

int cyl_row_hash(const cyl_column_t *cols, int32_t ncols, int64_t nrows, uint32_t *h, void *stream) {
  DCols k;
  CYL_TRY(make_dcols(cols, ncols, nrows, &k, "cyl_row_hash"));
  if (nrows == 0) return CYL_OK;
  row_hash_kernel<<<ew_grid(nrows), 256, 0, (cudaStream_t)stream>>>(k, nrows, h);
  CYL_LAUNCH_CHECK();
  return CYL_OK;
}

int cyl_hash_partition(const cyl_column_t *cols, int32_t ncols, int64_t nrows, uint32_t P, uint32_t *pid, uint64_t *hist, void *stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (ncols < 1) return fail(CYL_Invalid, "cyl_hash_partition: needs at least one key column");
  if (P == 0 || P > 4096) return fail(CYL_Invalid, "cyl_hash_partition: num_partitions must be in [1, 4096], got %u", P);
  PartArgs a;
  memset(&a, 0, sizeof(a));
  CYL_TRY(make_dcols(cols, ncols, nrows, &a.keys, "cyl_hash_partition"));
  CYL_CUDA(cudaMemsetAsync(hist, 0, P * sizeof(uint64_t), s));
  if (nrows == 0) return CYL_OK;
  a.pid_out = pid;
  a.n = nrows;
  a.P = P;
  a.pow2 = (P & (P - 1)) == 0;
  a.rounds = 32;
  a.ntiles = div_up(nrows, 32 * 32);
  a.global_hist = reinterpret_cast<unsigned long long *>(hist);
  return launch_part<PM_PID_HIST>(a, (size_t)kPartWarps * P * sizeof(uint32_t), s);
}

int cyl_scatter_map(const uint32_t *pid, int64_t nrows, uint32_t P, int64_t *dest, int64_t *part_offsets, void *stream) {
  if (nrows > 0 && (!pid || !dest)) return fail(CYL_Invalid, "cyl_scatter_map: null pid/dest");
  return partition_map_impl(nullptr, pid, nrows, P, nullptr, dest, nullptr, part_offsets, (cudaStream_t)stream, "cyl_scatter_map");
}

int cyl_scatter_map_static(const uint32_t *pid, int64_t nrows, uint32_t P, int64_t *dest, int64_t *part_offsets, void *stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (P == 0 || P > 1024) return fail(CYL_Invalid, "cyl_scatter_map_static: num_partitions must be in [1, 1024], got %u", P);
  if (nrows < 0) return fail(CYL_Invalid, "cyl_scatter_map_static: negative row count");
  if (!part_offsets || (nrows > 0 && (!pid || !dest))) return fail(CYL_Invalid, "cyl_scatter_map_static: null pid/dest/part_offsets");
  PartArgs a;
  memset(&a, 0, sizeof(a));
  a.pid_in = pid;
  a.n = nrows;
  a.P = P;
  a.pow2 = (P & (P - 1)) == 0;
  a.rounds = pick_rounds(nrows, P);
  a.ntiles = nrows == 0 ? 0 : div_up(nrows, (int64_t)a.rounds * 32);
  const int64_t ncounters = a.ntiles * P;
  if (ncounters > kStaticMaxCounters)
    return fail(CYL_CapacityError,
                "cyl_scatter_map_static: %lld rows x %u partitions need %lld counters, the static scratch holds %lld "
                "(about %lld rows at this partition count); use cyl_scatter_map for larger inputs",
                (long long)nrows, P, (long long)ncounters, (long long)kStaticMaxCounters,
                (long long)(kStaticMaxCounters / P * a.rounds * 32));
  StaticScatterScratch *ws = nullptr;
  CYL_TRY(static_scatter_scratch(&ws));  // first call on a device resolves the symbol address
  if (nrows == 0) {
    CYL_CUDA(cudaMemsetAsync(part_offsets, 0, (P + 1) * sizeof(int64_t), s));
    return CYL_OK;
  }
  // Every word of the scratch that is read below is written first by this call (each tile stores all P of its
  // counters, the scans overwrite their outputs), so nothing left behind by an earlier call can leak in.
  a.tile_hist = ws->hist;
  CYL_TRY(launch_part<PM_TILE_HIST>(a, (size_t)kPartWarps * P * sizeof(uint32_t), s));
  CYL_TRY(exclusive_scan_u32_in_workspace(ws->hist, ncounters, ws->base, ws->scan_tiles, s));
  extract_part_offsets<<<(P + 1 + 255) / 256, 256, 0, s>>>(ws->base, a.ntiles, P, part_offsets);
  CYL_LAUNCH_CHECK();
  a.tile_base = ws->base;
  a.dest = dest;
  CYL_TRY(launch_part<PM_DEST>(a, (size_t)kPartWarps * P * sizeof(int64_t), s));
  return CYL_OK;
}

int cyl_partition_map(const cyl_column_t *cols, int32_t ncols, int64_t nrows, uint32_t P, uint32_t *pid_out, int64_t *dest,
                      int64_t *part_offsets, void *stream) {
  if (ncols < 1) return fail(CYL_Invalid, "cyl_partition_map: needs at least one key column");
  DCols k;
  CYL_TRY(make_dcols(cols, ncols, nrows, &k, "cyl_partition_map"));
  return partition_map_impl(&k, nullptr, nrows, P, pid_out, dest, nullptr, part_offsets, (cudaStream_t)stream, "cyl_partition_map");
}

int cyl_scatter_fixed(const void *src, int32_t byte_width, const int64_t *dest, int64_t nrows, void *dst, void *stream) {
  return scatter_fixed(src, byte_width, dest, nrows, dst, (cudaStream_t)stream);
}

int cyl_scatter_varwidth(const int32_t *src_offsets, const uint8_t *src_data, const int64_t *dest, int64_t nrows, int32_t *dst_offsets,
                         uint8_t *dst_data, void *stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (nrows == 0) {
    CYL_CUDA(cudaMemsetAsync(dst_offsets, 0, sizeof(int32_t), s));
    return CYL_OK;
  }
  Scratch lens, offs64;
  CYL_TRY(lens.alloc((size_t)nrows * sizeof(int32_t), s));
  CYL_TRY(offs64.alloc((size_t)(nrows + 1) * sizeof(int64_t), s));
  scatter_lengths_kernel<<<ew_grid(nrows), 256, 0, s>>>(src_offsets, dest, nrows, lens.as<int32_t>());
  CYL_LAUNCH_CHECK();
  CYL_TRY(exclusive_scan<int32_t>(lens.as<int32_t>(), nrows, offs64.as<int64_t>(), s));
  narrow_offsets_kernel<<<ew_grid(nrows + 1), 256, 0, s>>>(offs64.as<int64_t>(), nrows + 1, dst_offsets);
  CYL_LAUNCH_CHECK();
  scatter_var_bytes_kernel<<<ew_grid(nrows * 4), 256, 0, s>>>(src_offsets, src_data, dest, nrows, dst_offsets, dst_data);
  CYL_LAUNCH_CHECK();
  return CYL_OK;
}

int cyl_scatter_validity(const uint8_t *vsrc, const int64_t *dest, int64_t nrows, uint8_t *vdst, void *stream) {
  cudaStream_t s = (cudaStream_t)stream;
  CYL_CUDA(cudaMemsetAsync(vdst, 0, (size_t)div_up(nrows, 32) * 4, s));
  if (nrows == 0) return CYL_OK;
  scatter_validity_kernel<<<ew_grid(nrows), 256, 0, s>>>(vsrc, dest, nrows, reinterpret_cast<uint32_t *>(vdst));
  CYL_LAUNCH_CHECK();
  return CYL_OK;
}

int cyl_partition_scatter_fixed(const cyl_column_t *cols, int32_t ncols, const int32_t *key_col_idx, int32_t nkeys, int64_t nrows,
                                uint32_t P, void *const *dst, int64_t *part_offsets, void *stream) {
  if (ncols < 1 || ncols > kMaxCols) return fail(CYL_Invalid, "cyl_partition_scatter_fixed: %d columns (1..%d)", ncols, kMaxCols);
  if (nkeys < 1 || nkeys > ncols) return fail(CYL_Invalid, "cyl_partition_scatter_fixed: %d key columns", nkeys);
  cyl_column_t kc[kMaxCols];
  for (int i = 0; i < nkeys; i++) {
    if (key_col_idx[i] < 0 || key_col_idx[i] >= ncols) return fail(CYL_IndexError, "cyl_partition_scatter_fixed: key column %d out of range", key_col_idx[i]);
    kc[i] = cols[key_col_idx[i]];
  }
  DCols k;
  CYL_TRY(make_dcols(kc, nkeys, nrows, &k, "cyl_partition_scatter_fixed"));
  ScatterCols sc;
  memset(&sc, 0, sizeof(sc));
  sc.n = ncols;
  for (int c = 0; c < ncols; c++) {
    if (cols[c].type == CYL_STRING || cols[c].type == CYL_BINARY || cols[c].byte_width <= 0)
      return fail(CYL_Invalid, "cyl_partition_scatter_fixed: column %d is not fixed width", c);
    if (cols[c].length != nrows) return fail(CYL_Invalid, "cyl_partition_scatter_fixed: column %d length mismatch", c);
    sc.src[c] = static_cast<const uint8_t *>(cols[c].data);
    sc.dst[c] = static_cast<uint8_t *>(dst[c]);
    sc.width[c] = cols[c].byte_width;
  }
  return partition_map_impl(&k, nullptr, nrows, P, nullptr, nullptr, &sc, part_offsets, (cudaStream_t)stream, "cyl_partition_scatter_fixed");
}

// ---- two-step form: count first, scatter later into per-partition destinations (fused partition + shuffle) -----------
struct cyl_partition_plan {
  PartArgs a;
  Scratch hist, base, multi;
  std::vector<int64_t> part_off;  // host copy, [P + 1]
  int ncols = 0;
  bool staged = false;
};

int cyl_partition_plan_create(const cyl_column_t *cols, int32_t ncols, const int32_t *key_col_idx, int32_t nkeys, int64_t nrows, uint32_t P,
                              cyl_partition_plan_t **plan, int64_t *part_rows, void *stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (!plan || !part_rows) return fail(CYL_Invalid, "cyl_partition_plan_create: null output");
  *plan = nullptr;
  if (ncols < 1 || ncols > kMaxCols) return fail(CYL_Invalid, "cyl_partition_plan_create: %d columns (1..%d)", ncols, kMaxCols);
  if (nkeys < 1 || nkeys > ncols) return fail(CYL_Invalid, "cyl_partition_plan_create: %d key columns", nkeys);
  if (P == 0 || P > 1024) return fail(CYL_Invalid, "cyl_partition_plan_create: num_partitions must be in [1, 1024], got %u", P);
  cyl_column_t kc[kMaxCols];
  for (int i = 0; i < nkeys; i++) {
    if (key_col_idx[i] < 0 || key_col_idx[i] >= ncols) return fail(CYL_IndexError, "cyl_partition_plan_create: key column %d out of range", key_col_idx[i]);
    kc[i] = cols[key_col_idx[i]];
  }
  std::unique_ptr<cyl_partition_plan> p(new cyl_partition_plan());
  PartArgs &a = p->a;
  memset(&a, 0, sizeof(a));
  CYL_TRY(make_dcols(kc, nkeys, nrows, &a.keys, "cyl_partition_plan_create"));
  a.sc.n = ncols;
  for (int c = 0; c < ncols; c++) {
    const int w = cols[c].byte_width;
    if (cols[c].type == CYL_STRING || cols[c].type == CYL_BINARY || !(w == 1 || w == 2 || w == 4 || w == 8 || w == 16))
      return fail(CYL_Invalid, "cyl_partition_plan_create: column %d is not a 1/2/4/8/16-byte fixed-width column", c);
    if (cols[c].length != nrows) return fail(CYL_Invalid, "cyl_partition_plan_create: column %d length mismatch", c);
    a.sc.src[c] = static_cast<const uint8_t *>(cols[c].data);
    a.sc.width[c] = w;
  }
  p->ncols = ncols;
  a.n = nrows;
  a.P = P;
  a.pow2 = (P & (P - 1)) == 0;
  a.rounds = pick_rounds(nrows, P);
  int row_bytes = 0;
  for (int c = 0; c < ncols; c++) row_bytes += a.sc.width[c];
  p->staged = P <= (uint32_t)kStageMaxP && row_bytes <= 24;  // 512-row warp tiles staged in <= 12 KB of shared memory each
  if (p->staged) a.rounds = 16;  // 512-row tiles; 128- and 256-row tiles measured the same (2.64-2.73 ms at 4 GPUs)
  a.ntiles = nrows == 0 ? 0 : div_up(nrows, (int64_t)a.rounds * 32);
  p->part_off.assign(P + 1, 0);
  if (nrows > 0) {
    const int64_t ncounters = a.ntiles * P;
    CYL_TRY(p->hist.alloc((size_t)ncounters * sizeof(uint32_t), s));
    CYL_TRY(p->base.alloc((size_t)(ncounters + 1) * sizeof(int64_t), s));
    a.tile_hist = p->hist.as<uint32_t>();
    {
      ProfScope ps("partition_tile_hist", s);
      if (p->staged && P <= 8 && single_key8(a.keys)) {
        partition_tile_hist_k8_kernel<<<grid_for(a.ntiles, 3), kPartThreads, 0, s>>>(a, a.keys.c[0].kind);
        CYL_LAUNCH_CHECK();
      } else {
        CYL_TRY(launch_part<PM_TILE_HIST>(a, (size_t)kPartWarps * P * sizeof(uint32_t), s));
      }
    }
    CYL_TRY(exclusive_scan<uint32_t>(p->hist.as<uint32_t>(), ncounters, p->base.as<int64_t>(), s));
    Scratch offs;
    CYL_TRY(offs.alloc((P + 1) * sizeof(int64_t), s));
    extract_part_offsets<<<(P + 1 + 255) / 256, 256, 0, s>>>(p->base.as<int64_t>(), a.ntiles, P, offs.as<int64_t>());
    CYL_LAUNCH_CHECK();
    CYL_CUDA(cudaMemcpyAsync(p->part_off.data(), offs.p, (P + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost, s));
    CYL_CUDA(cudaStreamSynchronize(s));
    a.tile_base = p->base.as<int64_t>();
  }
  for (uint32_t q = 0; q < P; q++) part_rows[q] = p->part_off[q + 1] - p->part_off[q];
  *plan = p.release();
  return CYL_OK;
}

int cyl_partition_plan_scatter(cyl_partition_plan_t *plan, void *const *dst, const int64_t *dst_row_base, void *stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (!plan || !dst || !dst_row_base) return fail(CYL_Invalid, "cyl_partition_plan_scatter: null argument");
  PartArgs a = plan->a;
  if (a.n == 0) return CYL_OK;
  std::vector<MultiDst> table(a.P);
  for (uint32_t p = 0; p < a.P; p++) {
    memset(&table[p], 0, sizeof(MultiDst));
    for (int c = 0; c < plan->ncols; c++) table[p].dst[c] = static_cast<uint8_t *>(dst[(size_t)p * plan->ncols + c]);
    table[p].delta = dst_row_base[p] - plan->part_off[p];
  }
  CYL_TRY(plan->multi.alloc(table.size() * sizeof(MultiDst), s));
  CYL_CUDA(cudaMemcpyAsync(plan->multi.p, table.data(), table.size() * sizeof(MultiDst), cudaMemcpyHostToDevice, s));
  CYL_CUDA(cudaStreamSynchronize(s));  // `table` is pageable host memory: finish the copy before it goes out of scope
  a.multi = plan->multi.as<MultiDst>();
  ProfScope ps("partition_scatter_peer", s);
  if (plan->staged) {
    StageArgs sa;
    memset(&sa, 0, sizeof(sa));
    sa.a = a;
    const int tile_rows = a.rounds * 32;
    sa.bulk = 1;
    for (int c = 0; c < plan->ncols; c++) sa.bulk &= (a.sc.width[c] == 8 || a.sc.width[c] == 16);
    const int stage_rows = tile_rows + (sa.bulk ? 2 * (int)a.P : 0);  // bulk mode pads every run to an even slot and size
    int off = 0;
    for (int c = 0; c < plan->ncols; c++) {
      sa.col_off[c] = off;
      off += (stage_rows * a.sc.width[c] + 15) / 16 * 16;
    }
    sa.warp_bytes = (2 * kStageMaxP + 4) * (int)sizeof(uint32_t) + off;
    const size_t smem = (size_t)sa.warp_bytes * kPartWarps;
    static PerDeviceOnce attr_once;
    if (auto once_guard = attr_once.first()) {
      CYL_CUDA(cudaFuncSetAttribute(partition_stage_scatter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024));
    }
    bool all8 = sa.bulk && plan->ncols <= 3 && a.rounds == 16;
    for (int c = 0; c < plan->ncols; c++) all8 = all8 && a.sc.width[c] == 8;
    sa.key_col = -1;
    if (single_key8(a.keys) && !a.pid_in && !a.digit_keys)
      for (int c = 0; c < plan->ncols; c++)
        if (a.sc.src[c] == a.keys.c[0].data) {
          sa.key_col = c;
          sa.key_kind = a.keys.c[0].kind;
        }
    if (all8) {
      void (*kern)(StageArgs) = nullptr;
      switch (plan->ncols * 4 + sa.key_col + 1) {  // (columns, key position + 1)
        case 4: kern = partition_stage_scatter8_kernel<1, -1>; break;
        case 5: kern = partition_stage_scatter8_kernel<1, 0>; break;
        case 8: kern = partition_stage_scatter8_kernel<2, -1>; break;
        case 9: kern = partition_stage_scatter8_kernel<2, 0>; break;
        case 10: kern = partition_stage_scatter8_kernel<2, 1>; break;
        case 12: kern = partition_stage_scatter8_kernel<3, -1>; break;
        case 13: kern = partition_stage_scatter8_kernel<3, 0>; break;
        case 14: kern = partition_stage_scatter8_kernel<3, 1>; break;
        default: kern = partition_stage_scatter8_kernel<3, 2>; break;
      }
      CYL_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024));
      kern<<<grid_for(a.ntiles, plan->ncols == 3 ? 1 : 2), kPartThreads, smem, s>>>(sa);
      CYL_LAUNCH_CHECK();
      return CYL_OK;
    }
    const int bps = (int)((200 * 1024) / (smem + 1024)) < 1 ? 1 : (int)((200 * 1024) / (smem + 1024));
    partition_stage_scatter_kernel<<<grid_for(a.ntiles, bps > 8 ? 8 : bps), kPartThreads, smem, s>>>(sa);
    CYL_LAUNCH_CHECK();
    return CYL_OK;
  }
  return launch_part<PM_DEST>(a, (size_t)kPartWarps * a.P * sizeof(int64_t), s);
}

void cyl_partition_plan_free(cyl_partition_plan_t *plan) { delete plan; }

}  // extern "C"
