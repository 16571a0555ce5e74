// scan.cuh -- device-wide exclusive prefix sum (int64 accumulation) used for output offsets everywhere:
// reduce per tile -> scan of tile sums (one CTA) -> rescan tiles with their base. All loads are coalesced:
// a warp walks its chunk 32 consecutive elements at a time and carries the running sum in a register.
#pragma once

#include "common.cuh"

namespace cylb {

constexpr int kScanThreads = 256;
constexpr int kScanRounds = 16;  // 32-element rounds per warp
constexpr int kScanTile = kScanThreads * kScanRounds;

#ifdef __CUDACC__
template <typename T>
__device__ __forceinline__ int64_t scan_load(const T *in, int64_t i) {
  return (int64_t)in[i];
}

__device__ __forceinline__ int64_t warp_inclusive_scan(int64_t v) {
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    int64_t t = __shfl_up_sync(0xffffffffu, v, d);
    if (lane_id() >= (unsigned)d) v += t;
  }
  return v;
}

__device__ __forceinline__ int64_t warp_reduce_sum(int64_t v) {
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
  return v;
}

#endif

// out has n+1 entries; out[n] = total. Scratch is taken from the stream-ordered pool.
template <typename T>
int exclusive_scan(const T *in, int64_t n, int64_t *out, cudaStream_t stream);

// non-allocating form: tile_workspace holds div_up(n, kScanTile) int64 values supplied by the caller
int exclusive_scan_u32_in_workspace(const uint32_t *in, int64_t n, int64_t *out, int64_t *tile_workspace, cudaStream_t stream);

}  // namespace cylb
