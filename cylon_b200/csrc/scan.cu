// scan.cu -- host drivers of the device-wide exclusive scan (see scan.cuh).
#include "scan.cuh"

namespace cylb {

template <typename T>
__global__ void __launch_bounds__(kScanThreads) scan_tile_reduce(const T *__restrict__ in, int64_t n, int64_t *__restrict__ tile_sums) {
  __shared__ int64_t wsum[kScanThreads / 32];
  const int64_t base = (int64_t)blockIdx.x * kScanTile;
  int64_t s = 0;
#pragma unroll 4
  for (int r = 0; r < kScanRounds; r++) {
    const int64_t i = base + (int64_t)r * kScanThreads + threadIdx.x;
    if (i < n) s += scan_load(in, i);
  }
  s = warp_reduce_sum(s);
  if (lane_id() == 0) wsum[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    int64_t t = 0;
    for (int w = 0; w < kScanThreads / 32; w++) t += wsum[w];
    tile_sums[blockIdx.x] = t;
  }
}

// one CTA: exclusive scan of the tile sums in place; writes the grand total to *total (may be null)
__global__ void __launch_bounds__(1024) scan_tile_sums(int64_t *__restrict__ tile_sums, int64_t ntiles, int64_t *__restrict__ total) {
  __shared__ int64_t wsum[32];
  __shared__ int64_t carry_s;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  for (int64_t base = 0; base < ntiles; base += 1024) {
    const int64_t i = base + threadIdx.x;
    const int64_t v = i < ntiles ? tile_sums[i] : 0;
    int64_t inc = warp_inclusive_scan(v);
    if (lane_id() == 31) wsum[threadIdx.x >> 5] = inc;
    __syncthreads();
    if (threadIdx.x < 32) {
      int64_t w = wsum[threadIdx.x];
      w = warp_inclusive_scan(w);
      wsum[threadIdx.x] = w;
    }
    __syncthreads();
    const int64_t carry = carry_s;
    const int64_t wbase = (threadIdx.x >> 5) ? wsum[(threadIdx.x >> 5) - 1] : 0;
    if (i < ntiles) tile_sums[i] = carry + wbase + inc - v;
    __syncthreads();
    if (threadIdx.x == 1023) carry_s = carry + wbase + inc;
    __syncthreads();
  }
  if (threadIdx.x == 0 && total) *total = carry_s;
}

template <typename T>
__global__ void __launch_bounds__(kScanThreads) scan_tile_write(const T *__restrict__ in, int64_t n, const int64_t *__restrict__ tile_base,
                                                                  int64_t *__restrict__ out) {
  __shared__ int64_t wsum[kScanThreads / 32];
  const int warp = threadIdx.x >> 5;
  const int64_t wstart = (int64_t)blockIdx.x * kScanTile + (int64_t)warp * (32 * kScanRounds);
  // pass 1: warp total
  int64_t vals[kScanRounds];
  int64_t s = 0;
#pragma unroll
  for (int r = 0; r < kScanRounds; r++) {
    const int64_t i = wstart + r * 32 + lane_id();
    vals[r] = i < n ? scan_load(in, i) : 0;
    s += vals[r];
  }
  s = warp_reduce_sum(s);
  if (lane_id() == 0) wsum[warp] = s;
  __syncthreads();
  int64_t carry = tile_base[blockIdx.x];
  for (int w = 0; w < warp; w++) carry += wsum[w];
#pragma unroll
  for (int r = 0; r < kScanRounds; r++) {
    const int64_t i = wstart + r * 32 + lane_id();
    const int64_t inc = warp_inclusive_scan(vals[r]);
    if (i < n) out[i] = carry + inc - vals[r];
    carry += __shfl_sync(0xffffffffu, inc, 31);
  }
}

template <typename T>
int exclusive_scan(const T *in, int64_t n, int64_t *out, cudaStream_t stream) {
  if (n < 0) return fail(CYL_Invalid, "exclusive_scan: negative length");
  if (n == 0) {
    CYL_CUDA(cudaMemsetAsync(out, 0, sizeof(int64_t), stream));
    return CYL_OK;
  }
  const int64_t ntiles = div_up(n, kScanTile);
  Scratch tiles;
  CYL_TRY(tiles.alloc((size_t)ntiles * sizeof(int64_t), stream));
  scan_tile_reduce<T><<<(unsigned)ntiles, kScanThreads, 0, stream>>>(in, n, tiles.as<int64_t>());
  CYL_LAUNCH_CHECK();
  scan_tile_sums<<<1, 1024, 0, stream>>>(tiles.as<int64_t>(), ntiles, out + n);
  CYL_LAUNCH_CHECK();
  scan_tile_write<T><<<(unsigned)ntiles, kScanThreads, 0, stream>>>(in, n, tiles.as<int64_t>(), out);
  CYL_LAUNCH_CHECK();
  return CYL_OK;
}

// Same scan, but the tile-sum workspace (div_up(n, kScanTile) int64 values) is supplied by the caller: nothing is
// allocated. Used by the entry point that must not allocate.
int exclusive_scan_u32_in_workspace(const uint32_t *in, int64_t n, int64_t *out, int64_t *tile_workspace, cudaStream_t stream) {
  if (n <= 0) return fail(CYL_Invalid, "exclusive_scan_u32_in_workspace: empty input");
  const int64_t ntiles = div_up(n, kScanTile);
  scan_tile_reduce<uint32_t><<<(unsigned)ntiles, kScanThreads, 0, stream>>>(in, n, tile_workspace);
  CYL_LAUNCH_CHECK();
  scan_tile_sums<<<1, 1024, 0, stream>>>(tile_workspace, ntiles, out + n);
  CYL_LAUNCH_CHECK();
  scan_tile_write<uint32_t><<<(unsigned)ntiles, kScanThreads, 0, stream>>>(in, n, tile_workspace, out);
  CYL_LAUNCH_CHECK();
  return CYL_OK;
}

template int exclusive_scan<int64_t>(const int64_t *, int64_t, int64_t *, cudaStream_t);
template int exclusive_scan<uint32_t>(const uint32_t *, int64_t, int64_t *, cudaStream_t);
template int exclusive_scan<int32_t>(const int32_t *, int64_t, int64_t *, cudaStream_t);
template int exclusive_scan<uint8_t>(const uint8_t *, int64_t, int64_t *, cudaStream_t);

}  // namespace cylb

extern "C" int cyl_exclusive_scan_i64(const int64_t *in, int64_t n, int64_t *out, void *stream) {
  if (in == out) return cylb::fail(CYL_Invalid, "cyl_exclusive_scan_i64: in-place scan is not supported");
  return cylb::exclusive_scan<int64_t>(in, n, out, (cudaStream_t)stream);
}
