// gather.cu -- materialisation by row index (copy_array_by_indices) and the order-independent table checksum.
//
// Replaces (reference, relative to cpp/src/cylon/): util/copy_arrray.cpp:24-110,142-292 as used by
// join::util::build_final_table (join/join_utils.cpp:126-181): out[i] = idx[i] == -1 ? null : src[idx[i]].
// The reads are random 8-byte touches of 32-byte sectors -- the HBM-unfriendly part of a join -- so the kernel
// keeps several independent loads in flight per thread and writes the output (and its validity words, built
// with one ballot per warp) fully coalesced.
#include "common.cuh"
#include "scan.cuh"

namespace cylb {

constexpr int kGatherThreads = 256;
constexpr int kGatherItems = 4;

template <typename T>
__global__ void __launch_bounds__(kGatherThreads) gather_fixed_kernel(const T *__restrict__ src, const int64_t *__restrict__ idx, int64_t m,
                                                                        T *__restrict__ dst, uint32_t *__restrict__ validity) {
  const int64_t tile = (int64_t)kGatherThreads * kGatherItems;
  for (int64_t base = (int64_t)blockIdx.x * tile; base < m; base += (int64_t)gridDim.x * tile) {
    int64_t id[kGatherItems];
#pragma unroll
    for (int k = 0; k < kGatherItems; k++) {
      const int64_t i = base + (int64_t)k * kGatherThreads + threadIdx.x;
      id[k] = i < m ? idx[i] : -1;
    }
    T v[kGatherItems];
#pragma unroll
    for (int k = 0; k < kGatherItems; k++) {
      T zero;
      memset(&zero, 0, sizeof(T));
      v[k] = id[k] >= 0 ? src[id[k]] : zero;
    }
#pragma unroll
    for (int k = 0; k < kGatherItems; k++) {
      const int64_t i = base + (int64_t)k * kGatherThreads + threadIdx.x;
      if (i < m) dst[i] = v[k];
      if (validity) {
        const unsigned bits = __ballot_sync(0xffffffffu, id[k] >= 0);  // i >= m gives id == -1 -> bit 0
        if (lane_id() == 0 && (i - lane_id()) < m) validity[i >> 5] = bits;
      }
    }
  }
}

__global__ void gather_bytes_kernel(const uint8_t *__restrict__ src, int32_t w, const int64_t *__restrict__ idx, int64_t m,
                                    uint8_t *__restrict__ dst, uint32_t *__restrict__ validity) {
  const int64_t mpad = (m + 31) & ~31ll;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < mpad; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t id = i < m ? idx[i] : -1;
    if (i < m) {
      uint8_t *t = dst + i * (int64_t)w;
      if (id >= 0) {
        const uint8_t *s = src + id * (int64_t)w;
        for (int b = 0; b < w; b++) t[b] = s[b];
      } else {
        for (int b = 0; b < w; b++) t[b] = 0;
      }
    }
    if (validity) {
      const unsigned bits = __ballot_sync(0xffffffffu, id >= 0);
      if (lane_id() == 0) validity[i >> 5] = bits;
    }
  }
}

__global__ void gather_lengths_kernel(const int32_t *__restrict__ off, const int64_t *__restrict__ idx, int64_t m, int32_t *__restrict__ len) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t id = idx[i];
    len[i] = id >= 0 ? off[id + 1] - off[id] : 0;
  }
}

__global__ void narrow_offsets_kernel2(const int64_t *__restrict__ in, int64_t n1, int32_t *__restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n1; i += (int64_t)gridDim.x * blockDim.x) out[i] = (int32_t)in[i];
}

// 4 lanes cooperate on one output row
__global__ void gather_var_bytes_kernel(const int32_t *__restrict__ soff, const uint8_t *__restrict__ sdata, const int64_t *__restrict__ idx,
                                        int64_t m, const int32_t *__restrict__ doff, uint8_t *__restrict__ ddata) {
  const int64_t gthread = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = gthread >> 2; i < m; i += nthreads >> 2) {
    const int64_t id = idx[i];
    if (id < 0) continue;
    const int32_t b = soff[id], len = soff[id + 1] - b, o = doff[i];
    for (int32_t t = (int32_t)(gthread & 3); t < len; t += 4) ddata[o + t] = sdata[b + t];
  }
}

__global__ void index_validity_kernel(const int64_t *__restrict__ idx, int64_t m, uint32_t *__restrict__ validity) {
  const int64_t mpad = (m + 31) & ~31ll;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < mpad; i += (int64_t)gridDim.x * blockDim.x) {
    const unsigned bits = __ballot_sync(0xffffffffu, i < m && idx[i] >= 0);
    if (lane_id() == 0) validity[i >> 5] = bits;
  }
}

__global__ void count_negative_kernel(const int64_t *__restrict__ idx, int64_t m, unsigned long long *__restrict__ out) {
  unsigned long long c = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (int64_t)gridDim.x * blockDim.x) c += idx[i] < 0;
  c = (unsigned long long)warp_reduce_sum((int64_t)c);
  if (lane_id() == 0 && c) atomicAdd(out, c);
}

__global__ void checksum_kernel(const __grid_constant__ DCols cols, int64_t n, unsigned long long *__restrict__ out) {
  unsigned long long total = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    unsigned long long acc = 0x9e3779b97f4a7c15ULL;
    for (int c = 0; c < cols.n; c++) {
      const DCol &col = cols.c[c];
      const bool valid = col.validity ? bit_is_set(col.validity, i) : true;
      unsigned long long bits = 0;
      if (valid) {
        if (col.kind == HK_MURMUR_VAR) {
          const int32_t b = col.offsets[i], e = col.offsets[i + 1];
          bits = ((unsigned long long)murmur3_32(col.data + b, e - b, 1) << 32) | murmur3_32(col.data + b, e - b, 0);
        } else if (col.width <= 8) {
          bits = load_bits(col, i);
        } else {
          const uint8_t *p = col.data + i * (int64_t)col.width;
          bits = ((unsigned long long)murmur3_32(p, col.width, 1) << 32) | murmur3_32(p, col.width, 0);
        }
      }
      acc = mix64(acc ^ bits) + (valid ? 1ull : 0ull);
    }
    total += acc;
  }
  total = (unsigned long long)warp_reduce_sum((int64_t)total);
  if (lane_id() == 0) atomicAdd(out, total);
}

static inline int grid1d(int64_t n, int per_block = 256) {
  int64_t b = div_up(n, per_block);
  const int64_t cap = (int64_t)num_sms() * 16;
  return (int)(b < 1 ? 1 : (b < cap ? b : cap));
}

int gather_fixed(const void *src, int32_t w, const int64_t *idx, int64_t m, void *dst, uint8_t *validity, cudaStream_t s) {
  if (m == 0) return CYL_OK;
  ProfScope ps("gather_fixed", s);
  uint32_t *v = reinterpret_cast<uint32_t *>(validity);
  const int g = grid1d(m, kGatherThreads * kGatherItems);
  switch (w) {
    case 1: gather_fixed_kernel<uint8_t><<<g, kGatherThreads, 0, s>>>((const uint8_t *)src, idx, m, (uint8_t *)dst, v); break;
    case 2: gather_fixed_kernel<uint16_t><<<g, kGatherThreads, 0, s>>>((const uint16_t *)src, idx, m, (uint16_t *)dst, v); break;
    case 4: gather_fixed_kernel<uint32_t><<<g, kGatherThreads, 0, s>>>((const uint32_t *)src, idx, m, (uint32_t *)dst, v); break;
    case 8: gather_fixed_kernel<uint64_t><<<g, kGatherThreads, 0, s>>>((const uint64_t *)src, idx, m, (uint64_t *)dst, v); break;
    case 16: gather_fixed_kernel<uint4><<<g, kGatherThreads, 0, s>>>((const uint4 *)src, idx, m, (uint4 *)dst, v); break;
    default:
      if (w <= 0) return fail(CYL_Invalid, "gather_fixed: byte_width %d", w);
      gather_bytes_kernel<<<grid1d(m), 256, 0, s>>>((const uint8_t *)src, w, idx, m, (uint8_t *)dst, v);
  }
  CYL_LAUNCH_CHECK();
  return CYL_OK;
}

}  // namespace cylb

using namespace cylb;

extern "C" {

int cyl_gather_fixed(const void *src, int32_t byte_width, const int64_t *idx, int64_t m, void *dst, uint8_t *dst_validity, void *stream) {
  if (m < 0) return fail(CYL_Invalid, "cyl_gather_fixed: negative length");
  return gather_fixed(src, byte_width, idx, m, dst, dst_validity, (cudaStream_t)stream);
}

int cyl_gather_varwidth_offsets(const int32_t *src_offsets, const int64_t *idx, int64_t m, int32_t *dst_offsets, int64_t *total_bytes,
                                void *stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (m == 0) {
    CYL_CUDA(cudaMemsetAsync(dst_offsets, 0, sizeof(int32_t), s));
    CYL_CUDA(cudaStreamSynchronize(s));
    *total_bytes = 0;
    return CYL_OK;
  }
  Scratch lens, offs64;
  CYL_TRY(lens.alloc((size_t)m * sizeof(int32_t), s));
  CYL_TRY(offs64.alloc((size_t)(m + 1) * sizeof(int64_t), s));
  gather_lengths_kernel<<<grid1d(m), 256, 0, s>>>(src_offsets, idx, m, lens.as<int32_t>());
  CYL_LAUNCH_CHECK();
  CYL_TRY(exclusive_scan<int32_t>(lens.as<int32_t>(), m, offs64.as<int64_t>(), s));
  narrow_offsets_kernel2<<<grid1d(m + 1), 256, 0, s>>>(offs64.as<int64_t>(), m + 1, dst_offsets);
  CYL_LAUNCH_CHECK();
  int64_t total = 0;
  CYL_CUDA(cudaMemcpyAsync(&total, offs64.as<int64_t>() + m, sizeof(int64_t), cudaMemcpyDeviceToHost, s));
  CYL_CUDA(cudaStreamSynchronize(s));
  if (total > 0x7FFFFFFFll) return fail(CYL_CapacityError, "gathered variable-width column needs %lld bytes (> 2^31-1, int32 offsets)", (long long)total);
  *total_bytes = total;
  return CYL_OK;
}

int cyl_gather_varwidth_data(const int32_t *src_offsets, const uint8_t *src_data, const int64_t *idx, int64_t m, const int32_t *dst_offsets,
                             uint8_t *dst_data, void *stream) {
  if (m == 0) return CYL_OK;
  gather_var_bytes_kernel<<<grid1d(m * 4), 256, 0, (cudaStream_t)stream>>>(src_offsets, src_data, idx, m, dst_offsets, dst_data);
  CYL_LAUNCH_CHECK();
  return CYL_OK;
}

int cyl_index_validity(const int64_t *idx, int64_t m, uint8_t *validity, void *stream) {
  if (m <= 0) return CYL_OK;
  index_validity_kernel<<<grid1d(m), 256, 0, (cudaStream_t)stream>>>(idx, m, reinterpret_cast<uint32_t *>(validity));
  CYL_LAUNCH_CHECK();
  return CYL_OK;
}

int cyl_count_negative(const int64_t *idx, int64_t m, int64_t *count, void *stream) {
  cudaStream_t s = (cudaStream_t)stream;
  Scratch acc;
  CYL_TRY(acc.alloc(sizeof(unsigned long long), s));
  CYL_CUDA(cudaMemsetAsync(acc.p, 0, sizeof(unsigned long long), s));
  if (m > 0) {
    count_negative_kernel<<<grid1d(m), 256, 0, s>>>(idx, m, acc.as<unsigned long long>());
    CYL_LAUNCH_CHECK();
  }
  unsigned long long h = 0;
  CYL_CUDA(cudaMemcpyAsync(&h, acc.p, sizeof(h), cudaMemcpyDeviceToHost, s));
  CYL_CUDA(cudaStreamSynchronize(s));
  *count = (int64_t)h;
  return CYL_OK;
}

int cyl_table_checksum(const cyl_column_t *cols, int32_t ncols, int64_t nrows, uint64_t *checksum, void *stream) {
  cudaStream_t s = (cudaStream_t)stream;
  DCols d;
  CYL_TRY(make_dcols(cols, ncols, nrows, &d, "cyl_table_checksum"));
  Scratch acc;
  CYL_TRY(acc.alloc(sizeof(unsigned long long), s));
  CYL_CUDA(cudaMemsetAsync(acc.p, 0, sizeof(unsigned long long), s));
  if (nrows > 0) {
    checksum_kernel<<<grid1d(nrows), 256, 0, s>>>(d, nrows, acc.as<unsigned long long>());
    CYL_LAUNCH_CHECK();
  }
  unsigned long long h = 0;
  CYL_CUDA(cudaMemcpyAsync(&h, acc.p, sizeof(h), cudaMemcpyDeviceToHost, s));
  CYL_CUDA(cudaStreamSynchronize(s));
  *checksum = h;
  return CYL_OK;
}

}  // extern "C"
