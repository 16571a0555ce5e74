// common.cuh -- shared host/device helpers of the cylon_b200 CUDA library (sm_100a only).
//
// Device-side column views, the reference-exact row hash (needed for per-rank partition parity) and the
// library's own 64-bit row tags / row equality used by every hash container on the path.
#pragma once

#include <cuda_runtime.h>

#include <mutex>
#include <stdint.h>
#include <stdio.h>

#include "../../include/cylon_b200.h"

namespace cylb {

// ------------------------------------------------------------------------------------------------
// errors: every C-ABI entry returns a cylon::Code and leaves a message for cyl_last_error()
// ------------------------------------------------------------------------------------------------
int fail(int code, const char *fmt, ...);
int cuda_fail(cudaError_t e, const char *expr, const char *file, int line);

#define CYL_CUDA(expr)                                                             \
  do {                                                                             \
    cudaError_t _e = (expr);                                                       \
    if (_e != cudaSuccess) return ::cylb::cuda_fail(_e, #expr, __FILE__, __LINE__); \
  } while (0)

#define CYL_TRY(expr)          \
  do {                         \
    int _rc = (expr);          \
    if (_rc != CYL_OK) return _rc; \
  } while (0)

// every kernel launch is followed by this: counts the launch (cyl_launch_count) and surfaces launch errors
void note_launch();
#define CYL_LAUNCH_CHECK()        \
  do {                            \
    ::cylb::note_launch();        \
    CYL_CUDA(cudaGetLastError()); \
  } while (0)

// Optional per-kernel timing with CUDA events on the launching stream (cyl_profile_enable). Off by default: the
// scope object then costs one branch.
struct ProfScope {
  int slot = -1;
  cudaStream_t s = nullptr;
  ProfScope(const char *name, cudaStream_t stream);
  ~ProfScope();
};

// Caching device allocator behind every buffer the library creates (scratch, join plans, table columns). Blocks are
// rounded to a size class and, when released, parked on a free list keyed by (device, stream, class); a block is only
// handed out again for work on the SAME stream, so stream order makes the reuse safe without events. After warm-up no
// call reaches cudaMalloc/cudaFree. (cudaMallocAsync's pool was measured to spend ~40 ms per step re-mapping physical
// memory once the allocation sizes of the shuffle and the join interleave.)
cudaError_t pool_malloc(void **ptr, size_t bytes, cudaStream_t stream);
cudaError_t pool_free(void *ptr, cudaStream_t stream);
void pool_trim();  // cudaFree every parked block
long long pool_generation();  // bumped by every pool_trim: cached mappings of this process's blocks are stale after a change

// Stream-ordered scratch buffer from that allocator.
struct Scratch {
  void *p = nullptr;
  cudaStream_t s = nullptr;
  Scratch() = default;
  Scratch(const Scratch &) = delete;
  Scratch &operator=(const Scratch &) = delete;
  ~Scratch() { release(); }
  int alloc(size_t bytes, cudaStream_t stream);
  void release();
  template <typename T>
  T *as() const { return reinterpret_cast<T *>(p); }
};

int num_sms();
// "once per device" latch for per-(function, device) settings such as cudaFuncSetAttribute (a per-process flag would leave
// a second device of the process unset). Use as `if (auto once = latch.first()) { ...set... }`: the guard is true for the
// thread that has to do the work and holds the latch's lock until the end of the if statement, so another host thread
// (contexts may be driven from several threads) cannot launch between "marked done" and "attribute set".
struct PerDeviceOnce {
  bool done[64] = {false};
  std::mutex mu;
  struct Guard {
    PerDeviceOnce *latch;
    int dev;
    bool need;
    explicit operator bool() const { return need; }
    ~Guard() {
      if (need && dev >= 0) latch->done[dev] = true;
      latch->mu.unlock();
    }
  };
  Guard first() {
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) dev = -1;
    mu.lock();
    return Guard{this, dev, dev < 0 || !done[dev]};
  }
};
static inline int64_t div_up(int64_t a, int64_t b) { return (a + b - 1) / b; }
static inline int64_t next_pow2(int64_t n) {
  int64_t c = 1;
  while (c < n) c <<= 1;
  return c;
}

// ------------------------------------------------------------------------------------------------
// device column views
// ------------------------------------------------------------------------------------------------
enum HashKind : int32_t { HK_UNSUPPORTED = 0, HK_MODULO = 1, HK_MURMUR_FIXED = 2, HK_MURMUR_VAR = 3 };

struct DCol {
  const uint8_t *data;
  const int32_t *offsets;
  const uint8_t *validity;
  int32_t type;
  int32_t width;
  int32_t kind;       // HashKind
  int32_t is_signed;  // integer kinds only
};

constexpr int kMaxCols = 24;
struct DCols {
  DCol c[kMaxCols];
  int32_t n;
};

// Host: validate and convert C-ABI columns. Returns CYL_OK or an error code (message set).
int make_dcols(const cyl_column_t *cols, int32_t ncols, int64_t nrows, DCols *out, const char *who);
// Host: true when the key is one fixed-width integer/temporal/float column of <= 8 bytes (its bits are the tag).
bool is_exact_key(const DCols &k);

__host__ __device__ inline HashKind hash_kind_of(int32_t type) {
  switch (type) {
    case CYL_UINT8: case CYL_INT8: case CYL_UINT16: case CYL_INT16: case CYL_UINT32: case CYL_INT32:
    case CYL_UINT64: case CYL_INT64: case CYL_DATE32: case CYL_DATE64: case CYL_TIMESTAMP:
    case CYL_TIME32: case CYL_TIME64:
      return HK_MODULO;
    case CYL_FLOAT: case CYL_DOUBLE: case CYL_FIXED_SIZE_BINARY:
      return HK_MURMUR_FIXED;
    case CYL_STRING: case CYL_BINARY:
      return HK_MURMUR_VAR;
    default:
      return HK_UNSUPPORTED;
  }
}
__host__ __device__ inline bool type_is_signed(int32_t type) {
  switch (type) {
    case CYL_INT8: case CYL_INT16: case CYL_INT32: case CYL_INT64: case CYL_DATE32: case CYL_DATE64:
    case CYL_TIMESTAMP: case CYL_TIME32: case CYL_TIME64:
      return true;
    default:
      return false;
  }
}

#ifdef __CUDACC__
// ------------------------------------------------------------------------------------------------
// device code
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t rotl32(uint32_t x, int r) { return __funnelshift_l(x, x, r); }

__device__ __forceinline__ uint32_t fmix32(uint32_t h) {
  h ^= h >> 16;
  h *= 0x85ebca6bu;
  h ^= h >> 13;
  h *= 0xc2b2ae35u;
  h ^= h >> 16;
  return h;
}

__device__ __forceinline__ uint64_t mix64(uint64_t x) {
  x ^= x >> 33;
  x *= 0xff51afd7ed558ccdULL;
  x ^= x >> 33;
  x *= 0xc4ceb9fe1a85ec53ULL;
  x ^= x >> 33;
  return x;
}

__device__ __forceinline__ uint32_t murmur_block(uint32_t h1, uint32_t k1) {
  k1 *= 0xcc9e2d51u;
  k1 = rotl32(k1, 15);
  k1 *= 0x1b873593u;
  h1 ^= k1;
  h1 = rotl32(h1, 13);
  return h1 * 5 + 0xe6546b64u;
}

// MurmurHash3_x86_32 (reference: util/murmur3.cpp:76-117). p may be unaligned.
__device__ inline uint32_t murmur3_32(const uint8_t *p, int len, uint32_t seed) {
  uint32_t h1 = seed;
  const int nblocks = len >> 2;
  if ((reinterpret_cast<uintptr_t>(p) & 3) == 0) {
    const uint32_t *b = reinterpret_cast<const uint32_t *>(p);
    for (int i = 0; i < nblocks; i++) h1 = murmur_block(h1, b[i]);
  } else {
    for (int i = 0; i < nblocks; i++) {
      const uint8_t *q = p + 4 * i;
      uint32_t k1 = (uint32_t)q[0] | ((uint32_t)q[1] << 8) | ((uint32_t)q[2] << 16) | ((uint32_t)q[3] << 24);
      h1 = murmur_block(h1, k1);
    }
  }
  const uint8_t *tail = p + nblocks * 4;
  uint32_t k1 = 0;
  switch (len & 3) {
    case 3: k1 ^= (uint32_t)tail[2] << 16;
    case 2: k1 ^= (uint32_t)tail[1] << 8;
    case 1:
      k1 ^= tail[0];
      k1 *= 0xcc9e2d51u;
      k1 = rotl32(k1, 15);
      k1 *= 0x1b873593u;
      h1 ^= k1;
  }
  h1 ^= (uint32_t)len;
  return fmix32(h1);
}

// murmur3 of one 8-byte / 4-byte value held in registers (float/double keys)
__device__ __forceinline__ uint32_t murmur3_32_u64(uint64_t v, uint32_t seed) {
  uint32_t h1 = murmur_block(seed, (uint32_t)v);
  h1 = murmur_block(h1, (uint32_t)(v >> 32));
  return fmix32(h1 ^ 8u);
}
__device__ __forceinline__ uint32_t murmur3_32_u32(uint32_t v, uint32_t seed) {
  return fmix32(murmur_block(seed, v) ^ 4u);
}

// raw bits of a fixed-width value of <= 8 bytes, zero-extended
__device__ __forceinline__ uint64_t load_bits(const DCol &c, int64_t i) {
  switch (c.width) {
    case 1: return c.data[i];
    case 2: return reinterpret_cast<const uint16_t *>(c.data)[i];
    case 4: return reinterpret_cast<const uint32_t *>(c.data)[i];
    default: return reinterpret_cast<const uint64_t *>(c.data)[i];
  }
}

// the low 32 bits of the value converted as C++ static_cast<uint32_t>(T) would
// (reference: arrow_partition_kernels.cpp:84,104 -- small signed ints are sign-extended)
__device__ __forceinline__ uint32_t modulo_hash(const DCol &c, int64_t i) {
  switch (c.width) {
    case 1: return c.is_signed ? (uint32_t)(int32_t)(int8_t)c.data[i] : (uint32_t)c.data[i];
    case 2: {
      uint16_t v = reinterpret_cast<const uint16_t *>(c.data)[i];
      return c.is_signed ? (uint32_t)(int32_t)(int16_t)v : (uint32_t)v;
    }
    case 4: return reinterpret_cast<const uint32_t *>(c.data)[i];
    default: return (uint32_t)reinterpret_cast<const uint64_t *>(c.data)[i];
  }
}

// per-column value hash of the reference (factory: arrow_partition_kernels.cpp:307-330); validity ignored (A6)
__device__ inline uint32_t ref_value_hash(const DCol &c, int64_t i) {
  if (c.kind == HK_MODULO) return modulo_hash(c, i);
  if (c.kind == HK_MURMUR_FIXED) {
    if (c.width == 8) return murmur3_32_u64(reinterpret_cast<const uint64_t *>(c.data)[i], 0);
    if (c.width == 4) return murmur3_32_u32(reinterpret_cast<const uint32_t *>(c.data)[i], 0);
    return murmur3_32(c.data + i * (int64_t)c.width, c.width, 0);
  }
  const int32_t b = c.offsets[i], e = c.offsets[i + 1];
  return murmur3_32(c.data + b, e - b, 0);
}

// h = 31*h + colhash, over all columns, starting from 0 (TableRowIndexHash, arrow_comparator.cpp:412-440)
__device__ inline uint32_t ref_row_hash(const DCols &k, int64_t i) {
  uint32_t h = 0;
  for (int c = 0; c < k.n; c++) h = 31u * h + ref_value_hash(k.c[c], i);
  return h;
}

__device__ __forceinline__ uint32_t ref_partition_of(uint32_t h, uint32_t p, bool pow2) {
  return pow2 ? (h & (p - 1)) : (h % p);
}

// 64 bits describing one value for the library's own hash containers (NOT the reference hash):
// exact bits for fixed width <= 8, two-seed murmur otherwise.
//
// Identity of floating-point keys is their BIT PATTERN, as it is in the reference: it hashes the bits (murmur3,
// arrow_partition_kernels.cpp:138-143) and its containers only ever compare keys that hashed alike, so -0.0 and +0.0
// are two keys and NaNs with identical bits are one key in Union/Unique/GroupBy and multi-column joins (run against
// the compiled reference in tests/test_gpu_float_keys.py). The one exception is the single-key join, whose reference
// comparator is operator== (arrow_comparator.cpp:240-243): there a NaN key matches nothing (key_bits_is_nan below).
__device__ inline uint64_t value_bits64(const DCol &c, int64_t i) {
  if (c.kind == HK_MURMUR_VAR) {
    const int32_t b = c.offsets[i], e = c.offsets[i + 1];
    return ((uint64_t)murmur3_32(c.data + b, e - b, 0x9747b28cu) << 32) | murmur3_32(c.data + b, e - b, 0);
  }
  if (c.width > 8) {
    const uint8_t *p = c.data + i * (int64_t)c.width;
    return ((uint64_t)murmur3_32(p, c.width, 0x9747b28cu) << 32) | murmur3_32(p, c.width, 0);
  }
  return load_bits(c, i);
}

// NaN test on the raw bits of a FLOAT / DOUBLE key (false for every other type)
__host__ __device__ __forceinline__ bool key_bits_is_nan(int32_t type, uint64_t bits) {
  if (type == CYL_DOUBLE) return (bits & 0x7FFFFFFFFFFFFFFFULL) > 0x7FF0000000000000ULL;
  if (type == CYL_FLOAT) return (bits & 0x7FFFFFFFULL) > 0x7F800000ULL;
  return false;
}

// 64-bit tag of a row over the key columns. exact==true: one fixed-width column, tag == value bits.
__device__ inline uint64_t row_tag(const DCols &k, int64_t i, bool exact, uint64_t seed) {
  if (exact) return value_bits64(k.c[0], i);
  uint64_t h = seed ^ 0x9e3779b97f4a7c15ULL;
  for (int c = 0; c < k.n; c++) h = mix64(h ^ value_bits64(k.c[c], i)) + 0x632be59bd9b4e019ULL;
  return h;
}

// value identity used to verify hashed candidates: numeric by bit pattern (see value_bits64), binary by bytes+length.
__device__ inline bool value_equal(const DCol &a, int64_t i, const DCol &b, int64_t j) {
  if (a.kind == HK_MURMUR_VAR) {
    const int32_t ab = a.offsets[i], al = a.offsets[i + 1] - ab;
    const int32_t bb = b.offsets[j], bl = b.offsets[j + 1] - bb;
    if (al != bl) return false;
    const uint8_t *p = a.data + ab, *q = b.data + bb;
    for (int32_t t = 0; t < al; t++)
      if (p[t] != q[t]) return false;
    return true;
  }
  if (a.width <= 8) return load_bits(a, i) == load_bits(b, j);
  const uint8_t *p = a.data + i * (int64_t)a.width, *q = b.data + j * (int64_t)b.width;
  for (int32_t t = 0; t < a.width; t++)
    if (p[t] != q[t]) return false;
  return true;
}

__device__ inline bool rows_equal(const DCols &a, int64_t i, const DCols &b, int64_t j) {
  for (int c = 0; c < a.n; c++)
    if (!value_equal(a.c[c], i, b.c[c], j)) return false;
  return true;
}

// three-way compare with the ordering the reference sorts by (numeric by value, unsigned for uint64,
// binary lexicographic then by length)
__device__ inline int value_compare(const DCol &a, int64_t i, const DCol &b, int64_t j) {
  if (a.kind == HK_MURMUR_VAR) {
    const int32_t ab = a.offsets[i], al = a.offsets[i + 1] - ab;
    const int32_t bb = b.offsets[j], bl = b.offsets[j + 1] - bb;
    const uint8_t *p = a.data + ab, *q = b.data + bb;
    const int32_t m = al < bl ? al : bl;
    for (int32_t t = 0; t < m; t++)
      if (p[t] != q[t]) return p[t] < q[t] ? -1 : 1;
    return (al > bl) - (al < bl);
  }
  if (a.type == CYL_DOUBLE) {
    double x = reinterpret_cast<const double *>(a.data)[i], y = reinterpret_cast<const double *>(b.data)[j];
    return (x > y) - (x < y);
  }
  if (a.type == CYL_FLOAT) {
    float x = reinterpret_cast<const float *>(a.data)[i], y = reinterpret_cast<const float *>(b.data)[j];
    return (x > y) - (x < y);
  }
  if (a.width > 8) {
    const uint8_t *p = a.data + i * (int64_t)a.width, *q = b.data + j * (int64_t)b.width;
    for (int32_t t = 0; t < a.width; t++)
      if (p[t] != q[t]) return p[t] < q[t] ? -1 : 1;
    return 0;
  }
  uint64_t x = load_bits(a, i), y = load_bits(b, j);
  if (a.is_signed) {
    const int sh = 64 - 8 * a.width;
    int64_t sx = (int64_t)(x << sh) >> sh, sy = (int64_t)(y << sh) >> sh;
    return (sx > sy) - (sx < sy);
  }
  return (x > y) - (x < y);
}

__device__ __forceinline__ bool bit_is_set(const uint8_t *bitmap, int64_t i) { return (bitmap[i >> 3] >> (i & 7)) & 1; }

__device__ __forceinline__ unsigned lane_id() { return threadIdx.x & 31; }
__device__ __forceinline__ unsigned lanemask_lt() {
  unsigned m;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}

// streaming 128-bit load/store that do not pollute L1 (inputs are read once)
__device__ __forceinline__ uint4 ld_stream_u4(const uint4 *p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ uint64_t ld_stream_u64(const uint64_t *p) {
  uint64_t r;
  asm volatile("ld.global.nc.L1::no_allocate.u64 %0, [%1];" : "=l"(r) : "l"(p));
  return r;
}
#endif  // __CUDACC__

}  // namespace cylb
