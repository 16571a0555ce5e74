// runtime.cu -- error plumbing, device memory helpers and column validation of the C-ABI.
#include <stdarg.h>
#include <string.h>

#include <atomic>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "common.cuh"

namespace cylb {

static thread_local char g_err[1024] = "";

int fail(int code, const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

int cuda_fail(cudaError_t e, const char *expr, const char *file, int line) {
  const char *base = strrchr(file, '/');
  snprintf(g_err, sizeof(g_err), "CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString(e), base ? base + 1 : file, line, expr);
  return e == cudaErrorMemoryAllocation ? CYL_OutOfMemory : CYL_ExecutionError;
}

// ---- caching allocator ------------------------------------------------------------------------------------------
namespace {
struct PoolKey {
  int dev;
  cudaStream_t stream;
  size_t cls;
  bool operator<(const PoolKey &o) const {
    if (dev != o.dev) return dev < o.dev;
    if (stream != o.stream) return stream < o.stream;
    return cls < o.cls;
  }
};
std::mutex g_pool_mu;
std::map<PoolKey, std::vector<void *>> g_pool_free;
std::map<void *, std::pair<int, size_t>> g_pool_live;  // ptr -> (device, class)

size_t size_class(size_t bytes) {
  if (bytes < 512) return 512;
  // keep the top 3 bits after the leading one: classes are at most 12.5 % larger than the request
  int lead = 63 - __builtin_clzll((unsigned long long)bytes);
  const size_t step = (size_t)1 << (lead > 3 ? lead - 3 : 0);
  return (bytes + step - 1) / step * step;
}
}  // namespace

static std::atomic<long long> g_pool_generation{0};
long long pool_generation() { return g_pool_generation.load(); }

void pool_trim() {
  g_pool_generation.fetch_add(1);  // addresses handed to peers (CUDA IPC) may be recycled for other allocations from here on
  std::lock_guard<std::mutex> lk(g_pool_mu);
  for (auto &kv : g_pool_free)
    for (void *p : kv.second) {
      cudaFree(p);
      g_pool_live.erase(p);
    }
  g_pool_free.clear();
}

cudaError_t pool_malloc(void **ptr, size_t bytes, cudaStream_t stream) {
  *ptr = nullptr;
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  const size_t cls = size_class(bytes);
  {
    std::lock_guard<std::mutex> lk(g_pool_mu);
    auto it = g_pool_free.find(PoolKey{dev, stream, cls});
    if (it != g_pool_free.end() && !it->second.empty()) {
      *ptr = it->second.back();
      it->second.pop_back();
      return cudaSuccess;
    }
  }
  void *p = nullptr;
  e = cudaMalloc(&p, cls);
  if (e == cudaErrorMemoryAllocation) {  // give parked blocks back to the driver and try once more
    (void)cudaGetLastError();
    pool_trim();
    e = cudaMalloc(&p, cls);
  }
  if (e != cudaSuccess) return e;
  std::lock_guard<std::mutex> lk(g_pool_mu);
  g_pool_live[p] = std::make_pair(dev, cls);
  *ptr = p;
  return cudaSuccess;
}

cudaError_t pool_free(void *ptr, cudaStream_t stream) {
  if (!ptr) return cudaSuccess;
  std::lock_guard<std::mutex> lk(g_pool_mu);
  auto it = g_pool_live.find(ptr);
  if (it == g_pool_live.end()) return cudaFree(ptr);  // not ours
  g_pool_free[PoolKey{it->second.first, stream, it->second.second}].push_back(ptr);
  return cudaSuccess;
}

int Scratch::alloc(size_t bytes, cudaStream_t stream) {
  release();
  s = stream;
  if (bytes == 0) bytes = 16;
  cudaError_t e = pool_malloc(&p, bytes, stream);
  if (e != cudaSuccess) {
    p = nullptr;
    return cuda_fail(e, "pool_malloc(scratch)", __FILE__, __LINE__);
  }
  return CYL_OK;
}

void Scratch::release() {
  if (p) {
    pool_free(p, s);
    p = nullptr;
  }
}

// ---- launch counter + event profiler ------------------------------------------------------------------------
static std::atomic<long long> g_launches{0};
void note_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

struct ProfRecord {
  const char *name;
  cudaEvent_t start, stop;
};
static std::atomic<int> g_prof_enabled{0};
static std::mutex g_prof_mu;
static std::vector<ProfRecord> g_prof_records;

ProfScope::ProfScope(const char *name, cudaStream_t stream) : s(stream) {
  if (!g_prof_enabled.load(std::memory_order_relaxed)) return;
  ProfRecord r;
  r.name = name;
  if (cudaEventCreate(&r.start) != cudaSuccess || cudaEventCreate(&r.stop) != cudaSuccess) return;
  cudaEventRecord(r.start, stream);
  std::lock_guard<std::mutex> lk(g_prof_mu);
  g_prof_records.push_back(r);
  slot = (int)g_prof_records.size() - 1;
}

ProfScope::~ProfScope() {
  if (slot < 0) return;
  std::lock_guard<std::mutex> lk(g_prof_mu);
  cudaEventRecord(g_prof_records[slot].stop, s);
}

int num_sms() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

int make_dcols(const cyl_column_t *cols, int32_t ncols, int64_t nrows, DCols *out, const char *who) {
  if (ncols < 0 || ncols > kMaxCols) return fail(CYL_Invalid, "%s: %d columns (max %d)", who, ncols, kMaxCols);
  if (ncols > 0 && cols == nullptr) return fail(CYL_Invalid, "%s: null column array", who);
  if (nrows < 0 || nrows > 0xFFFFFFFEll) return fail(CYL_CapacityError, "%s: %lld rows (max 2^32-2 per GPU)", who, (long long)nrows);
  memset(out, 0, sizeof(*out));
  out->n = ncols;
  for (int c = 0; c < ncols; c++) {
    const cyl_column_t &s = cols[c];
    DCol &d = out->c[c];
    d.kind = hash_kind_of(s.type);
    if (d.kind == HK_UNSUPPORTED)
      return fail(CYL_ExecutionError, "%s: column %d has unsupported type id %d", who, c, s.type);  // partition.cpp:102-104
    if (s.length != nrows) return fail(CYL_Invalid, "%s: column %d has %lld rows, expected %lld", who, c, (long long)s.length, (long long)nrows);
    if (d.kind == HK_MURMUR_VAR) {
      if (nrows > 0 && s.offsets == nullptr) return fail(CYL_Invalid, "%s: column %d is variable width but has no offsets", who, c);
      d.width = 0;
    } else {
      d.width = s.byte_width;
      if (d.kind == HK_MODULO || s.type == CYL_FLOAT || s.type == CYL_DOUBLE) {
        int want = 0;
        switch (s.type) {
          case CYL_UINT8: case CYL_INT8: want = 1; break;
          case CYL_UINT16: case CYL_INT16: want = 2; break;
          case CYL_UINT32: case CYL_INT32: case CYL_FLOAT: case CYL_DATE32: case CYL_TIME32: want = 4; break;
          default: want = 8; break;
        }
        if (d.width != want) return fail(CYL_Invalid, "%s: column %d type %d needs byte_width %d, got %d", who, c, s.type, want, d.width);
      } else if (d.width <= 0) {
        return fail(CYL_Invalid, "%s: column %d fixed-size binary needs byte_width > 0", who, c);
      }
    }
    if (nrows > 0 && s.data == nullptr && d.kind != HK_MURMUR_VAR) return fail(CYL_Invalid, "%s: column %d has no data", who, c);
    d.data = static_cast<const uint8_t *>(s.data);
    d.offsets = s.offsets;
    d.validity = s.validity;
    d.type = s.type;
    d.is_signed = type_is_signed(s.type) ? 1 : 0;
  }
  return CYL_OK;
}

bool is_exact_key(const DCols &k) {
  return k.n == 1 && k.c[0].kind != HK_MURMUR_VAR && k.c[0].width <= 8 && k.c[0].type != CYL_FIXED_SIZE_BINARY;
}

}  // namespace cylb

using namespace cylb;

extern "C" {

int cyl_version(void) { return 100; }

const char *cyl_last_error(void) { return g_err; }

int cyl_init(int32_t device) {
  CYL_CUDA(cudaSetDevice(device));
  CYL_CUDA(cudaFree(nullptr));  // create the context now, not inside the first operator
  return CYL_OK;
}

long long cyl_launch_count(void) { return g_launches.load(); }

int cyl_profile_enable(int32_t on) {
  g_prof_enabled.store(on ? 1 : 0);
  return CYL_OK;
}

// Synchronises the device, folds the recorded scopes into "name,launches,total_ms\n" lines and clears them.
int cyl_profile_report(char *buf, int64_t buf_len) {
  CYL_CUDA(cudaDeviceSynchronize());
  std::lock_guard<std::mutex> lk(g_prof_mu);
  std::map<std::string, std::pair<long long, double>> agg;
  for (auto &r : g_prof_records) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, r.start, r.stop) == cudaSuccess) {
      auto &a = agg[r.name];
      a.first += 1;
      a.second += ms;
    }
    cudaEventDestroy(r.start);
    cudaEventDestroy(r.stop);
  }
  g_prof_records.clear();
  std::string out;
  for (auto &kv : agg) {
    char line[256];
    snprintf(line, sizeof(line), "%s,%lld,%.6f\n", kv.first.c_str(), kv.second.first, kv.second.second);
    out += line;
  }
  if (buf && buf_len > 0) {
    const size_t n = out.size() < (size_t)buf_len - 1 ? out.size() : (size_t)buf_len - 1;
    memcpy(buf, out.data(), n);
    buf[n] = 0;
  }
  return CYL_OK;
}

int cyl_device_count(int32_t *count) {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess) {
    *count = 0;
    return cuda_fail(e, "cudaGetDeviceCount", __FILE__, __LINE__);
  }
  *count = n;
  return CYL_OK;
}

int cyl_malloc(void **ptr, int64_t bytes) {
  if (bytes < 0) return fail(CYL_Invalid, "cyl_malloc: negative size");
  *ptr = nullptr;
  CYL_CUDA(cudaMalloc(ptr, bytes == 0 ? 16 : (size_t)bytes));
  return CYL_OK;
}

int cyl_free(void *ptr) {
  if (ptr) CYL_CUDA(cudaFree(ptr));
  return CYL_OK;
}

int cyl_memset(void *ptr, int32_t value, int64_t bytes, void *stream) {
  if (bytes > 0) CYL_CUDA(cudaMemsetAsync(ptr, value, (size_t)bytes, (cudaStream_t)stream));
  return CYL_OK;
}

int cyl_memcpy_h2d(void *dst, const void *src, int64_t bytes, void *stream) {
  if (bytes > 0) CYL_CUDA(cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyHostToDevice, (cudaStream_t)stream));
  return CYL_OK;
}

int cyl_memcpy_d2h(void *dst, const void *src, int64_t bytes, void *stream) {
  if (bytes > 0) CYL_CUDA(cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  return CYL_OK;
}

int cyl_memcpy_d2d(void *dst, const void *src, int64_t bytes, void *stream) {
  if (bytes > 0) CYL_CUDA(cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  return CYL_OK;
}

int cyl_stream_sync(void *stream) {
  CYL_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  return CYL_OK;
}

int cyl_host_alloc_pinned(void **ptr, int64_t bytes) {
  *ptr = nullptr;
  CYL_CUDA(cudaHostAlloc(ptr, bytes <= 0 ? 16 : (size_t)bytes, cudaHostAllocDefault));
  return CYL_OK;
}

int cyl_host_free_pinned(void *ptr) {
  if (ptr) CYL_CUDA(cudaFreeHost(ptr));
  return CYL_OK;
}

}  // extern "C"

// C-ABI doors onto the caching allocator (for hosts that want their table buffers to come from the same pool)
extern "C" int cyl_pool_alloc(void **ptr, int64_t bytes, void *stream) {
  if (bytes < 0) return cylb::fail(CYL_Invalid, "cyl_pool_alloc: negative size");
  CYL_CUDA(cylb::pool_malloc(ptr, (size_t)bytes, (cudaStream_t)stream));
  return CYL_OK;
}
extern "C" int cyl_pool_free(void *ptr, void *stream) {
  CYL_CUDA(cylb::pool_free(ptr, (cudaStream_t)stream));
  return CYL_OK;
}
extern "C" int cyl_pool_trim(void) {
  cylb::pool_trim();
  return CYL_OK;
}
