// util.hpp -- internal helpers of the C++ host layer (not installed).
#pragma once
#include <cuda_runtime.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <string>

#include "cylon/status.hpp"
#include "cylon_b200.h"
#include "../csrc/common.cuh"

namespace cylon {
namespace detail {

inline Status FromCode(int rc) { return rc == CYL_OK ? Status::OK() : Status(rc, cyl_last_error()); }

#define CYL_RETURN_IF_FAILED(expr)                              \
  do {                                                          \
    int _rc = (expr);                                           \
    if (_rc != CYL_OK) return ::cylon::Status(_rc, cyl_last_error()); \
  } while (0)

#define CUDA_RETURN_IF_FAILED(expr)                                                                                  \
  do {                                                                                                               \
    cudaError_t _e = (expr);                                                                                         \
    if (_e != cudaSuccess)                                                                                           \
      return ::cylon::Status(_e == cudaErrorMemoryAllocation ? ::cylon::Code::OutOfMemory : ::cylon::Code::ExecutionError, \
                             std::string("CUDA error: ") + cudaGetErrorString(_e) + " in " #expr);                   \
  } while (0)

// stream-ordered device buffer with RAII
struct DevBuf {
  void *p = nullptr;
  cudaStream_t s = nullptr;
  DevBuf() = default;
  DevBuf(const DevBuf &) = delete;
  DevBuf &operator=(const DevBuf &) = delete;
  ~DevBuf() {
    if (p) cylb::pool_free(p, s);
  }
  Status alloc(size_t bytes, void *stream) {
    if (p) cylb::pool_free(p, s);
    p = nullptr;
    s = static_cast<cudaStream_t>(stream);
    CUDA_RETURN_IF_FAILED(cylb::pool_malloc(&p, bytes ? bytes : 16, s));
    return Status::OK();
  }
  template <typename T>
  T *as() const { return static_cast<T *>(p); }
};

// Stage tracing for humans (the reference logs per-stage milliseconds with glog, e.g. table.cpp:170-176): with
// CYLON_B200_TRACE=1 every stage boundary synchronises the stream and prints the wall time since the previous one.
struct StageTrace {
  bool on;
  cudaStream_t s;
  std::chrono::steady_clock::time_point t;
  const char *what;
  StageTrace(const char *op, void *stream) : s(static_cast<cudaStream_t>(stream)), what(op) {
    const char *e = getenv("CYLON_B200_TRACE");
    on = e && e[0] == '1';
    if (on) {
      cudaStreamSynchronize(s);
      t = std::chrono::steady_clock::now();
    }
  }
  void mark(const char *stage) {
    if (!on) return;
    cudaStreamSynchronize(s);
    auto now = std::chrono::steady_clock::now();
    fprintf(stderr, "[cylon_b200] %s: %s %.3f ms\n", what, stage, std::chrono::duration<double, std::milli>(now - t).count());
    t = now;
  }
};

inline size_t validity_bytes(int64_t rows) { return (size_t)((rows + 31) / 32) * 4 + 4; }

}  // namespace detail
}  // namespace cylon
