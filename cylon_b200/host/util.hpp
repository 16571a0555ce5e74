// util.hpp -- internal helpers of the C++ host layer (not installed).
#pragma once
#include <cuda_runtime.h>

#include <cstdio>
#include <string>

#include "cylon/status.hpp"
#include "cylon_b200.h"

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
    if (p) cudaFreeAsync(p, s);
  }
  Status alloc(size_t bytes, void *stream) {
    if (p) cudaFreeAsync(p, s);
    p = nullptr;
    s = static_cast<cudaStream_t>(stream);
    CUDA_RETURN_IF_FAILED(cudaMallocAsync(&p, bytes ? bytes : 16, s));
    return Status::OK();
  }
  template <typename T>
  T *as() const { return static_cast<T *>(p); }
};

inline size_t validity_bytes(int64_t rows) { return (size_t)((rows + 31) / 32) * 4 + 4; }

}  // namespace detail
}  // namespace cylon
