// Host-side helpers shared by every translation unit of libb200dp.so:
// thread-local error string and a CUDA error-check macro that records where it failed.
#pragma once

#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>

namespace b200dp {

char* error_buffer();  // thread-local, 1 KiB
constexpr int kErrorBufferSize = 1024;

inline void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(error_buffer(), kErrorBufferSize, fmt, ap);
  va_end(ap);
}

}  // namespace b200dp

#define B200DP_CUDA_TRY(expr)                                                                              \
  do {                                                                                                     \
    cudaError_t _e = (expr);                                                                               \
    if (_e != cudaSuccess) {                                                                               \
      ::b200dp::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e));           \
      return static_cast<int>(_e);                                                                         \
    }                                                                                                      \
  } while (0)
