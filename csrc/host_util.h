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

namespace b200dp {

// Programmatic dependent launch: the kernel may start (and run its prologue: barrier init, TMEM
// allocation, descriptor prefetch) while the previous kernel on the stream is still draining; inside
// the kernel griddepcontrol.wait blocks before the first access to global memory.  Captured into a
// CUDA graph this becomes a programmatic dependency edge.  B200DP_PDL=0 turns it off (A/B runs).
bool pdl_enabled();

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
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
