// Shared helpers for libwoft_b200.so (host side of the C ABI + a few device utilities).
#pragma once
#include <cuda_runtime.h>
#include <cstdarg>
#include <cstdio>
#include <cstdint>

#include "../../include/woft_b200.h"

namespace woft {

// thread-local message behind woft_last_error().  The only other mutable library state: the per-device address of
// the built-in scratch of woft_corr_pyramid_static (corr_static.cu) and the debug trace pointer of corr_tc.cu.
char* last_error_buffer();
int fail(int code, const char* fmt, ...);

#define WOFT_CUDA_CHECK(expr)                                                                         \
  do {                                                                                                \
    cudaError_t err__ = (expr);                                                                       \
    if (err__ != cudaSuccess)                                                                         \
      return ::woft::fail(WOFT_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(err__),  \
                          __FILE__, __LINE__);                                                        \
  } while (0)

#define WOFT_REQUIRE(cond, ...)                                             \
  do {                                                                      \
    if (!(cond)) return ::woft::fail(WOFT_ERR_INVALID_ARGUMENT, __VA_ARGS__); \
  } while (0)

inline int launch_status(const char* what) {
  cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) return fail(WOFT_ERR_CUDA, "%s launch failed: %s", what, cudaGetErrorString(err));
  return WOFT_OK;
}

constexpr int kMaxLevels = 6;

struct LevelPtrs {
  float* ptr[kMaxLevels];
  int h[kMaxLevels];
  int w[kMaxLevels];
};

__host__ __device__ inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

// Programmatic dependent launch for the kernels that follow one another inside a refinement iteration: launched through
// launch_pdl a kernel may be scheduled while its predecessor in the stream drains; it calls pdl_sync() before it reads
// anything the predecessor wrote (or writes anything it read).  WOFT_PDL=0 (read once per process) switches it off.
bool pdl_enabled();

#ifdef __CUDACC__
__device__ __forceinline__ void pdl_sync() {
  asm volatile("griddepcontrol.wait;" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              Args... args) {
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
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
#endif

}  // namespace woft
