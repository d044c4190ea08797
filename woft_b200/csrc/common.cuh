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

}  // namespace woft
