// Host-side helpers shared by all translation units of libsinga_b200.
#pragma once
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>

namespace sg {

void set_error(const char* fmt, ...);

inline int check_cuda(cudaError_t e, const char* what) {
  if (e == cudaSuccess) return 0;
  set_error("%s: %s", what, cudaGetErrorString(e));
  return 1;
}

inline int num_sms() {
  static thread_local int cached_dev = -1;
  static thread_local int cached = 0;
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev != cached_dev) {
    cudaDeviceGetAttribute(&cached, cudaDevAttrMultiProcessorCount, dev);
    cached_dev = dev;
  }
  return cached > 0 ? cached : 148;
}

}  // namespace sg

#define SG_CHECK_LAUNCH(name)                                   \
  do {                                                          \
    cudaError_t e__ = cudaGetLastError();                       \
    if (e__ != cudaSuccess) {                                   \
      sg::set_error("%s launch: %s", name, cudaGetErrorString(e__)); \
      return 1;                                                 \
    }                                                           \
  } while (0)

#define SG_REQUIRE(cond, ...)      \
  do {                             \
    if (!(cond)) {                 \
      sg::set_error(__VA_ARGS__);  \
      return 2;                    \
    }                              \
  } while (0)
