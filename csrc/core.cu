// Error reporting, version and capability probes for libsinga_b200.
#include <string.h>

#include "../include/singa_b200.h"
#include "common.h"

namespace sg {
static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
}  // namespace sg

extern "C" const char* sg_last_error(void) { return sg::g_err; }
extern "C" int sg_version(void) { return 100; }
extern "C" int sg_device_cc(int device) {
  int major = 0, minor = 0;
  if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device) != cudaSuccess ||
      cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, device) != cudaSuccess) {
    sg::set_error("sg_device_cc: no CUDA device %d", device);
    cudaGetLastError();
    return -1;
  }
  return major * 10 + minor;
}
