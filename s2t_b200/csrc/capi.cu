// Status / introspection entry points of the C ABI (include/s2t_b200.h).
#include "common.cuh"

#include <stdarg.h>
#include <stdio.h>

#include "../../include/s2t_b200.h"

namespace s2t {
static thread_local char g_err[1024] = "";

void set_last_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
}  // namespace s2t

extern "C" const char* s2t_last_error(void) { return s2t::g_err; }

extern "C" int s2t_abi_version(void) { return S2T_ABI_VERSION; }

extern "C" int s2t_device_arch(void) {
  int dev = 0, major = 0, minor = 0;
  S2T_CHECK_CUDA(cudaGetDevice(&dev));
  S2T_CHECK_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
  S2T_CHECK_CUDA(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev));
  return major * 10 + minor;
}
