// Status / introspection entry points of the C ABI (include/s2t_b200.h).
#include "common.cuh"

#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>

#include "../../include/s2t_b200.h"

namespace s2t {
static thread_local char g_err[1024] = "";

void set_last_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

static int g_pdl = -1;   // -1: read S2T_PDL on first use
bool pdl_enabled() {
  if (g_pdl < 0) {
    const char* e = getenv("S2T_PDL");
    g_pdl = (e != nullptr && e[0] == '0') ? 0 : 1;
  }
  return g_pdl != 0;
}
}  // namespace s2t

extern "C" int s2t_set_pdl(int on) {
  const int prev = s2t::pdl_enabled() ? 1 : 0;
  s2t::g_pdl = on ? 1 : 0;
  return prev;
}

extern "C" const char* s2t_last_error(void) { return s2t::g_err; }

extern "C" int s2t_abi_version(void) { return S2T_ABI_VERSION; }

extern "C" int s2t_device_arch(void) {
  int dev = 0, major = 0, minor = 0;
  S2T_CHECK_CUDA(cudaGetDevice(&dev));
  S2T_CHECK_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
  S2T_CHECK_CUDA(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev));
  return major * 10 + minor;
}
