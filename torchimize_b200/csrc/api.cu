// api.cu — housekeeping entry points of the C ABI (include/nlls_b200.h).
#include <cstdarg>
#include "common.cuh"

namespace nlls {
static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
}  // namespace nlls

extern "C" int nlls_abi_version(void) { return NLLS_B200_ABI_VERSION; }
extern "C" const char* nlls_last_error(void) { return nlls::g_err; }

extern "C" int nlls_device_check(void) {
  int dev = 0, major = 0, minor = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e == cudaSuccess) e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
  if (e == cudaSuccess) e = cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev);
  if (e != cudaSuccess) {
    nlls::set_error("nlls_device_check: %s", cudaGetErrorString(e));
    return (int)e;
  }
  if (major != 10) {
    nlls::set_error("built for sm_100a only; current device is sm_%d%d", major, minor);
    return NLLS_E_ARCH;
  }
  return 0;
}
