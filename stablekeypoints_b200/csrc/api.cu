// Library-level entry points: version, error string, device probe.
#include "common.cuh"

namespace skp {
static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
}  // namespace skp

extern "C" int skp_version(void) { return 101; }
extern "C" const char* skp_last_error(void) { return skp::g_err; }
extern "C" int skp_device_arch(int device) {
  cudaDeviceProp p;
  SKP_CUDA(cudaGetDeviceProperties(&p, device));
  return p.major * 10 + p.minor;
}
