// Library-level entry points: version, error string, device probe.
#include <atomic>

#include "common.cuh"

namespace skp {
static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
static std::atomic<int> g_sm_count[64];     // 0 = not queried yet
int num_sms() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  int n = g_sm_count[dev].load(std::memory_order_relaxed);
  if (n == 0) {
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    g_sm_count[dev].store(n, std::memory_order_relaxed);
  }
  return n;
}
static std::atomic<long long> g_launches{0};
int note_launch() {
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return 0;
}
}  // namespace skp

extern "C" int skp_version(void) { return 107; }
extern "C" long long skp_kernel_launches(void) { return skp::g_launches.load(std::memory_order_relaxed); }
extern "C" const char* skp_last_error(void) { return skp::g_err; }
extern "C" int skp_device_arch(int device) {
  cudaDeviceProp p;
  SKP_CUDA(cudaGetDeviceProperties(&p, device));
  return p.major * 10 + p.minor;
}
