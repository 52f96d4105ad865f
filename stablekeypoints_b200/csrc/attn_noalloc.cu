// skp_attn_map_fwd_static: the fused attention-map forward for callers that cannot allocate (code under CUDA graph
// capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.
//
// skp_attn_map_fwd stages its tcgen05 operands (bf16 hi + lo splits of q / k, the native-resolution logits s_low and
// the interpolation tables W: attn_tc_lr.cu) in a workspace the CALLER allocates.  This variant takes no workspace:
// the operands go into the file-scope device array below, SKP_STATIC_WS_BYTES (96 MiB) of it, which every GPU that
// loads the library holds until the process exits.  96 MiB covers one BASELINE-size image (SD-2.1 capture set,
// T = 500, R = 128: 83.6 MiB); larger problems are refused with SKP_ERR_WORKSPACE.
//
// In host code `g_static_ws` is only a host-side stand-in for the symbol, NOT a device address: the address is looked
// up once per device with cudaGetSymbolAddress (that device current) and that pointer is what the kernels get.
// The entry point calls no cudaMalloc / cudaFree / allocating library routine.  All calls on a device share the array
// and it keeps the previous call's contents: the forward's staging kernels overwrite every byte the main kernel
// reads (and TMA zero-fills out-of-range rows), so nothing has to start at zero and nothing is cleared; but two calls
// that overlap on the GPU (different streams, or graph branches without a dependency) race on it.
#include <mutex>

#include "common.cuh"

namespace skp {
constexpr size_t kStaticWsBytes = static_cast<size_t>(SKP_STATIC_WS_BYTES);
static __device__ __align__(256) unsigned char g_static_ws[kStaticWsBytes];

constexpr int kMaxDevices = 64;
static void* g_static_ws_addr[kMaxDevices];   // device address of g_static_ws per device ordinal, nullptr = not looked up
static std::mutex g_static_ws_mutex;

// device address of the scratch array on the CURRENT device (looked up on the first call per device)
static int static_ws_address(void** out) {
  int dev = -1;
  SKP_CUDA(cudaGetDevice(&dev));
  SKP_REQUIRE(dev >= 0 && dev < kMaxDevices, "skp_attn_map_fwd_static: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_static_ws_mutex);
  if (!g_static_ws_addr[dev]) {
    void* p = nullptr;
    SKP_CUDA(cudaGetSymbolAddress(&p, g_static_ws));
    SKP_REQUIRE(p != nullptr, "skp_attn_map_fwd_static: cudaGetSymbolAddress returned a null address");
    g_static_ws_addr[dev] = p;
  }
  *out = g_static_ws_addr[dev];
  return SKP_OK;
}
}  // namespace skp

using namespace skp;

extern "C" int skp_attn_map_fwd_static(const skp_attn_layer* h_layers, int n_layers, int B, int T, int R,
                                       const int64_t* token_idx, int n_idx, int64_t idx_stride, float* maps,
                                       float* row_stats, int precision, void* stream) {
  SKP_REQUIRE(h_layers && n_layers >= 1, "skp_attn_map_fwd_static: no layers");
  const size_t need = skp_attn_map_workspace_bytes_for(h_layers, n_layers, B, T, R, precision, SKP_WS_FWD);
  if (need > kStaticWsBytes) {
    set_error("skp_attn_map_fwd_static: this problem needs %zu B of scratch, the static array holds %zu B "
              "(use skp_attn_map_fwd with a caller-allocated workspace)", need, kStaticWsBytes);
    return SKP_ERR_WORKSPACE;
  }
  void* ws = nullptr;
  const int rc = static_ws_address(&ws);
  if (rc != SKP_OK) return rc;
  return skp_attn_map_fwd(h_layers, n_layers, B, T, R, token_idx, n_idx, idx_stride, maps, row_stats, precision, ws,
                          kStaticWsBytes, stream);
}
