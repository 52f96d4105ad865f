// placeholder until the tcgen05 kernel lands
#include "common.cuh"
namespace skp {
bool attn_map_tc_supported(const skp_attn_layer*, int, int, int, int, int) { return false; }
size_t attn_map_tc_workspace(const skp_attn_layer*, int, int, int, int, int) { return 0; }
int attn_map_fwd_tc(const skp_attn_layer*, int, int, int, int, const int64_t*, int, float*, float*, int, void*, size_t, cudaStream_t) {
  set_error("tcgen05 kernel not built");
  return SKP_ERR_UNSUPPORTED;
}
}
