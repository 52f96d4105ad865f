// C-ABI dispatch for the attention-map forward / backward (include/skp_b200.h section (1)).
#include "common.cuh"

namespace skp {
int attn_map_fwd_simt(const skp_attn_layer*, int, int, int, int, const int64_t*, int, int64_t, float*, float*, cudaStream_t);
int attn_map_bwd_simt(const skp_attn_layer*, int, int, int, int, const int64_t*, int, int64_t, const float*, void*, size_t, cudaStream_t);
size_t attn_map_bwd_simt_workspace(const skp_attn_layer*, int, int, int);
// tcgen05 path (attn_tc.cu)
bool attn_map_tc_supported(const skp_attn_layer*, int, int, int, int, int n_idx);
size_t attn_map_tc_workspace(const skp_attn_layer*, int, int, int, int, int precision);
int attn_map_fwd_tc(const skp_attn_layer*, int, int, int, int, const int64_t*, int, int64_t, float*, float*, int precision, void*,
                    size_t, cudaStream_t);
// tcgen05 backward (attn_tc_bwd.cu)
bool attn_map_bwd_tc_supported(const skp_attn_layer*, int, int, int, int, int n_idx);
size_t attn_map_bwd_tc_workspace(const skp_attn_layer*, int, int, int, int, int precision);
int attn_map_bwd_tc(const skp_attn_layer*, int, int, int, int, const int64_t*, int, int64_t, const float*, const float*, int precision,
                    void*, size_t, cudaStream_t, const void* fwd_workspace, size_t fwd_workspace_bytes);
}  // namespace skp

using namespace skp;

static int resolve_precision(const skp_attn_layer* h_layers, int n_layers, int B, int T, int R, int n_idx, int precision) {
  if (precision == SKP_PREC_AUTO)
    return attn_map_tc_supported(h_layers, n_layers, B, T, R, n_idx) ? SKP_PREC_TC_BF16X3 : SKP_PREC_SIMT_F32;
  return precision;
}

extern "C" size_t skp_attn_map_workspace_bytes_for(const skp_attn_layer* h_layers, int n_layers, int B, int T, int R,
                                                   int precision, int call) {
  if (!h_layers || n_layers < 1) return 0;
  const int p = resolve_precision(h_layers, n_layers, B, T, R, 0, precision);
  const bool tc = p != SKP_PREC_SIMT_F32;
  const size_t fwd = tc ? attn_map_tc_workspace(h_layers, n_layers, B, T, R, p) : 0;
  // skp_attn_map_bwd dispatch: tcgen05 with row_stats when the shape is supported, else the CUDA-core kernel
  const bool bwd_tc = tc && attn_map_bwd_tc_supported(h_layers, n_layers, B, T, R, 1);
  const size_t bwd_stats = bwd_tc ? attn_map_bwd_tc_workspace(h_layers, n_layers, B, T, R, p) : attn_map_bwd_simt_workspace(h_layers, n_layers, B, R);
  const size_t bwd_nostats = attn_map_bwd_simt_workspace(h_layers, n_layers, B, R);
  switch (call) {
    case SKP_WS_FWD: return fwd;
    case SKP_WS_BWD_STATS: return bwd_stats;
    case SKP_WS_BWD_NOSTATS: return bwd_nostats;
    default: {
      size_t m = fwd > bwd_stats ? fwd : bwd_stats;
      return m > bwd_nostats ? m : bwd_nostats;
    }
  }
}
extern "C" size_t skp_attn_map_workspace_bytes(const skp_attn_layer* h_layers, int n_layers, int B, int T, int R, int precision) {
  return skp_attn_map_workspace_bytes_for(h_layers, n_layers, B, T, R, precision, SKP_WS_ANY);
}

extern "C" int skp_attn_map_fwd(const skp_attn_layer* h_layers, int n_layers, int B, int T, int R, const int64_t* token_idx,
                                int n_idx, int64_t idx_stride, float* maps, float* row_stats, int precision, void* workspace,
                                size_t workspace_bytes, void* stream) {
  SKP_REQUIRE(maps, "skp_attn_map_fwd: maps is null");
  SKP_REQUIRE(!token_idx || n_idx >= 1, "skp_attn_map_fwd: token_idx given but n_idx < 1");
  int p = resolve_precision(h_layers, n_layers, B, T, R, token_idx ? n_idx : 0, precision);
  if (p == SKP_PREC_SIMT_F32)
    return attn_map_fwd_simt(h_layers, n_layers, B, T, R, token_idx, n_idx, idx_stride, maps, row_stats, as_stream(stream));
  if (p == SKP_PREC_TC_BF16 || p == SKP_PREC_TC_BF16X3) {
    if (!attn_map_tc_supported(h_layers, n_layers, B, T, R, token_idx ? n_idx : 0)) {
      set_error("skp_attn_map_fwd: shape not supported by the tcgen05 kernel (need dim_head %% 16 == 0, T <= 512, R*R %% 128 == 0)");
      return SKP_ERR_UNSUPPORTED;
    }
    return attn_map_fwd_tc(h_layers, n_layers, B, T, R, token_idx, n_idx, idx_stride, maps, row_stats, p, workspace, workspace_bytes,
                           as_stream(stream));
  }
  set_error("skp_attn_map_fwd: unknown precision %d", precision);
  return SKP_ERR_INVALID;
}

extern "C" int skp_attn_map_bwd_reuse(const skp_attn_layer* h_layers, int n_layers, int B, int T, int R, const int64_t* token_idx,
                                      int n_idx, int64_t idx_stride, const float* dmaps_sel, const float* row_stats, int precision,
                                      void* workspace, size_t workspace_bytes, const void* fwd_workspace,
                                      size_t fwd_workspace_bytes, void* stream) {
  int p = resolve_precision(h_layers, n_layers, B, T, R, n_idx, precision);
  if (p != SKP_PREC_SIMT_F32 && row_stats && attn_map_bwd_tc_supported(h_layers, n_layers, B, T, R, n_idx))
    return attn_map_bwd_tc(h_layers, n_layers, B, T, R, token_idx, n_idx, idx_stride, dmaps_sel, row_stats, p, workspace,
                           workspace_bytes, as_stream(stream), fwd_workspace, fwd_workspace_bytes);
  return attn_map_bwd_simt(h_layers, n_layers, B, T, R, token_idx, n_idx, idx_stride, dmaps_sel, workspace, workspace_bytes, as_stream(stream));
}

extern "C" int skp_attn_map_bwd(const skp_attn_layer* h_layers, int n_layers, int B, int T, int R, const int64_t* token_idx,
                                int n_idx, int64_t idx_stride, const float* dmaps_sel, const float* row_stats, int precision,
                                void* workspace, size_t workspace_bytes, void* stream) {
  return skp_attn_map_bwd_reuse(h_layers, n_layers, B, T, R, token_idx, n_idx, idx_stride, dmaps_sel, row_stats, precision, workspace,
                                workspace_bytes, nullptr, 0, stream);
}
