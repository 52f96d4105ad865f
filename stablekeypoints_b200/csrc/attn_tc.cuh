// Declarations shared by the tcgen05 attention-map kernels (attn_tc.cu forward, attn_tc_bwd.cu backward).
#pragma once

#include <cuda.h>
#include <cuda_bf16.h>
#include <cstring>

#include "common.cuh"
#include "ptx_sm100.cuh"

// 1: the backward's window flush accumulates dz in bf16 (TMA reduce-add on a bf16 tensor map); the native-resolution GEMMs then
// read it as is (dq: MN-major A) and lr_dz_convert_kernel is not launched.  0: fp32 accumulation + conversion.
#ifndef SKP_DZ_BF16
#define SKP_DZ_BF16 1
#endif

namespace skp {

constexpr int TC_MAX_LAYERS = 12;
constexpr int TC_BM = 128;       // pixels per tile (UMMA M)
constexpr int TC_BN = 256;       // tokens per CTA (UMMA N)
constexpr int TC_BK = 64;        // bf16 per k-block = one 128-byte swizzle row
constexpr int TC_THREADS = 320;  // warp 0 TMA, warp 1 MMA, warps 2..9 softmax
constexpr int TC_SOFTMAX_THREADS = 256;
constexpr int TC_MAX_SEL = 32;   // selected-token planes per forward launch
constexpr int TC_A_BYTES = TC_BM * TC_BK * 2;   // 16 KB
constexpr int TC_B_BYTES = TC_BN * TC_BK * 2;   // 32 KB

// bf16 operands of every captured layer, staged in the workspace by tc_prepare():
//   q' = bicubic(q) * scale*log2e as [b*h + head][pixel][d] hi (+ lo residual), k as [bk*h + head][token][d]
struct TcOperands {
  CUtensorMap q_map[TC_MAX_LAYERS][2];   // [layer][hi, lo]: dims (d, R*R, B*h), box (64, 128, 1)
  CUtensorMap k_map[TC_MAX_LAYERS][2];   // [layer][hi, lo]: dims (d, T,   Bk*h), box (64, 256, 1)
  __nv_bfloat16* k_ptr[TC_MAX_LAYERS][2];
  __nv_bfloat16* s_ptr[TC_MAX_LAYERS][2];   // lr: native-resolution logits hi, lo [b*h + head][token][s*s]
  int heads[TC_MAX_LAYERS], d[TC_MAX_LAYERS], k_shared[TC_MAX_LAYERS], head_base[TC_MAX_LAYERS];
  float weight[TC_MAX_LAYERS];           // layer weight / heads
  float scale[TC_MAX_LAYERS];
  int n_layers, total_heads;
  // "upsample after the logits" mode (attn_tc_lr.cu): q_map = interpolation tables W [tile][128 px][d[l] taps],
  // k_map = native-resolution logits s_low [b*h + head][token][s*s]; the B tile of (tile, kb) starts at source pixel
  // lr_window_row0(tile) * s + 64 kb
  int lr, R;
  int win_s[TC_MAX_LAYERS], win_nr[TC_MAX_LAYERS];
  // lr: order-preserving integer keys of (max s_low, max -s_low) over all tokens and source pixels of every (image, head):
  // [layer][(b * heads + head) * 2 + {0, 1}] (lr_range_key / lr_range_value).  The range bounds the logits of that (image,
  // head) at every feature pixel -- bicubic weights (A = -0.75) sum to 1 with at most 0.445 of negative mass over both axes,
  // so S lies within [min - 0.445 range, max + 0.445 range] -- which lets the forward skip the row-max pass when exp2 of
  // the logits shifted by the range's centre cannot over/underflow.
  const int* absmax[TC_MAX_LAYERS];
};

// float <-> int key whose signed-integer order is the float order (atomicMax on floats of either sign)
__device__ __host__ __forceinline__ int lr_range_key(float v) {
#ifdef __CUDA_ARCH__
  const int b = __float_as_int(v);
#else
  int b; memcpy(&b, &v, 4);
#endif
  return b >= 0 ? b : b ^ 0x7fffffff;
}
__device__ __forceinline__ float lr_range_value(int key) { return __int_as_float(key >= 0 ? key : key ^ 0x7fffffff); }

// window of source rows [r0, r0 + nr) a 128-pixel tile of the R x R output interpolates from (s x s source)
struct LrGeom {
  int nr, kw;   // rows in the window; taps per tile = nr * s rounded up to a multiple of 64
};
__device__ __host__ __forceinline__ int lr_window_row0(int tile, int s, int R, int nr) {
  int ia;
  float w[4];
  bicubic_coeffs(tile * TC_BM / R, static_cast<float>(s) / R, &ia, w);
  const int lo = ia - 1 > 0 ? ia - 1 : 0;
  return lo < s - nr ? lo : s - nr;
}
LrGeom lr_geometry(int s, int R);
bool attn_map_lr_supported(const skp_attn_layer* h_layers, int n_layers, int B, int T, int R, int n_idx);
size_t attn_map_lr_workspace(const skp_attn_layer* h_layers, int n_layers, int B, int T, int R);
// operands of one layer as tc_prepare_lr stages them in a workspace
struct LrLayerBufs {
  __nv_bfloat16 *q_hi, *q_lo, *k_hi, *k_lo, *s_hi, *s_lo, *w_hi, *w_lo;
  int* range;
};
// carve-up of an "upsample after the logits" operand workspace over `n_layers` layers (no launches): the layout
// tc_prepare_lr writes, so that a later call can find the operands an earlier call staged.  -> first free byte
char* lr_carve(const skp_attn_layer* h_layers, int n_layers, int B, int T, int R, void* workspace, LrLayerBufs* out);
// bn: token rows per TMA box of the logit window (k_map).  ready (optional, n_layers entries): operands ALREADY staged for
// these layers (by the forward of the same q / k): nothing is launched, no workspace is consumed, only the maps are encoded.
int tc_prepare_lr(const skp_attn_layer* h_layers, int n_layers, int B, int T, int R, void* workspace, cudaStream_t st,
                  TcOperands* out, char** ws_end, int bn = TC_BN, const LrLayerBufs* ready = nullptr);

bool attn_map_tc_supported(const skp_attn_layer* h_layers, int n_layers, int B, int T, int R, int n_idx);
size_t attn_map_tc_workspace(const skp_attn_layer* h_layers, int n_layers, int B, int T, int R, int precision);
size_t attn_map_staged_workspace(const skp_attn_layer* h_layers, int n_layers, int B, int T, int R, int precision);
// runs the prep kernels on `st`, encodes the tensor maps; *ws_end = first free byte of the workspace
int tc_prepare(const skp_attn_layer* h_layers, int n_layers, int B, int T, int R, bool x3, void* workspace, cudaStream_t st,
               TcOperands* out, char** ws_end);
int tc_encode_map(CUtensorMap* m, void* ptr, int d, int rows, int batch, int box_rows);
int tc_encode_map_f32(CUtensorMap* m, void* ptr, int inner, int rows, int batch, int box_rows);

inline size_t tc_align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

}  // namespace skp
