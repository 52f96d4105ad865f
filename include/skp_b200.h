/*
 * skp_b200.h — C ABI of the B200-native StableKeypoints hot path (libskp_b200.so).
 *
 * The reference (damaggu/StableKeypoints) is pure Python on PyTorch and has no FFI; the
 * "interface each entry point replaces" is therefore the Python function (file:line under
 * /root/reference/unsupervised_keypoints/) whose ATen op sequence the entry point fuses.
 * The Python host mirror in stablekeypoints_b200/ binds these symbols with ctypes
 * (INTEGRATION.md shows the stub a maintainer of the reference would add).
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless its name starts with h_ (host);
 *   - tensors are dense row-major fp32 unless stated; indices are int64;
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream);
 *   - all calls are asynchronous on `stream`, never synchronise, never allocate;
 *   - return value: 0 on success, negative skp_status on error; skp_last_error() gives the
 *     message of the last failure on the calling thread.  There is no CPU fallback.
 */
#ifndef SKP_B200_H
#define SKP_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  SKP_OK = 0,
  SKP_ERR_INVALID = -1,     /* bad argument (shape, null pointer, unsupported size) */
  SKP_ERR_CUDA = -2,        /* CUDA runtime / driver error, see skp_last_error()    */
  SKP_ERR_UNSUPPORTED = -3, /* shape not supported by the requested kernel          */
  SKP_ERR_WORKSPACE = -4    /* workspace too small                                  */
} skp_status;

/* Arithmetic used for the Q.K^T contraction of the attention-map kernels. */
typedef enum {
  SKP_PREC_AUTO = 0,    /* tcgen05 bf16x3 when the shape allows it, else SIMT fp32            */
  SKP_PREC_SIMT_F32 = 1,/* CUDA-core fp32 FMA (validation kernel, any shape)                  */
  SKP_PREC_TC_BF16 = 2, /* tcgen05 kind::f16, bf16 operands, fp32 accumulate (fast mode)      */
  SKP_PREC_TC_BF16X3 = 3/* tcgen05, operands split hi+lo bf16, 3 products (~fp32, parity mode) */
} skp_precision;

int skp_version(void);
const char* skp_last_error(void);
/* compute capability major*10+minor of `device`, or negative status */
int skp_device_arch(int device);
/* number of CUDA kernels this library has launched in this process so far (diagnostics:
 * bench.py reports the difference over its timed region as gpu_launches) */
long long skp_kernel_launches(void);

/* ---------------------------------------------------------------------------------------
 * (1) Fused cross-attention map.
 * Replaces: ptp_utils.py:508-538 (captured branch of the patched CrossAttention.forward:
 * bicubic upsample to R, to_q, q.k^T*scale, softmax over tokens) + optimize.py:27-79
 * (collect_maps: token gather, head/layer mean; the bilinear upsample is skp_bilinear_upsample
 * or is folded into the consumers skp_tta_accumulate / skp_plane_*).
 *
 * One captured layer.  q is the module's own low-resolution projection to_q(x)
 * (ptp_utils.py:483), k = to_k(context) (ptp_utils.py:487); the kernel upsamples q
 * bicubically (align_corners=False, A=-0.75) to R x R, which equals to_q(upsample(x)) because
 * to_q is linear and bias-free.  s == R feeds an already-upsampled q.
 * ------------------------------------------------------------------------------------- */
typedef struct {
  const float* q;   /* (B, s*s, C)                                                       */
  const float* k;   /* (B, T, C) or (1, T, C) when k_shared != 0                         */
  float* dq;        /* backward only: (B, s*s, C), ACCUMULATED into (caller zero-fills)  */
  float* dk;        /* backward only: same shape as k, ACCUMULATED into                  */
  int32_t s;        /* source side of q; R % s == 0                                      */
  int32_t C;        /* heads * dim_head                                                  */
  int32_t heads;
  int32_t k_shared; /* 1: one k for all images (context.repeat(B,1,1), ptp_utils.py:229) */
  float scale;      /* dim_head ** -0.5 (ptp_utils.py:534)                               */
  float weight;     /* weight of this layer's head-mean in the layer mean (1/L)          */
} skp_attn_layer;

/* bytes of scratch the forward/backward need for this problem (0 for SIMT) */
size_t skp_attn_map_workspace_bytes(const skp_attn_layer* h_layers, int n_layers, int B, int T, int R,
                                    int precision);
/* The same per call: the query above covers every call the caller might make with one buffer, including the
 * fp32 CUDA-core backward (row_stats == NULL), whose upsampled-gradient scratch is the largest term by far.
 * `call` = SKP_WS_FWD (skp_attn_map_fwd), SKP_WS_BWD_STATS (skp_attn_map_bwd with row_stats),
 * SKP_WS_BWD_NOSTATS (skp_attn_map_bwd with row_stats == NULL), SKP_WS_ANY (= the query above). */
enum { SKP_WS_ANY = 0, SKP_WS_FWD = 1, SKP_WS_BWD_STATS = 2, SKP_WS_BWD_NOSTATS = 3 };
size_t skp_attn_map_workspace_bytes_for(const skp_attn_layer* h_layers, int n_layers, int B, int T, int R,
                                        int precision, int call);

/* maps (B, T', R, R): T' = n_idx when token_idx != NULL (gather, optimize.py:58-59) else T.
 * token_idx holds n_idx ids for image b at token_idx + b*idx_stride (idx_stride 0: the same
 * ids for every image, the reference's case; n_idx: one row per image, used when the images
 * of a batch carry their own selected tokens, optimize.py:403-424).
 * out[b,t,p] = sum_l weight_l * mean_head softmax_t(scale_l * q'_l[b,head,p,:].k_l[b,head,t,:]).
 * row_stats (optional, may be NULL): (B, sum_l heads_l, R*R, 2) = (row max, sum exp) in log2
 * units, consumed by skp_attn_map_bwd. */
int skp_attn_map_fwd(const skp_attn_layer* h_layers, int n_layers, int B, int T, int R,
                     const int64_t* token_idx, int n_idx, int64_t idx_stride, float* maps, float* row_stats,
                     int precision, void* workspace, size_t workspace_bytes, void* stream);

/* skp_attn_map_fwd without a caller workspace and without any allocation.
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot
 * call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  The tcgen05
 * operands are staged in a file-scope `static __device__` array of SKP_STATIC_WS_BYTES inside the library
 * (held by every GPU that loads it until the process exits).  Same arguments and results as
 * skp_attn_map_fwd minus (workspace, workspace_bytes); problems whose forward workspace
 * (skp_attn_map_workspace_bytes_for(..., SKP_WS_FWD)) exceeds SKP_STATIC_WS_BYTES return
 * SKP_ERR_WORKSPACE (96 MiB holds one BASELINE-size image: SD-2.1 capture set, T=500, R=128 = 83.6 MiB).
 * Calls cudaGetSymbolAddress once per device (first call on that device, that device current) and
 * otherwise no cudaMalloc/cudaFree/allocating routine.  All calls on a device share the array: calls
 * that overlap on the GPU (different streams without a dependency) race; nothing in it has to start at
 * zero (the staging kernels overwrite everything the main kernel reads). */
#define SKP_STATIC_WS_BYTES (96u << 20)
int skp_attn_map_fwd_static(const skp_attn_layer* h_layers, int n_layers, int B, int T, int R,
                            const int64_t* token_idx, int n_idx, int64_t idx_stride, float* maps, float* row_stats,
                            int precision, void* stream);

/* Backward of the above for a gradient that is non-zero on n_idx token planes only
 * (the losses read attn_map[top_k indices], optimize.py:420-424).
 * dmaps_sel (B, n_idx, R, R) = d loss / d maps[b, token_idx[j]].  Accumulates into
 * layer.dq / layer.dk.  Linear in dmaps_sel, so callers may chunk planes.
 * row_stats: the (B, sum_l heads_l, R*R, 2) tensor skp_attn_map_fwd wrote for the SAME q/k, or
 * NULL.  With it (and dim_head <= 64, n_idx <= 16, T <= 512) the backward runs on tcgen05
 * (attn_tc_bwd.cu); without it the fp32 CUDA-core kernel recomputes the softmax statistics.
 * Arithmetic: the tcgen05 backward recomputes the logits fp32-grade (bf16x3) but takes SINGLE-bf16 operands in its
 * gradient products (dS, the interpolation table / q' / k) with fp32 accumulation, and adds the window sums of the logit
 * gradient in bf16: dq / dk are within 1.5e-2 of each gradient's max of fp64 autograd (measured 3e-3 .. 7e-3;
 * tests/test_full_size_parity_gpu.py), looser than the forward's 1e-3.  SKP_PREC_SIMT_F32 is the fp32 backward (2e-3). */
int skp_attn_map_bwd(const skp_attn_layer* h_layers, int n_layers, int B, int T, int R,
                     const int64_t* token_idx, int n_idx, int64_t idx_stride, const float* dmaps_sel,
                     const float* row_stats, int precision, void* workspace, size_t workspace_bytes, void* stream);

/* skp_attn_map_bwd that does not recompute what the forward already staged (ABI 107).  fwd_workspace / fwd_workspace_bytes:
 * the workspace skp_attn_map_fwd was given for the SAME layers, q, k, B, T, R and precision (all tokens or selected ones),
 * still intact.  Layers on the upsample-after-logits path then read their native-resolution logits and interpolation tables
 * from it instead of running the two head splits, the logit GEMM and the table kernel again; other layers, or a forward
 * that staged upsampled q' instead, are unaffected (the pointer is ignored).  NULL == skp_attn_map_bwd. */
int skp_attn_map_bwd_reuse(const skp_attn_layer* h_layers, int n_layers, int B, int T, int R,
                           const int64_t* token_idx, int n_idx, int64_t idx_stride, const float* dmaps_sel,
                           const float* row_stats, int precision, void* workspace, size_t workspace_bytes,
                           const void* fwd_workspace, size_t fwd_workspace_bytes, void* stream);

/* ---------------------------------------------------------------------------------------
 * Resampling (SURVEY.md Appendix A semantics == torch.nn.functional, align_corners=False)
 * ------------------------------------------------------------------------------------- */
/* F.interpolate(bilinear): in (N, h, w) -> out (N, H, W).  Replaces optimize.py:63-70. */
int skp_bilinear_upsample(const float* in, int N, int h, int w, float* out, int H, int W, void* stream);
/* F.interpolate(bicubic) on channels-last data: in (B, s, s, C) -> out (B, R, R, C).
 * Replaces ptp_utils.py:513-529 (exposed for tests / API completeness). */
int skp_bicubic_upsample_nhwc(const float* in, int B, int s, int C, float* out, int R, void* stream);
/* F.grid_sample(img, F.affine_grid(theta)) bilinear/zeros.  img (N,C,H,W), theta (N,2,3)
 * DEVICE fp32.  Replaces invertable_transform.py:64-68 and :86-90. */
int skp_affine_warp(const float* img, const float* theta, int N, int C, int H, int W, float* out, void* stream);
/* adjoint of skp_affine_warp w.r.t. img: dimg (N,C,H,W) is ACCUMULATED into (atomics). */
int skp_affine_warp_bwd(const float* dout, const float* theta, int N, int C, int H, int W, float* dimg, void* stream);

/* ---------------------------------------------------------------------------------------
 * (2) Readout / per-plane statistics.  maps (N, H, W).
 * ------------------------------------------------------------------------------------- */
/* eval.py:39-60 find_max_pixel: first flat argmax per plane.  flat_idx (N) int64; maxval (N)
 * optional.  (row, col) = (idx / W + 0.5, idx % W + 0.5) is formed by the caller or by
 * skp_find_k_max_pixels. */
int skp_plane_argmax(const float* maps, int N, int H, int W, int64_t* flat_idx, float* maxval, void* stream);
/* eval.py:62-111 find_k_max_pixels + mask_radius: `num` successive maxima with a disc of radius
 * 0.05*H zeroed between picks.  out (num, N, 2) fp32 (row+0.5, col+0.5). num <= 8. */
int skp_find_k_max_pixels(const float* maps, int N, int H, int W, int num, float* out, void* stream);
/* eval.py:113-155 pixel_from_weighted_avg (windowed soft-argmax, distance px; <0: no window).
 * out (N,2) fp32 (row, col).  If mutate != 0 the input is zeroed outside the window exactly
 * like the reference's in-place edit (eval.py:137). */
int skp_weighted_avg_readout(float* maps, int N, int H, int W, float distance, int mutate, float* out, void* stream);
/* ptp_utils.py:97-108: KL(gaussian(argmax) || softmax_pixels(map)) per plane -> kl (N).
 * ptp_utils.py:179-182: entropy of softmax over pixels -> entropy (N).  Either may be NULL.
 * argmax_out (N) int64 optional (flat index of the first maximum). */
int skp_plane_stats(const float* maps, int N, int H, int W, float sigma, float epsilon, int num_subjects,
                    float* kl, float* entropy, int64_t* argmax_out, void* stream);
/* optimize_token.py:204-242 gaussian_circles: pos (S, N, 2) in [0,1] (row, col) -> out (N, size, size) */
int skp_gaussian_circles(const float* pos, int S, int N, int size, float sigma, float* out, void* stream);

/* ---------------------------------------------------------------------------------------
 * (3) Token selection
 * ------------------------------------------------------------------------------------- */
/* torch.argsort(score, descending=False)[:top_k] (ptp_utils.py:110-112, :185-187); ties by index.
 * N <= 4096. */
int skp_rank_ascending(const float* score, int N, int top_k, int64_t* out, void* stream);
/* B score vectors in one launch: score (B, N) -> out (B, top_k).  The token optimisation ranks the tokens of every
 * image of the batch (optimize.py:403-418 runs per image; one CTA per image here). */
int skp_rank_ascending_batched(const float* score, int B, int N, int top_k, int64_t* out, void* stream);
/* ptp_utils.py:115-159 furthest_point_sampling on the per-plane argmax locations.
 * argmax_flat (N) int64 from skp_plane_argmax; candidates (n_cand) int64 token ids; out (top_k)
 * int64 in selection order, n_out (1) int32 = how many were written (== top_k unless the
 * candidate list runs out, as in the reference).  fp32 distance arithmetic identical to the
 * reference's tensor ops, so indices are bit-exact. n_cand <= 1024. */
int skp_furthest_point_sampling(const int64_t* argmax_flat, int N, int H, int W, const int64_t* candidates,
                                int n_cand, int top_k, int64_t* out, int32_t* n_out, void* stream);
/* B images in one launch: argmax_flat (B, N), candidates (B, n_cand) -> out (B, top_k), n_out (B). */
int skp_furthest_point_sampling_batched(const int64_t* argmax_flat, int B, int N, int H, int W, const int64_t* candidates,
                                        int n_cand, int top_k, int64_t* out, int32_t* n_out, void* stream);

/* ---------------------------------------------------------------------------------------
 * (4) Losses.  Replaces optimize.py:157-206 (+ invertable_transform.py:72-92 for the un-warp).
 * maps_a / maps_b (T, H, W) are the maps of the image and of its transformed copy;
 * idx (K) int64 the selected tokens; theta_inv (2,3) DEVICE fp32 = inverse of the forward affine.
 * loss_parts (K, 2): per selected plane [sum (a-g)^2, sum (a-unwarp(b))^2]; the caller divides
 * by K*H*W.  Either map pointer pair may be used alone (maps_b == NULL: sharpening only).
 * ------------------------------------------------------------------------------------- */
int skp_keypoint_losses_fwd(const float* maps_a, const float* maps_b, const int64_t* idx, int K, int H, int W,
                            const float* theta_inv, float sigma, int num_subjects, float* loss_parts, void* stream);
/* gradients w.r.t. the SELECTED planes: d_a_sel (K,H,W) written; d_b_sel (K,H,W) accumulated
 * (atomics; caller zero-fills).  g_sharp / g_equiv are d total / d (mean sharpening / mean
 * equivariance) as host scalars. */
int skp_keypoint_losses_bwd(const float* maps_a, const float* maps_b, const int64_t* idx, int K, int H, int W,
                            const float* theta_inv, float sigma, int num_subjects, float g_sharp, float g_equiv,
                            float* d_a_sel, float* d_b_sel, void* stream);

/* The same for the B images of a batch in ONE launch (the loop over images of optimize.py:403-429).
 * maps_a / maps_b (B, planes, H, W); idx (B, K) plane ids inside each image's stack (idx_stride = K) or one shared row
 * (idx_stride = 0); theta_inv (B, 2, 3) DEVICE; loss_parts (B, K, 2).
 * Backward: g (B, 2) DEVICE fp32 = d total / d (mean sharpening, mean equivariance) of every image -- a device pointer, so
 * the autograd glue never synchronises the stream to read upstream scalars; d_a_sel written, d_b_sel accumulated
 * (caller zero-fills), both (B, K, H, W). */
int skp_keypoint_losses_batched_fwd(const float* maps_a, const float* maps_b, const int64_t* idx, int64_t idx_stride, int B,
                                    int planes, int K, int H, int W, const float* theta_inv, float sigma, int num_subjects,
                                    float* loss_parts, void* stream);
int skp_keypoint_losses_batched_bwd(const float* maps_a, const float* maps_b, const int64_t* idx, int64_t idx_stride, int B,
                                    int planes, int K, int H, int W, const float* theta_inv, float sigma, int num_subjects,
                                    const float* g, float* d_a_sel, float* d_b_sel, void* stream);

/* ---------------------------------------------------------------------------------------
 * (5) Test-time-augmentation ensemble.  Replaces eval.py:221-222, 260-266, 327-330 and the
 * bilinear upsample of optimize.py:63-70 in front of it.
 * maps (A, K, h, w): per augmentation the K selected planes at feature resolution (h == H
 * means already upsampled); theta_inv (A,2,3) DEVICE.  sum_ / num_ (K, H, W) are
 * read-modify-written: sum_ += unwarp(upsample(maps)), num_ += unwarp(ones).
 * ------------------------------------------------------------------------------------- */
int skp_tta_accumulate(const float* maps, const float* theta_inv, int A, int K, int h, int w,
                       float* sum_, float* num_, int H, int W, void* stream);
/* out = sum_/num_ with NaN -> 0 (eval.py:327-330) */
int skp_tta_finalize(const float* sum_, const float* num_, int64_t n, float* out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SKP_B200_H */
