// "Upsample after the logits" operands for the tcgen05 attention-map kernels (sm_100a).
//
// The reference upsamples the layer input bicubically and THEN projects / multiplies by k
// (ptp_utils.py:513-534).  Bicubic interpolation is linear and to_q has no bias, so the logits of a
// 128-pixel tile are an interpolation of the logits at the module's native resolution:
//
//   S[p, t] = sum_tap W[p, tap] * s_low[t, tap],      s_low[t, sp] = scale*log2e * q[sp, :] . k[t, :]
//
// where the taps are the source pixels of the few source rows the tile touches (a window of NR rows x s
// columns, 64 taps at 16^2 and 128 at 32^2 for R = 128) and W holds the products of the clamped bicubic row /
// column weights.  The main kernels keep their M=128 x N=256 x K MMA shape with A = W (one small table per
// source size, L2 resident) and B = a window of s_low, so the 2.35 GB bf16 staging of the upsampled q' (and
// the prep pass that wrote it) disappears; s_low itself is a 1.7 GFLOP/image GEMM at native resolution.
//
//  * lr_split_heads_kernel : fp32 (Bx, rows, C) -> bf16 hi + lo residual, [bx*h + head][row][d]
//  * lr_logits_kernel      : s_low = k . q^T per (image, head) on tcgen05 (bf16x3, fp32 accumulate), written as
//                            bf16 hi + lo in [b*h + head][token][source pixel] (K-major for the main kernel)
//  * lr_weights_kernel     : W tables [tile][128 px][KW taps] hi + lo
#include "attn_tc.cuh"

namespace skp {

constexpr int LR_MAX_K = 128;          // taps per tile the main kernels accept (two 64-wide K blocks)
constexpr int LRG_THREADS = 192;       // warp 0: TMA + MMA issue, warp 1 idle, warps 2-5: epilogue
constexpr int LRG_BM = 128;            // tokens per CTA
constexpr int LRG_BN = 256;            // source pixels per CTA
constexpr int LRG_A_BYTES = LRG_BM * TC_BK * 2;
constexpr int LRG_B_BYTES = LRG_BN * TC_BK * 2;
constexpr int LRG_STAGE = 2 * (LRG_A_BYTES + LRG_B_BYTES);     // 96 KB: two CTAs per SM
constexpr int LRG_DYNAMIC = LRG_STAGE + 64 + 1024;
constexpr int LRG_STG_PITCH = 128 * 2 + 16;                    // epilogue staging: 128 bf16 per row + 16 B pad
constexpr int LRG_STG_BYTES = 32 * LRG_STG_PITCH;              // per warp, per hi / lo
static_assert(4 * 2 * LRG_STG_BYTES <= LRG_STAGE, "epilogue staging must fit in the operand tiles");

struct LrGemmParams {
  CUtensorMap k_map[2];   // (d, T, Bk*h), box (64, 128, 1)
  CUtensorMap q_map[2];   // (d, s*s, B*h), box (64, 256, 1)
  __nv_bfloat16* out_hi;  // [B*h][T][s*s]
  __nv_bfloat16* out_lo;
  int* absmax;            // [B*h][2]: lr_range_key of (max s_low, max -s_low) of the (image, head), via atomicMax
  int T, s2, d, heads, k_shared;
};

LrGeom lr_geometry(int s, int R) {
  LrGeom g;
  const int tiles = R * R / TC_BM;
  const float ratio = static_cast<float>(s) / R;
  int span = 0;
  for (int t = 0; t < tiles; ++t) {
    int ia, ib;
    float w[4];
    bicubic_coeffs(t * TC_BM / R, ratio, &ia, w);
    bicubic_coeffs((t * TC_BM + TC_BM - 1) / R, ratio, &ib, w);
    span = ib - ia > span ? ib - ia : span;
  }
  g.nr = span + 4 < s ? span + 4 : s;
  g.kw = (g.nr * s + TC_BK - 1) / TC_BK * TC_BK;
  return g;
}

bool attn_map_lr_supported(const skp_attn_layer* h_layers, int n_layers, int B, int T, int R, int n_idx) {
  if (!attn_map_tc_supported(h_layers, n_layers, B, T, R, n_idx)) return false;
  for (int l = 0; l < n_layers; ++l) {
    const int s = h_layers[l].s;
    if ((s * s) % 8 != 0) return false;                 // 16-byte rows for TMA and the vector stores
    if (lr_geometry(s, R).kw > LR_MAX_K) return false;  // wide windows (64^2 sources) cost more MMAs than the staging saves
  }
  return true;
}

size_t attn_map_lr_workspace(const skp_attn_layer* h_layers, int n_layers, int B, int T, int R) {
  size_t total = 0;
  const size_t tiles = static_cast<size_t>(R) * R / TC_BM;
  for (int l = 0; l < n_layers; ++l) {
    const skp_attn_layer& a = h_layers[l];
    const size_t s2 = static_cast<size_t>(a.s) * a.s;
    total += 2 * tc_align_up(static_cast<size_t>(B) * s2 * a.C * 2, 1024);                     // q hi, lo
    total += 2 * tc_align_up(static_cast<size_t>(a.k_shared ? 1 : B) * T * a.C * 2, 1024);     // k hi, lo
    total += 2 * tc_align_up(static_cast<size_t>(B) * a.heads * T * s2 * 2, 1024);             // s_low hi, lo
    total += 2 * tc_align_up(tiles * TC_BM * lr_geometry(a.s, R).kw * 2, 1024);                // W hi, lo
    total += tc_align_up(static_cast<size_t>(B) * a.heads * 2 * sizeof(int), 1024);            // range of s_low per (image, head)
  }
  return total + 1024;
}

// src (Bx, rows, C) fp32 -> mult * src as bf16 hi (+ lo residual), [bx*h + head][row][d]
__global__ void lr_split_heads_kernel(const float* __restrict__ src, int rows, int C, int heads, float mult,
                                      __nv_bfloat16* __restrict__ hi, __nv_bfloat16* __restrict__ lo) {
  const int r = blockIdx.x, bx = blockIdx.y;
  const int d = C / heads;
  const float* p = src + (static_cast<int64_t>(bx) * rows + r) * C;
  for (int cp = threadIdx.x; cp < C / 2; cp += blockDim.x) {
    const int c = cp * 2;
    float2 v = *reinterpret_cast<const float2*>(p + c);
    v.x *= mult;
    v.y *= mult;
    const int head = c / d, dd = c % d;
    const int64_t o = ((static_cast<int64_t>(bx) * heads + head) * rows + r) * d + dd;
    const __nv_bfloat16 hx = __float2bfloat16_rn(v.x), hy = __float2bfloat16_rn(v.y);
    *reinterpret_cast<__nv_bfloat162*>(hi + o) = __halves2bfloat162(hx, hy);
    *reinterpret_cast<__nv_bfloat162*>(lo + o) =
        __halves2bfloat162(__float2bfloat16_rn(v.x - __bfloat162float(hx)), __float2bfloat16_rn(v.y - __bfloat162float(hy)));
  }
}

// W[tile][px][tap]: tap = (r - r0) * s + c over the tile's window of source rows [r0, r0 + nr); the clamped
// border taps of the reference (ATen upsample_bicubic2d, align_corners=False) fold into the edge rows / columns
__global__ void lr_weights_kernel(int s, int R, int nr, int kw, __nv_bfloat16* __restrict__ hi, __nv_bfloat16* __restrict__ lo) {
  extern __shared__ float lr_w_sm[];          // ey[128][nr], ex[128][s]
  float* ey = lr_w_sm;
  float* ex = lr_w_sm + TC_BM * nr;
  const int tile = blockIdx.x;
  const float ratio = static_cast<float>(s) / R;
  const int r0 = lr_window_row0(tile, s, R, nr);
  for (int px = threadIdx.x; px < TC_BM; px += blockDim.x) {
    const int p = tile * TC_BM + px, y = p / R, x = p - y * R;
    int iy, ix;
    float wy[4], wx[4];
    bicubic_coeffs(y, ratio, &iy, wy);
    bicubic_coeffs(x, ratio, &ix, wx);
    for (int r = 0; r < nr; ++r) {
      float e = 0.0f;
#pragma unroll
      for (int a = 0; a < 4; ++a) e += (min(max(iy - 1 + a, 0), s - 1) == r0 + r) ? wy[a] : 0.0f;
      ey[px * nr + r] = e;
    }
    for (int c = 0; c < s; ++c) {
      float e = 0.0f;
#pragma unroll
      for (int a = 0; a < 4; ++a) e += (min(max(ix - 1 + a, 0), s - 1) == c) ? wx[a] : 0.0f;
      ex[px * s + c] = e;
    }
  }
  __syncthreads();
  const int kw2 = kw >> 1;
  for (int e = threadIdx.x; e < TC_BM * kw2; e += blockDim.x) {
    const int px = e / kw2, tap = (e - px * kw2) * 2;
    float w[2];
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int r = (tap + j) / s, c = (tap + j) - r * s;
      w[j] = tap + j < nr * s ? ey[px * nr + r] * ex[px * s + c] : 0.0f;
    }
    const int64_t o = (static_cast<int64_t>(tile) * TC_BM + px) * kw + tap;
    const __nv_bfloat16 h0 = __float2bfloat16_rn(w[0]), h1 = __float2bfloat16_rn(w[1]);
    *reinterpret_cast<__nv_bfloat162*>(hi + o) = __halves2bfloat162(h0, h1);
    *reinterpret_cast<__nv_bfloat162*>(lo + o) =
        __halves2bfloat162(__float2bfloat16_rn(w[0] - __bfloat162float(h0)), __float2bfloat16_rn(w[1] - __bfloat162float(h1)));
  }
}

int lr_launch_weights(int s, int R, const LrGeom& g, __nv_bfloat16* hi, __nv_bfloat16* lo, cudaStream_t st) {
  lr_weights_kernel<<<(::skp::note_launch(), R * R / TC_BM), 256, static_cast<size_t>(TC_BM) * (g.nr + s) * sizeof(float), st>>>(s, R, g.nr, g.kw, hi, lo);
  SKP_LAUNCH_CHECK();
  return SKP_OK;
}

// s_low[z][t][sp] = sum_d k[kz][t][d] * q[z][sp][d]   (q already carries scale * log2e), bf16x3
__global__ void __launch_bounds__(LRG_THREADS) lr_logits_kernel(const __grid_constant__ LrGemmParams P) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = ptx::smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  uint8_t* base_ptr = smem_raw + (base - raw);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tt = blockIdx.x, st = blockIdx.y, z = blockIdx.z;
  const int b = z / P.heads, head = z - b * P.heads;
  const int kz = (P.k_shared ? 0 : b * P.heads) + head;
  const uint32_t a_hi = base, b_hi = base + LRG_A_BYTES, a_lo = b_hi + LRG_B_BYTES, b_lo = a_lo + LRG_A_BYTES;
  const uint32_t bar_full = base + LRG_STAGE, bar_done = bar_full + 8, bar_acc = bar_full + 16;
  volatile uint32_t* tmem_ptr = reinterpret_cast<volatile uint32_t*>(base_ptr + LRG_STAGE + 32);

  if (threadIdx.x == 0) {
    ptx::mbar_init(bar_full, 1);
    ptx::mbar_init(bar_done, 1);
    ptx::mbar_init(bar_acc, 1);
    ptx::fence_barrier_init();
  }
  if (warp == 0) {
    ptx::tmem_alloc(ptx::smem_u32(const_cast<uint32_t*>(tmem_ptr)), LRG_BN);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == 0) {
    if (lane == 0) {
      constexpr uint32_t idesc = ptx::umma_idesc_bf16(LRG_BM, LRG_BN);
      const int nkb = (P.d + TC_BK - 1) / TC_BK;
      for (int kb = 0; kb < nkb; ++kb) {
        if (kb > 0) ptx::mbar_wait(bar_done, (kb - 1) & 1u);      // the previous block's MMAs are done with the tiles
        ptx::mbar_arrive_expect_tx(bar_full, LRG_STAGE);
        ptx::tma_load_3d(a_hi, &P.k_map[0], bar_full, kb * TC_BK, tt * LRG_BM, kz);
        ptx::tma_load_3d(b_hi, &P.q_map[0], bar_full, kb * TC_BK, st * LRG_BN, z);
        ptx::tma_load_3d(a_lo, &P.k_map[1], bar_full, kb * TC_BK, tt * LRG_BM, kz);
        ptx::tma_load_3d(b_lo, &P.q_map[1], bar_full, kb * TC_BK, st * LRG_BN, z);
        ptx::mbar_wait(bar_full, kb & 1u);
        ptx::tc_fence_after();
        const int nk16 = (min(TC_BK, P.d - kb * TC_BK) + 15) / 16;
        for (int kk = 0; kk < nk16; ++kk) {
          const uint64_t dah = ptx::umma_desc_sw128(a_hi + kk * 32), dbh = ptx::umma_desc_sw128(b_hi + kk * 32);
          const uint64_t dal = ptx::umma_desc_sw128(a_lo + kk * 32), dbl = ptx::umma_desc_sw128(b_lo + kk * 32);
          ptx::umma_bf16(tmem_base, dah, dbl, idesc, (kb | kk) != 0);
          ptx::umma_bf16(tmem_base, dal, dbh, idesc, 1u);
          ptx::umma_bf16(tmem_base, dah, dbh, idesc, 1u);
        }
        ptx::umma_commit(bar_done);
      }
      ptx::umma_commit(bar_acc);
    }
  } else if (warp >= 2) {
    ptx::mbar_wait_hint(bar_acc, 0u, 1000u);
    ptx::tc_fence_after();
    // TMEM lane = token, columns = source pixels, but the output rows [token][s*s] want a warp to write whole rows:
    // each warp transposes 32 tokens x 128 pixels at a time through its own slice of the (now idle) operand tiles
    // (row pitch 272 B: the 16-byte stores of 32 lanes spread over all banks), then writes 2 full 256-byte rows per
    // instruction.  Direct stores from the TMEM layout (32 lanes x 16 B at a 512 B stride) ran the kernel at 4x its
    // HBM time.
    const int quad = warp & 3;
    const uint32_t t_lane = tmem_base + (static_cast<uint32_t>(quad * 32) << 16);
    uint8_t* stg_hi = base_ptr + (warp - 2) * (2 * LRG_STG_BYTES);
    uint8_t* stg_lo = stg_hi + LRG_STG_BYTES;
    const int t0 = tt * LRG_BM + quad * 32;
    // rows / columns past T / s*s hold zeros (TMA zero fill): they can only widen the range towards 0, which stays valid
    float vmax = -INFINITY, vmin = INFINITY;
#pragma unroll 1
    for (int half = 0; half < 2; ++half) {
      const int col0 = st * LRG_BN + half * 128;
      if (col0 >= P.s2) break;
#pragma unroll 1
      for (int c4 = 0; c4 < 4; ++c4) {
        uint32_t v[32];
        ptx::tmem_ld32(t_lane + half * 128 + c4 * 32, v);
        ptx::tmem_ld_wait();
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          uint32_t ph[4], pl[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float f0 = __uint_as_float(v[8 * q + 2 * j]), f1 = __uint_as_float(v[8 * q + 2 * j + 1]);
            vmax = fmaxf(vmax, fmaxf(f0, f1));
            vmin = fminf(vmin, fminf(f0, f1));
            const __nv_bfloat16 h0 = __float2bfloat16_rn(f0), h1 = __float2bfloat16_rn(f1);
            const __nv_bfloat162 hh = __halves2bfloat162(h0, h1);
            const __nv_bfloat162 ll = __halves2bfloat162(__float2bfloat16_rn(f0 - __bfloat162float(h0)),
                                                         __float2bfloat16_rn(f1 - __bfloat162float(h1)));
            ph[j] = *reinterpret_cast<const uint32_t*>(&hh);
            pl[j] = *reinterpret_cast<const uint32_t*>(&ll);
          }
          const int off = lane * LRG_STG_PITCH + (c4 * 32 + q * 8) * 2;
          *reinterpret_cast<uint4*>(stg_hi + off) = make_uint4(ph[0], ph[1], ph[2], ph[3]);
          *reinterpret_cast<uint4*>(stg_lo + off) = make_uint4(pl[0], pl[1], pl[2], pl[3]);
        }
      }
      __syncwarp();
#pragma unroll 4
      for (int it = 0; it < 16; ++it) {
        const int r = it * 2 + (lane >> 4), cb = (lane & 15) * 8;       // row of the slice, first of 8 pixels
        const int t = t0 + r;
        if (t < P.T && col0 + cb + 8 <= P.s2) {
          const int64_t o = (static_cast<int64_t>(z) * P.T + t) * P.s2 + col0 + cb;
          *reinterpret_cast<uint4*>(P.out_hi + o) = *reinterpret_cast<const uint4*>(stg_hi + r * LRG_STG_PITCH + cb * 2);
          *reinterpret_cast<uint4*>(P.out_lo + o) = *reinterpret_cast<const uint4*>(stg_lo + r * LRG_STG_PITCH + cb * 2);
        }
      }
      __syncwarp();
    }
    vmax = warp_max(vmax);
    vmin = -warp_max(-vmin);
    if (lane == 0 && P.absmax && vmax >= vmin) {
      atomicMax(P.absmax + 2 * z, lr_range_key(vmax));
      atomicMax(P.absmax + 2 * z + 1, lr_range_key(-vmin));
    }
  }
  __syncwarp();
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, LRG_BN);
  }
}

char* lr_carve(const skp_attn_layer* h_layers, int n_layers, int B, int T, int R, void* workspace, LrLayerBufs* out) {
  char* ws = reinterpret_cast<char*>(tc_align_up(reinterpret_cast<size_t>(workspace), 1024));
  const int tiles = R * R / TC_BM;
  for (int l = 0; l < n_layers; ++l) {
    const skp_attn_layer& a = h_layers[l];
    const int s2 = a.s * a.s, Bk = a.k_shared ? 1 : B;
    const LrGeom g = lr_geometry(a.s, R);
    auto take = [&](size_t bytes) {
      __nv_bfloat16* p = reinterpret_cast<__nv_bfloat16*>(ws);
      ws += tc_align_up(bytes, 1024);
      return p;
    };
    const size_t qb = static_cast<size_t>(B) * s2 * a.C * 2, kb = static_cast<size_t>(Bk) * T * a.C * 2;
    const size_t sb = static_cast<size_t>(B) * a.heads * T * s2 * 2, wb = static_cast<size_t>(tiles) * TC_BM * g.kw * 2;
    LrLayerBufs& o = out[l];
    o.q_hi = take(qb); o.q_lo = take(qb); o.k_hi = take(kb); o.k_lo = take(kb); o.s_hi = take(sb); o.s_lo = take(sb);
    o.w_hi = take(wb); o.w_lo = take(wb);            // reserved even when an earlier layer's table is reused
    o.range = reinterpret_cast<int*>(take(static_cast<size_t>(B) * a.heads * 2 * sizeof(int)));
    for (int m = 0; m < l; ++m)                        // one interpolation table per source size: the first layer's
      if (h_layers[m].s == a.s) {
        o.w_hi = out[m].w_hi;
        o.w_lo = out[m].w_lo;
        break;
      }
  }
  return ws;
}

// Stages the operands of the "upsample after the logits" formulation in `workspace` (launches on `st`) and fills
// `O` so that the main kernels see A = W tables (q_map) and B = s_low (k_map) with d[l] = taps per tile.
int tc_prepare_lr(const skp_attn_layer* h_layers, int n_layers, int B, int T, int R, void* workspace, cudaStream_t st,
                  TcOperands* O, char** ws_end, int bn, const LrLayerBufs* ready) {
  LrLayerBufs own[TC_MAX_LAYERS];
  char* ws = reinterpret_cast<char*>(tc_align_up(reinterpret_cast<size_t>(workspace), 1024));
  if (!ready) ws = lr_carve(h_layers, n_layers, B, T, R, workspace, own);
  const LrLayerBufs* bufs = ready ? ready : own;
  const int tiles = R * R / TC_BM;
  int base = 0;
  if (!ready) SKP_CUDA(cudaFuncSetAttribute(lr_logits_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, LRG_DYNAMIC));
  for (int l = 0; l < n_layers; ++l) {
    const skp_attn_layer& a = h_layers[l];
    SKP_REQUIRE(a.q && a.k, "skp_attn_map: layer %d has null q/k", l);
    const int d = a.C / a.heads, s2 = a.s * a.s, Bk = a.k_shared ? 1 : B;
    const LrGeom g = lr_geometry(a.s, R);
    __nv_bfloat16 *q_hi = bufs[l].q_hi, *q_lo = bufs[l].q_lo, *k_hi = bufs[l].k_hi, *k_lo = bufs[l].k_lo;
    __nv_bfloat16 *s_hi = bufs[l].s_hi, *s_lo = bufs[l].s_lo, *w_hi = bufs[l].w_hi, *w_lo = bufs[l].w_lo;
    int* absmax = bufs[l].range;
    if (!ready) {
    SKP_CUDA(cudaMemsetAsync(absmax, 0x80, static_cast<size_t>(B) * a.heads * 2 * sizeof(int), st));   // keys of ~ -3e38: below every logit
    const float mult = a.scale * 1.4426950408889634f;
    lr_split_heads_kernel<<<(::skp::note_launch(), dim3(s2, B)), 256, 0, st>>>(a.q, s2, a.C, a.heads, mult, q_hi, q_lo);
    lr_split_heads_kernel<<<(::skp::note_launch(), dim3(T, Bk)), 256, 0, st>>>(a.k, T, a.C, a.heads, 1.0f, k_hi, k_lo);
    bool own_table = true;
    for (int m = 0; m < l; ++m)
      if (h_layers[m].s == a.s) own_table = false;
    if (own_table)
      lr_weights_kernel<<<(::skp::note_launch(), tiles), 256, static_cast<size_t>(TC_BM) * (g.nr + a.s) * sizeof(float), st>>>(a.s, R, g.nr, g.kw, w_hi, w_lo);
    SKP_LAUNCH_CHECK();
    LrGemmParams G;
    memset(&G, 0, sizeof(G));
    int rc = tc_encode_map(&G.k_map[0], k_hi, d, T, Bk * a.heads, LRG_BM);
    if (rc == SKP_OK) rc = tc_encode_map(&G.k_map[1], k_lo, d, T, Bk * a.heads, LRG_BM);
    if (rc == SKP_OK) rc = tc_encode_map(&G.q_map[0], q_hi, d, s2, B * a.heads, LRG_BN);
    if (rc == SKP_OK) rc = tc_encode_map(&G.q_map[1], q_lo, d, s2, B * a.heads, LRG_BN);
    if (rc != SKP_OK) return rc;
    G.out_hi = s_hi; G.out_lo = s_lo; G.absmax = absmax; G.T = T; G.s2 = s2; G.d = d; G.heads = a.heads; G.k_shared = a.k_shared;
    dim3 grid((T + LRG_BM - 1) / LRG_BM, (s2 + LRG_BN - 1) / LRG_BN, B * a.heads);
    lr_logits_kernel<<<(::skp::note_launch(), grid), LRG_THREADS, LRG_DYNAMIC, st>>>(G);
    SKP_LAUNCH_CHECK();
    }
    int rc = tc_encode_map(&O->q_map[l][0], w_hi, g.kw, TC_BM, tiles, TC_BM);
    if (rc == SKP_OK) rc = tc_encode_map(&O->q_map[l][1], w_lo, g.kw, TC_BM, tiles, TC_BM);
    if (rc == SKP_OK) rc = tc_encode_map(&O->k_map[l][0], s_hi, s2, T, B * a.heads, bn);
    if (rc == SKP_OK) rc = tc_encode_map(&O->k_map[l][1], s_lo, s2, T, B * a.heads, bn);
    if (rc != SKP_OK) return rc;
    O->k_ptr[l][0] = k_hi; O->k_ptr[l][1] = k_lo;
    O->s_ptr[l][0] = s_hi; O->s_ptr[l][1] = s_lo;
    O->heads[l] = a.heads; O->d[l] = g.kw; O->k_shared[l] = 0; O->head_base[l] = base;
    O->weight[l] = a.weight / a.heads; O->scale[l] = a.scale;
    O->win_s[l] = a.s; O->win_nr[l] = g.nr; O->absmax[l] = absmax;
    base += a.heads;
  }
  O->n_layers = n_layers; O->total_heads = base; O->lr = 1; O->R = R;
  if (ws_end) *ws_end = ws;
  return SKP_OK;
}

}  // namespace skp
