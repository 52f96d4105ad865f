// CUDA-core (fp32 FMA) attention-map kernels: forward and backward.
//
// These are the validation / any-shape kernels behind SKP_PREC_SIMT_F32.  They implement the
// same contract as the tcgen05 kernels in attn_tc.cu with plain fp32 arithmetic, so they serve
// as the on-device cross-check for the tensor-core path and cover shapes it does not
// (dim_head not a multiple of 16, T > 512, ...).
//
// Forward: one CTA owns (image b, 16 consecutive feature pixels) and loops over every captured
// (layer, head): builds the bicubically upsampled, pre-scaled q' tile from the low-resolution
// q, streams k through shared memory in 128-token chunks, keeps the 16 x T logits in shared
// memory, takes the softmax over tokens and accumulates weight/heads * P into a 16 x T'
// accumulator that lives in shared memory and is written once.  Nothing of size (h, R^2, T)
// ever reaches HBM (the reference stores 262-655 MB per layer: ptp_utils.py:535-538).
#include "common.cuh"

namespace skp {

constexpr int SIMT_TP = 16;        // pixels per CTA
constexpr int SIMT_TT = 128;       // tokens per k chunk
constexpr int SIMT_THREADS = 256;
constexpr int SIMT_MAX_LAYERS = 16;
constexpr float kLog2e = 1.4426950408889634f;

struct SimtLayers {
  const float* q[SIMT_MAX_LAYERS];
  const float* k[SIMT_MAX_LAYERS];
  float* dq_up[SIMT_MAX_LAYERS];   // backward: (B, R*R, C) gradient w.r.t. the UPSAMPLED q
  float* dk[SIMT_MAX_LAYERS];
  int s[SIMT_MAX_LAYERS], C[SIMT_MAX_LAYERS], heads[SIMT_MAX_LAYERS], k_shared[SIMT_MAX_LAYERS];
  int head_base[SIMT_MAX_LAYERS];  // prefix sum of heads (row_stats / grid indexing)
  float scale[SIMT_MAX_LAYERS], weight[SIMT_MAX_LAYERS];
  int n_layers, total_heads, dmax;
};

// q' tile: Qs[px][c] = scale*log2e * bicubic(q_low)[p0+px][head*d + c]
__device__ __forceinline__ void build_q_tile(const float* __restrict__ q, int s, int C, int d, int head, int R, int p0,
                                             float mult, float* Qs, int ldq) {
  const float ratio = static_cast<float>(s) / R;
  for (int e = threadIdx.x; e < SIMT_TP * d; e += blockDim.x) {
    const int px = e / d, c = e % d;
    const int p = p0 + px;
    const int oy = p / R, ox = p % R;
    float v;
    if (s == R) {
      v = q[static_cast<int64_t>(p) * C + head * d + c];
    } else {
      int iy, ix;
      float wy[4], wx[4];
      bicubic_coeffs(oy, ratio, &iy, wy);
      bicubic_coeffs(ox, ratio, &ix, wx);
      const float* src = q + head * d + c;
      v = 0.0f;
#pragma unroll
      for (int a = 0; a < 4; ++a) {
        const int yy = min(max(iy - 1 + a, 0), s - 1);
        float row = 0.0f;
#pragma unroll
        for (int b = 0; b < 4; ++b) {
          const int xx = min(max(ix - 1 + b, 0), s - 1);
          row += wx[b] * src[static_cast<int64_t>(yy * s + xx) * C];
        }
        v += wy[a] * row;
      }
    }
    Qs[px * ldq + c] = v * mult;
  }
}

// Ss[px][t] for all t: streams k in chunks through Ks.  Leaves Ss in log2 units.
__device__ __forceinline__ void logits_tile(const float* __restrict__ k, int C, int d, int head, int T, const float* Qs,
                                            int ldq, float* Ks, int ldk, float* Ss, int lds) {
  const int ty = threadIdx.x >> 5;   // 8 groups of 2 pixels
  const int tx = threadIdx.x & 31;   // tokens tx + 32 j
  for (int t0 = 0; t0 < T; t0 += SIMT_TT) {
    const int nt = min(SIMT_TT, T - t0);
    __syncthreads();
    for (int e = threadIdx.x; e < SIMT_TT * d; e += blockDim.x) {
      const int t = e / d, c = e % d;
      Ks[t * ldk + c] = t < nt ? k[static_cast<int64_t>(t0 + t) * C + head * d + c] : 0.0f;
    }
    __syncthreads();
    float acc[2][4];
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = 0.0f;
    for (int c = 0; c < d; ++c) {
      const float q0 = Qs[(2 * ty) * ldq + c], q1 = Qs[(2 * ty + 1) * ldq + c];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float kv = Ks[(tx + 32 * j) * ldk + c];
        acc[0][j] = fmaf(q0, kv, acc[0][j]);
        acc[1][j] = fmaf(q1, kv, acc[1][j]);
      }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int t = t0 + tx + 32 * j;
      if (t < T) {
        Ss[(2 * ty) * lds + t] = acc[0][j];
        Ss[(2 * ty + 1) * lds + t] = acc[1][j];
      }
    }
  }
  __syncthreads();
}

// softmax over tokens, in place (Ss -> P); one warp per 2 rows.  stats (optional): (max, sum)
__device__ __forceinline__ void softmax_rows(float* Ss, int lds, int T, float* stat_m, float* stat_l) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int r = 2 * warp; r < 2 * warp + 2; ++r) {
    float* row = Ss + r * lds;
    float m = -INFINITY;
    for (int t = lane; t < T; t += 32) m = fmaxf(m, row[t]);
    m = warp_max(m);
    float l = 0.0f;
    for (int t = lane; t < T; t += 32) {
      float e = exp2f(row[t] - m);
      row[t] = e;
      l += e;
    }
    l = warp_sum(l);
    const float inv = 1.0f / l;
    for (int t = lane; t < T; t += 32) row[t] *= inv;
    if (lane == 0 && stat_m) {
      stat_m[r] = m;
      stat_l[r] = l;
    }
  }
  __syncthreads();
}

__global__ void __launch_bounds__(SIMT_THREADS, 1)
attn_map_fwd_simt_kernel(const SimtLayers L, int T, int R, const int64_t* __restrict__ token_idx, int n_idx,
                         int64_t idx_stride, float* __restrict__ maps, float* __restrict__ row_stats) {
  extern __shared__ float smem[];
  const int Tout = token_idx ? n_idx : T;
  const int ldq = L.dmax + 1, ldk = L.dmax + 1, lds = T + 1;
  float* Qs = smem;                          // [TP][ldq]
  float* Ks = Qs + SIMT_TP * ldq;            // [TT][ldk]
  float* Ss = Ks + SIMT_TT * ldk;            // [TP][lds]
  float* Acc = Ss + SIMT_TP * lds;           // [Tout][TP]
  __shared__ float st_m[SIMT_TP], st_l[SIMT_TP];
  const int b = blockIdx.y;
  const int p0 = blockIdx.x * SIMT_TP;
  const int R2 = R * R;
  if (token_idx) token_idx += b * idx_stride;
  for (int e = threadIdx.x; e < Tout * SIMT_TP; e += blockDim.x) Acc[e] = 0.0f;
  for (int l = 0; l < L.n_layers; ++l) {
    const int C = L.C[l], h = L.heads[l], d = C / h, s = L.s[l];
    const float* q = L.q[l] + static_cast<int64_t>(b) * s * s * C;
    const float* k = L.k[l] + (L.k_shared[l] ? 0 : static_cast<int64_t>(b) * T * C);
    const float wgt = L.weight[l] / h;
    for (int head = 0; head < h; ++head) {
      __syncthreads();
      build_q_tile(q, s, C, d, head, R, p0, L.scale[l] * kLog2e, Qs, ldq);
      logits_tile(k, C, d, head, T, Qs, ldq, Ks, ldk, Ss, lds);
      softmax_rows(Ss, lds, T, row_stats ? st_m : nullptr, st_l);
      if (row_stats && threadIdx.x < SIMT_TP) {
        float* rs = row_stats + ((static_cast<int64_t>(b) * L.total_heads + L.head_base[l] + head) * R2 + p0 + threadIdx.x) * 2;
        rs[0] = st_m[threadIdx.x];
        rs[1] = st_l[threadIdx.x];
      }
      for (int e = threadIdx.x; e < Tout * SIMT_TP; e += blockDim.x) {
        const int j = e / SIMT_TP, px = e % SIMT_TP;
        const int t = token_idx ? static_cast<int>(token_idx[j]) : j;
        Acc[e] = fmaf(wgt, Ss[px * lds + t], Acc[e]);
      }
    }
  }
  __syncthreads();
  for (int e = threadIdx.x; e < Tout * SIMT_TP; e += blockDim.x) {
    const int j = e / SIMT_TP, px = e % SIMT_TP;
    maps[(static_cast<int64_t>(b) * Tout + j) * R2 + p0 + px] = Acc[e];
  }
}

// Backward: grid (tiles, total_heads, B).  Recomputes P for its (layer, head, 16 pixels), forms
// dS = P * (dP - delta) with dP non-zero on the n_idx selected tokens only, then
//   dq'[px][c] = scale * sum_t dS[px][t] k[t][c]          (written, each element owned once)
//   dk[t][c]  += scale * sum_px dS[px][t] q'[px][c]       (atomics across pixel tiles / images)
__global__ void __launch_bounds__(SIMT_THREADS, 1)
attn_map_bwd_simt_kernel(const SimtLayers L, int T, int R, const int64_t* __restrict__ token_idx, int n_idx,
                         int64_t idx_stride, const float* __restrict__ dmaps_sel) {
  extern __shared__ float smem[];
  const int ldq = L.dmax + 1, ldk = L.dmax + 1, lds = T + 1;
  float* Qs = smem;                          // [TP][ldq]   (unscaled-by-log2e copy kept in Qr)
  float* Ks = Qs + SIMT_TP * ldq;            // [TT][ldk]
  float* Ss = Ks + SIMT_TT * ldk;            // [TP][lds]
  float* dQ = Ss + SIMT_TP * lds;            // [TP][ldq]
  __shared__ float delta[SIMT_TP];
  const int b = blockIdx.z;
  const int p0 = blockIdx.x * SIMT_TP;
  const int R2 = R * R;
  token_idx += b * idx_stride;
  int l = 0;
  while (l + 1 < L.n_layers && static_cast<int>(blockIdx.y) >= L.head_base[l + 1]) ++l;
  const int head = blockIdx.y - L.head_base[l];
  const int C = L.C[l], h = L.heads[l], d = C / h, s = L.s[l];
  const float* q = L.q[l] + static_cast<int64_t>(b) * s * s * C;
  const float* k = L.k[l] + (L.k_shared[l] ? 0 : static_cast<int64_t>(b) * T * C);
  float* dk = L.dk[l] + (L.k_shared[l] ? 0 : static_cast<int64_t>(b) * T * C);
  float* dq_up = L.dq_up[l] + static_cast<int64_t>(b) * R2 * C;
  const float wgt = L.weight[l] / h;
  const float scale = L.scale[l];

  build_q_tile(q, s, C, d, head, R, p0, scale * kLog2e, Qs, ldq);
  logits_tile(k, C, d, head, T, Qs, ldq, Ks, ldk, Ss, lds);
  softmax_rows(Ss, lds, T, nullptr, nullptr);
  // delta[px] = sum_j P[px][idx_j] * dP[px][j]
  if (threadIdx.x < SIMT_TP) {
    const int px = threadIdx.x;
    float dl = 0.0f;
    for (int j = 0; j < n_idx; ++j) {
      const float dp = wgt * dmaps_sel[(static_cast<int64_t>(b) * n_idx + j) * R2 + p0 + px];
      dl = fmaf(Ss[px * lds + static_cast<int>(token_idx[j])], dp, dl);
    }
    delta[px] = dl;
  }
  __syncthreads();
  // dS = scale * P * (dP - delta): dense part first, then the sparse correction.  The sparse
  // term needs the ORIGINAL P of the selected tokens, so stash them before overwriting.
  float* Psel = dQ;   // reuse dQ storage temporarily: [TP][n_idx] (n_idx <= ldq guaranteed by host)
  for (int e = threadIdx.x; e < SIMT_TP * n_idx; e += blockDim.x) {
    const int px = e / n_idx, j = e % n_idx;
    Psel[px * n_idx + j] = Ss[px * lds + static_cast<int>(token_idx[j])];
  }
  __syncthreads();
  for (int e = threadIdx.x; e < SIMT_TP * T; e += blockDim.x) {
    const int px = e / T, t = e % T;
    Ss[px * lds + t] *= -scale * delta[px];
  }
  __syncthreads();
  if (threadIdx.x < SIMT_TP) {   // one thread per row: duplicates in token_idx accumulate correctly
    const int px = threadIdx.x;
    for (int j = 0; j < n_idx; ++j) {
      const float dp = wgt * dmaps_sel[(static_cast<int64_t>(b) * n_idx + j) * R2 + p0 + px];
      Ss[px * lds + static_cast<int>(token_idx[j])] += scale * Psel[px * n_idx + j] * dp;
    }
  }
  __syncthreads();
  for (int e = threadIdx.x; e < SIMT_TP * ldq; e += blockDim.x) dQ[e] = 0.0f;
  // Qs holds q' * scale*log2e; dk needs plain q'
  const float unscale = 1.0f / (scale * kLog2e);
  for (int t0 = 0; t0 < T; t0 += SIMT_TT) {
    const int nt = min(SIMT_TT, T - t0);
    __syncthreads();
    for (int e = threadIdx.x; e < SIMT_TT * d; e += blockDim.x) {
      const int t = e / d, c = e % d;
      Ks[t * ldk + c] = t < nt ? k[static_cast<int64_t>(t0 + t) * C + head * d + c] : 0.0f;
    }
    __syncthreads();
    // dq'[px][c] += sum_t dS[px][t0+t] * K[t][c]
    for (int e = threadIdx.x; e < SIMT_TP * d; e += blockDim.x) {
      const int px = e / d, c = e % d;
      float a = 0.0f;
      for (int t = 0; t < nt; ++t) a = fmaf(Ss[px * lds + t0 + t], Ks[t * ldk + c], a);
      dQ[px * ldq + c] += a;
    }
    // dk[t][c] += sum_px dS[px][t0+t] * q'[px][c]
    for (int e = threadIdx.x; e < nt * d; e += blockDim.x) {
      const int t = e / d, c = e % d;
      float a = 0.0f;
#pragma unroll
      for (int px = 0; px < SIMT_TP; ++px) a = fmaf(Ss[px * lds + t0 + t], Qs[px * ldq + c], a);
      atomicAdd(dk + static_cast<int64_t>(t0 + t) * C + head * d + c, a * unscale);
    }
  }
  __syncthreads();
  for (int e = threadIdx.x; e < SIMT_TP * d; e += blockDim.x) {
    const int px = e / d, c = e % d;
    dq_up[static_cast<int64_t>(p0 + px) * C + head * d + c] = dQ[px * ldq + c];
  }
}

// Adjoint of the bicubic upsample, separable and gather-form (deterministic, no atomics).
// pass 1: tmp[b][y][sx][c] = sum_x wx(x -> sx) * g[b][y][x][c]
__global__ void bicubic_adjoint_x_kernel(const float* __restrict__ g, int s, int C, int R, float* __restrict__ tmp) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int sx = blockIdx.y % s, y = blockIdx.y / s;
  const int b = blockIdx.z;
  if (c >= C) return;
  const float ratio = static_cast<float>(s) / R;
  const int up = R / s;
  const int xlo = max(0, (sx - 2) * up), xhi = min(R, (sx + 3) * up);
  const float* row = g + (static_cast<int64_t>(b) * R + y) * R * C + c;
  float acc = 0.0f;
  for (int x = xlo; x < xhi; ++x) {
    int ix;
    float w[4];
    bicubic_coeffs(x, ratio, &ix, w);
    float wsum = 0.0f;
#pragma unroll
    for (int a = 0; a < 4; ++a) wsum += (min(max(ix - 1 + a, 0), s - 1) == sx) ? w[a] : 0.0f;
    if (wsum != 0.0f) acc = fmaf(wsum, row[static_cast<int64_t>(x) * C], acc);
  }
  tmp[((static_cast<int64_t>(b) * R + y) * s + sx) * C + c] = acc;
}
// pass 2: dq[b][sy][sx][c] += sum_y wy(y -> sy) * tmp[b][y][sx][c]
__global__ void bicubic_adjoint_y_kernel(const float* __restrict__ tmp, int s, int C, int R, float* __restrict__ dq) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int sx = blockIdx.y % s, sy = blockIdx.y / s;
  const int b = blockIdx.z;
  if (c >= C) return;
  const float ratio = static_cast<float>(s) / R;
  const int up = R / s;
  const int ylo = max(0, (sy - 2) * up), yhi = min(R, (sy + 3) * up);
  float acc = 0.0f;
  for (int y = ylo; y < yhi; ++y) {
    int iy;
    float w[4];
    bicubic_coeffs(y, ratio, &iy, w);
    float wsum = 0.0f;
#pragma unroll
    for (int a = 0; a < 4; ++a) wsum += (min(max(iy - 1 + a, 0), s - 1) == sy) ? w[a] : 0.0f;
    if (wsum != 0.0f) acc = fmaf(wsum, tmp[((static_cast<int64_t>(b) * R + y) * s + sx) * C + c], acc);
  }
  dq[(static_cast<int64_t>(b) * s * s + sy * s + sx) * C + c] += acc;
}
__global__ void add_inplace_kernel(float* __restrict__ dst, const float* __restrict__ src, int64_t n) {
  int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < n) dst[i] += src[i];
}

int fill_simt_layers(const skp_attn_layer* h_layers, int n_layers, int B, int T, int R, SimtLayers* L) {
  SKP_REQUIRE(h_layers && n_layers >= 1 && n_layers <= SIMT_MAX_LAYERS, "attn_map: 1 <= n_layers <= %d", SIMT_MAX_LAYERS);
  SKP_REQUIRE(B >= 1 && T >= 1 && R >= 1 && (R * R) % SIMT_TP == 0, "attn_map: need B,T >= 1 and R*R %% 16 == 0");
  int base = 0, dmax = 0;
  for (int l = 0; l < n_layers; ++l) {
    const skp_attn_layer& a = h_layers[l];
    SKP_REQUIRE(a.q && a.k, "attn_map: layer %d has null q/k", l);
    SKP_REQUIRE(a.heads >= 1 && a.C >= a.heads && a.C % a.heads == 0, "attn_map: layer %d: C %% heads != 0", l);
    SKP_REQUIRE(a.s >= 1 && R % a.s == 0, "attn_map: layer %d: R %% s != 0", l);
    L->q[l] = a.q; L->k[l] = a.k; L->dq_up[l] = nullptr; L->dk[l] = a.dk;
    L->s[l] = a.s; L->C[l] = a.C; L->heads[l] = a.heads; L->k_shared[l] = a.k_shared;
    L->scale[l] = a.scale; L->weight[l] = a.weight; L->head_base[l] = base;
    base += a.heads;
    dmax = dmax > a.C / a.heads ? dmax : a.C / a.heads;
  }
  L->n_layers = n_layers; L->total_heads = base; L->dmax = dmax;
  return SKP_OK;
}

int attn_map_fwd_simt(const skp_attn_layer* h_layers, int n_layers, int B, int T, int R, const int64_t* token_idx,
                      int n_idx, int64_t idx_stride, float* maps, float* row_stats, cudaStream_t st) {
  SimtLayers L;
  int rc = fill_simt_layers(h_layers, n_layers, B, T, R, &L);
  if (rc != SKP_OK) return rc;
  const int Tout = token_idx ? n_idx : T;
  const size_t smem = sizeof(float) * (static_cast<size_t>(SIMT_TP) * (L.dmax + 1) + SIMT_TT * (L.dmax + 1) +
                                       SIMT_TP * (T + 1) + static_cast<size_t>(Tout) * SIMT_TP);
  SKP_REQUIRE(smem <= 227 * 1024, "attn_map_fwd (SIMT): T=%d, dim_head=%d need %zu B of shared memory (> 227 KB)", T, L.dmax, smem);
  SKP_CUDA(cudaFuncSetAttribute(attn_map_fwd_simt_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
  dim3 grid(R * R / SIMT_TP, B);
  attn_map_fwd_simt_kernel<<<(::skp::note_launch(), grid), SIMT_THREADS, smem, st>>>(L, T, R, token_idx, n_idx, idx_stride, maps, row_stats);
  SKP_LAUNCH_CHECK();
  return SKP_OK;
}

size_t attn_map_bwd_simt_workspace(const skp_attn_layer* h_layers, int n_layers, int B, int R) {
  // per layer: dq_up (B,R*R,C) when s != R, plus one shared tmp (B,R,s,C) sized for the largest layer
  size_t total = 0, tmp = 0;
  for (int l = 0; l < n_layers; ++l) {
    if (h_layers[l].s != R) {
      total += sizeof(float) * static_cast<size_t>(B) * R * R * h_layers[l].C;
      size_t t = sizeof(float) * static_cast<size_t>(B) * R * h_layers[l].s * h_layers[l].C;
      tmp = tmp > t ? tmp : t;
    } else {
      total += sizeof(float) * static_cast<size_t>(B) * R * R * h_layers[l].C;
    }
  }
  return total + tmp;
}

int attn_map_bwd_simt(const skp_attn_layer* h_layers, int n_layers, int B, int T, int R, const int64_t* token_idx,
                      int n_idx, int64_t idx_stride, const float* dmaps_sel, void* workspace, size_t workspace_bytes,
                      cudaStream_t st) {
  SimtLayers L;
  int rc = fill_simt_layers(h_layers, n_layers, B, T, R, &L);
  if (rc != SKP_OK) return rc;
  SKP_REQUIRE(token_idx && dmaps_sel && n_idx >= 1 && n_idx <= L.dmax + 1,
              "attn_map_bwd (SIMT): need 1 <= n_idx <= dim_head+1 (chunk the planes; the op is linear in dmaps)");
  const size_t need = attn_map_bwd_simt_workspace(h_layers, n_layers, B, R);
  if (workspace_bytes < need || !workspace) {
    set_error("attn_map_bwd: workspace %zu B < required %zu B", workspace_bytes, need);
    return SKP_ERR_WORKSPACE;
  }
  char* ws = static_cast<char*>(workspace);
  for (int l = 0; l < n_layers; ++l) {
    SKP_REQUIRE(h_layers[l].dq && h_layers[l].dk, "attn_map_bwd: layer %d has null dq/dk", l);
    L.dq_up[l] = reinterpret_cast<float*>(ws);
    ws += sizeof(float) * static_cast<size_t>(B) * R * R * h_layers[l].C;
  }
  float* tmp = reinterpret_cast<float*>(ws);
  const size_t smem = sizeof(float) * (static_cast<size_t>(2 * SIMT_TP) * (L.dmax + 1) + SIMT_TT * (L.dmax + 1) + SIMT_TP * (T + 1));
  SKP_REQUIRE(smem <= 227 * 1024, "attn_map_bwd (SIMT): T=%d, dim_head=%d need %zu B of shared memory (> 227 KB)", T, L.dmax, smem);
  SKP_CUDA(cudaFuncSetAttribute(attn_map_bwd_simt_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
  SKP_REQUIRE(L.total_heads <= 65535 && B <= 65535, "attn_map_bwd: grid too large");
  dim3 grid(R * R / SIMT_TP, L.total_heads, B);
  attn_map_bwd_simt_kernel<<<(::skp::note_launch(), grid), SIMT_THREADS, smem, st>>>(L, T, R, token_idx, n_idx, idx_stride, dmaps_sel);
  SKP_LAUNCH_CHECK();
  for (int l = 0; l < n_layers; ++l) {
    const skp_attn_layer& a = h_layers[l];
    if (a.s == R) {
      const int64_t n = static_cast<int64_t>(B) * R * R * a.C;
      add_inplace_kernel<<<(::skp::note_launch(), static_cast<unsigned>((n + 255) / 256)), 256, 0, st>>>(a.dq, L.dq_up[l], n);
    } else {
      dim3 g1((a.C + 127) / 128, R * a.s, B), g2((a.C + 127) / 128, a.s * a.s, B);
      SKP_REQUIRE(R * a.s <= 65535, "attn_map_bwd: R*s too large for the adjoint grid");
      bicubic_adjoint_x_kernel<<<(::skp::note_launch(), g1), 128, 0, st>>>(L.dq_up[l], a.s, a.C, R, tmp);
      bicubic_adjoint_y_kernel<<<(::skp::note_launch(), g2), 128, 0, st>>>(tmp, a.s, a.C, R, a.dq);
    }
    SKP_LAUNCH_CHECK();
  }
  return SKP_OK;
}

}  // namespace skp
