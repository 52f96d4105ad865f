// Native-resolution half of the "upsample after the logits" backward (sm_100a).
//
// attn_map_bwd_tc_kernel<.., LR> accumulates dz_low[b*h + head][token][source pixel] = W^T dS (fp32, atomics), the
// gradient w.r.t. the module's native-resolution logits z = scale * q . k.  What is left is two small dense products
// per (image, head):
//     dq[b, sp, head*d + c] += scale * sum_t  dz_low[t, sp] * k[t, head*d + c]
//     dk[t, head*d + c]     += scale * sum_b sum_sp dz_low[t, sp] * q[b, sp, head*d + c]
// (1.7 GFLOP each per SD-2.1 image instead of the 147 GFLOP dQ'/dK products at 128^2), done on tcgen05 with bf16
// operands like the full-resolution gradient products of the staged path:
//  * lr_dz_convert_kernel    : dz_low fp32 -> bf16 in both orientations ([t][sp] and [sp][t], tiled transpose)
//  * lr_heads_transposed_kernel : fp32 (Bx, rows, C) -> bf16 [bx*h + head][d][rows_padded]
//  * lr_grad_gemm_kernel     : out[m, n] += alpha * sum_z sum_k A_z[m, k] B_z[n, k], both operands K-major,
//                              M tile 128, N = dim_head <= 64, fp32 accumulate in TMEM, plain read-modify-write
//                              epilogue (every output element is owned by one CTA)
#include "attn_tc.cuh"

namespace skp {

constexpr int LGG_THREADS = 192;       // warp 0: TMA + MMA issue, warps 2-5: epilogue
constexpr int LGG_BN = 64;             // dim_head columns
constexpr int LGG_A_BYTES = TC_BM * TC_BK * 2;      // 16 KB
constexpr int LGG_B_BYTES = LGG_BN * TC_BK * 2;     // 8 KB
constexpr int LGG_STAGES = 4;
constexpr int LGG_STAGE = LGG_A_BYTES + LGG_B_BYTES;
constexpr int LGG_DYNAMIC = LGG_STAGES * LGG_STAGE + 128 + 1024;

struct LrGradGemmParams {
  CUtensorMap a_map;      // (K, M_rows, Z), box (64, 128, 1)
  CUtensorMap b_map;      // (K, d, Zb), box (64, 64, 1)
  float* out;             // out + zo * out_z + m * ld + head * d + n
  int64_t out_z;          // stride between output batches (0 for a shared k)
  int ld, d, heads, m_valid, K;
  int nsum, a_zstride, b_zstride;   // the sum over z: A batch = za + i * a_zstride, B batch = zb + i * b_zstride
  int b_shared;           // B batch ignores the image (dq product with a shared k)
  int a_mn;               // A is stored [k][m] (m contiguous): a_map = (M, K_rows, Z) with box (64, 64, 1), two 64-wide M slabs per stage
  int split;              // the sum over z is cut into `split` CTAs per output tile (blockIdx.y = zg * split + part); their
                          // epilogues add with vector reductions instead of the single owner's read-modify-write
  float alpha;
};

// dz fp32 [z][T][s2] -> bf16 [z][T][s2] and bf16 [z][s2][Tp] (Tp = T padded to 8, pad columns zero)
// (a hi + lo split of dz was measured: no change of the gradient error, which is set by the bf16 dS / W operands of the
// main kernel, so dz is rounded once)
__global__ void __launch_bounds__(256) lr_dz_convert_kernel(const float* __restrict__ dz, int T, int s2, int Tp, __nv_bfloat16* __restrict__ out,
                                                            __nv_bfloat16* __restrict__ outT) {
  // 64 x 64 tile, 16 x 16 threads: 16-byte loads, 8-byte stores in both orientations (s2 and Tp are multiples of 4)
  __shared__ float tile[64][65];
  const int z = blockIdx.z, t0 = blockIdx.y * 64, c0 = blockIdx.x * 64;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const float* src = dz + static_cast<int64_t>(z) * T * s2;
  auto pack4 = [](float a, float b, float c, float d) {
    const __nv_bfloat162 lo = __floats2bfloat162_rn(a, b), hi = __floats2bfloat162_rn(c, d);
    return make_uint2(*reinterpret_cast<const uint32_t*>(&lo), *reinterpret_cast<const uint32_t*>(&hi));
  };
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int r = ty + 16 * i, t = t0 + r, c = c0 + 4 * tx;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (t < T && c < s2) {
      v = *reinterpret_cast<const float4*>(src + static_cast<int64_t>(t) * s2 + c);
      *reinterpret_cast<uint2*>(out + (static_cast<int64_t>(z) * T + t) * s2 + c) = pack4(v.x, v.y, v.z, v.w);
    }
    tile[r][4 * tx] = v.x; tile[r][4 * tx + 1] = v.y; tile[r][4 * tx + 2] = v.z; tile[r][4 * tx + 3] = v.w;
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int cr = ty + 16 * i, c = c0 + cr, t = t0 + 4 * tx;
    if (c < s2 && t < Tp)
      *reinterpret_cast<uint2*>(outT + (static_cast<int64_t>(z) * s2 + c) * Tp + t) =
          pack4(tile[4 * tx][cr], tile[4 * tx + 1][cr], tile[4 * tx + 2][cr], tile[4 * tx + 3][cr]);
  }
}

// src (Bx, rows, C) fp32 -> mult * src as bf16 [bx*h + head][dd][rows_p] (rows contiguous; pad columns zero)
__global__ void lr_heads_transposed_kernel(const float* __restrict__ src, int rows, int rows_p, int C, int heads, float mult,
                                           __nv_bfloat16* __restrict__ out) {
  __shared__ float tile[32][33];
  const int bx = blockIdx.z, r0 = blockIdx.y * 32, c0 = blockIdx.x * 32;
  const int d = C / heads;
  for (int r = threadIdx.y; r < 32; r += blockDim.y) {
    const int rr = r0 + r, c = c0 + threadIdx.x;
    tile[r][threadIdx.x] = (rr < rows && c < C) ? src[(static_cast<int64_t>(bx) * rows + rr) * C + c] * mult : 0.0f;
  }
  __syncthreads();
  for (int r = threadIdx.y; r < 32; r += blockDim.y) {
    const int c = c0 + r, rr = r0 + threadIdx.x;
    if (c < C && rr < rows_p) {
      const int head = c / d, dd = c - head * d;
      out[((static_cast<int64_t>(bx) * heads + head) * d + dd) * rows_p + rr] = __float2bfloat16_rn(tile[threadIdx.x][r]);
    }
  }
}

__global__ void __launch_bounds__(LGG_THREADS) lr_grad_gemm_kernel(const __grid_constant__ LrGradGemmParams P) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = ptx::smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  uint8_t* base_ptr = smem_raw + (base - raw);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int split = P.split > 1 ? P.split : 1;
  const int mt = blockIdx.x, zg = blockIdx.y / split, part = blockIdx.y - zg * split;   // zg = output batch * heads + head
  const int n0 = blockIdx.z * LGG_BN;                   // first dim_head column of this CTA (dim_head > 64: several N chunks)
  const int zo = zg / P.heads, head = zg - zo * P.heads;
  auto bar_full = [&](int s) { return base + LGG_STAGES * LGG_STAGE + 8u * s; };
  auto bar_empty = [&](int s) { return base + LGG_STAGES * LGG_STAGE + 8u * (LGG_STAGES + s); };
  const uint32_t bar_acc = base + LGG_STAGES * LGG_STAGE + 8u * (2 * LGG_STAGES);
  volatile uint32_t* tmem_ptr = reinterpret_cast<volatile uint32_t*>(base_ptr + LGG_STAGES * LGG_STAGE + 8 * (2 * LGG_STAGES + 1));

  if (threadIdx.x == 0) {
    for (int s = 0; s < LGG_STAGES; ++s) {
      ptx::mbar_init(bar_full(s), 1);
      ptx::mbar_init(bar_empty(s), 1);
    }
    ptx::mbar_init(bar_acc, 1);
    ptx::fence_barrier_init();
  }
  if (warp == 0) {
    ptx::tmem_alloc(ptx::smem_u32(const_cast<uint32_t*>(tmem_ptr)), LGG_BN);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  const int nkb = (P.K + TC_BK - 1) / TC_BK;
  const int i_lo = part * P.nsum / split, i_hi = (part + 1) * P.nsum / split;      // my slice of the sum over z
  const int total = nkb * (i_hi - i_lo);

  if (warp == 0) {
    if (lane == 0) {
      // one thread runs both the TMA ring (LGG_STAGES blocks ahead) and the MMA issue
      // instruction descriptor of ptx::umma_idesc_bf16 (+ bit 15: A is MN-major); MN-major descriptor as in attn_tc_bwd.cu (bw_desc)
      const uint32_t idesc = ptx::umma_idesc_bf16(TC_BM, LGG_BN) | (P.a_mn ? (1u << 15) : 0u);
      auto mn_desc = [](uint32_t addr, uint32_t lbo, uint32_t sbo) {
        uint64_t d = 0;
        d |= static_cast<uint64_t>((addr & 0x3FFFFu) >> 4);
        d |= static_cast<uint64_t>(lbo >> 4) << 16;
        d |= static_cast<uint64_t>(sbo >> 4) << 32;
        d |= static_cast<uint64_t>(1) << 46;
        d |= static_cast<uint64_t>(2) << 61;
        return d;
      };
      auto load = [&](int it) {
        const int i = i_lo + it / nkb, kb = it % nkb;
        const int s = it % LGG_STAGES;
        const int za = (zo * P.heads + head) + i * P.a_zstride;
        const int zb = (P.b_shared ? head : zo * P.heads + head) + i * P.b_zstride;
        ptx::mbar_wait(bar_empty(s), ((it / LGG_STAGES) & 1u) ^ 1u);
        ptx::mbar_arrive_expect_tx(bar_full(s), LGG_STAGE);
        const uint32_t st = base + s * LGG_STAGE;
        if (P.a_mn) {         // [k rows][64 m] slabs: m 0-63, then m 64-127
          ptx::tma_load_3d(st, &P.a_map, bar_full(s), mt * TC_BM, kb * TC_BK, za);
          ptx::tma_load_3d(st + LGG_A_BYTES / 2, &P.a_map, bar_full(s), mt * TC_BM + 64, kb * TC_BK, za);
        } else {
          ptx::tma_load_3d(st, &P.a_map, bar_full(s), kb * TC_BK, mt * TC_BM, za);
        }
        ptx::tma_load_3d(st + LGG_A_BYTES, &P.b_map, bar_full(s), kb * TC_BK, n0, zb);
      };
      const int ahead = total < LGG_STAGES ? total : LGG_STAGES;
      for (int it = 0; it < ahead; ++it) load(it);
      for (int it = 0; it < total; ++it) {
        const int s = it % LGG_STAGES;
        ptx::mbar_wait(bar_full(s), (it / LGG_STAGES) & 1u);
        ptx::tc_fence_after();
        const uint32_t st = base + s * LGG_STAGE;
        const int kb = it % nkb;
        const int nk16 = (min(TC_BK, P.K - kb * TC_BK) + 15) / 16;
        for (int kk = 0; kk < nk16; ++kk)
          ptx::umma_bf16(tmem_base, P.a_mn ? mn_desc(st + kk * 2048, LGG_A_BYTES / 2, 1024) : ptx::umma_desc_sw128(st + kk * 32),
                         ptx::umma_desc_sw128(st + LGG_A_BYTES + kk * 32), idesc, (it | kk) != 0);
        ptx::umma_commit(bar_empty(s));
        if (it + LGG_STAGES < total) load(it + LGG_STAGES);
      }
      ptx::umma_commit(bar_acc);
    }
  } else if (warp >= 2) {
    ptx::mbar_wait_hint(bar_acc, 0u, 1000u);
    ptx::tc_fence_after();
    const int quad = warp & 3;
    const int m = mt * TC_BM + quad * 32 + lane;
    const uint32_t t_lane = tmem_base + (static_cast<uint32_t>(quad * 32) << 16);
    float* o = P.out + static_cast<int64_t>(zo) * P.out_z + static_cast<int64_t>(m) * P.ld + head * P.d + n0;
    const int dn = P.d - n0;                              // columns of this chunk that exist
#pragma unroll 1
    for (int c = 0; c < LGG_BN / 32; ++c) {
      if (c * 32 >= dn) break;
      uint32_t v[32];
      ptx::tmem_ld32(t_lane + c * 32, v);
      ptx::tmem_ld_wait();
      if (m < P.m_valid) {
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          if (c * 32 + q * 4 < dn) {
            float4* p4 = reinterpret_cast<float4*>(o + c * 32 + q * 4);
            if (split > 1) {
              asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p4), "f"(P.alpha * __uint_as_float(v[4 * q])),
                           "f"(P.alpha * __uint_as_float(v[4 * q + 1])), "f"(P.alpha * __uint_as_float(v[4 * q + 2])),
                           "f"(P.alpha * __uint_as_float(v[4 * q + 3])) : "memory");
            } else {
              float4 cur = *p4;
              cur.x = fmaf(P.alpha, __uint_as_float(v[4 * q]), cur.x);
              cur.y = fmaf(P.alpha, __uint_as_float(v[4 * q + 1]), cur.y);
              cur.z = fmaf(P.alpha, __uint_as_float(v[4 * q + 2]), cur.z);
              cur.w = fmaf(P.alpha, __uint_as_float(v[4 * q + 3]), cur.w);
              *p4 = cur;
            }
          }
        }
      }
    }
  }
  __syncwarp();
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, LGG_BN);
  }
}

size_t lr_bwd_post_workspace(const skp_attn_layer& a, int B, int T) {
  const size_t s2 = static_cast<size_t>(a.s) * a.s, Tp = (T + 7) / 8 * 8, Bk = a.k_shared ? 1 : B;
  size_t total = 0;
  total += tc_align_up(static_cast<size_t>(B) * a.heads * T * s2 * 4, 1024);        // dz fp32
  total += tc_align_up(static_cast<size_t>(B) * a.heads * T * s2 * 2, 1024);        // dz bf16 [t][sp]
  total += tc_align_up(static_cast<size_t>(B) * a.heads * s2 * Tp * 2, 1024);       // dz bf16 [sp][t]
  total += tc_align_up(Bk * static_cast<size_t>(a.C) * Tp * 2, 1024);               // k^T
  total += tc_align_up(static_cast<size_t>(B) * a.C * s2 * 2, 1024);                // q^T
  return total;
}

// dz (fp32, [B*h][T][s2], complete) -> accumulate dq and dk of this layer.  `ws` must hold lr_bwd_post_workspace minus the
// dz buffer itself (the four bf16 arrays), 1024-aligned.
// dz_bf16: dz already is bf16 [B*h][T][s2] (accumulated that way by the main kernel): no conversion, the dq product reads it as
// an MN-major A operand.
int lr_bwd_post(const skp_attn_layer& a, int B, int T, const float* dz, char* ws, cudaStream_t st, bool dz_bf16) {
  const int s2 = a.s * a.s, d = a.C / a.heads, Tp = (T + 7) / 8 * 8, Bk = a.k_shared ? 1 : B, Z = B * a.heads;
  auto take = [&](size_t bytes) {
    __nv_bfloat16* p = reinterpret_cast<__nv_bfloat16*>(ws);
    ws += tc_align_up(bytes, 1024);
    return p;
  };
  __nv_bfloat16* dzb = take(static_cast<size_t>(Z) * T * s2 * 2);
  __nv_bfloat16* dzT = take(static_cast<size_t>(Z) * s2 * Tp * 2);
  __nv_bfloat16* kT = take(static_cast<size_t>(Bk) * a.C * Tp * 2);
  __nv_bfloat16* qT = take(static_cast<size_t>(B) * a.C * s2 * 2);
  const dim3 tb(32, 8);
  if (dz_bf16) dzb = reinterpret_cast<__nv_bfloat16*>(const_cast<float*>(dz));
  else lr_dz_convert_kernel<<<(::skp::note_launch(), dim3((s2 + 63) / 64, (Tp + 63) / 64, Z)), 256, 0, st>>>(dz, T, s2, Tp, dzb, dzT);
  lr_heads_transposed_kernel<<<(::skp::note_launch(), dim3((a.C + 31) / 32, (Tp + 31) / 32, Bk)), tb, 0, st>>>(a.k, T, Tp, a.C, a.heads, 1.0f, kT);
  lr_heads_transposed_kernel<<<(::skp::note_launch(), dim3((a.C + 31) / 32, (s2 + 31) / 32, B)), tb, 0, st>>>(a.q, s2, s2, a.C, a.heads, 1.0f, qT);
  SKP_LAUNCH_CHECK();
  SKP_CUDA(cudaFuncSetAttribute(lr_grad_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, LGG_DYNAMIC));
  // dq[b, sp, head*d + c] += scale * sum_t dzT[z][sp][t] * kT[kz][c][t]
  {
    LrGradGemmParams G;
    memset(&G, 0, sizeof(G));
    int rc = dz_bf16 ? tc_encode_map(&G.a_map, dzb, s2, T, Z, 64) : tc_encode_map(&G.a_map, dzT, Tp, s2, Z, TC_BM);
    G.a_mn = dz_bf16 ? 1 : 0;
    if (rc == SKP_OK) rc = tc_encode_map(&G.b_map, kT, Tp, d, Bk * a.heads, LGG_BN);
    if (rc != SKP_OK) return rc;
    G.out = a.dq; G.out_z = static_cast<int64_t>(s2) * a.C; G.ld = a.C; G.d = d; G.heads = a.heads; G.m_valid = s2; G.K = Tp;
    G.nsum = 1; G.a_zstride = 0; G.b_zstride = 0; G.b_shared = a.k_shared; G.alpha = a.scale;
    lr_grad_gemm_kernel<<<(::skp::note_launch(), dim3((s2 + TC_BM - 1) / TC_BM, Z, (d + LGG_BN - 1) / LGG_BN)), LGG_THREADS, LGG_DYNAMIC, st>>>(G);
    SKP_LAUNCH_CHECK();
  }
  // dk[bk, t, head*d + c] += scale * sum_b sum_sp dzb[z][t][sp] * qT[z][c][sp]
  {
    LrGradGemmParams G;
    memset(&G, 0, sizeof(G));
    int rc = tc_encode_map(&G.a_map, dzb, s2, T, Z, TC_BM);
    if (rc == SKP_OK) rc = tc_encode_map(&G.b_map, qT, s2, d, Z, LGG_BN);
    if (rc != SKP_OK) return rc;
    G.out = a.dk; G.out_z = static_cast<int64_t>(T) * a.C; G.ld = a.C; G.d = d; G.heads = a.heads; G.m_valid = T; G.K = s2;
    G.nsum = a.k_shared ? B : 1; G.a_zstride = a.heads; G.b_zstride = a.heads; G.b_shared = 0; G.alpha = a.scale;
    // a shared k sums over the images: with one CTA per (token tile, head) that is 80 CTAs walking 8 images each; up to four
    // CTAs share the sum
    G.split = G.nsum >= 4 ? 4 : (G.nsum >= 2 ? 2 : 1);
    lr_grad_gemm_kernel<<<(::skp::note_launch(), dim3((T + TC_BM - 1) / TC_BM, Bk * a.heads * G.split, (d + LGG_BN - 1) / LGG_BN)), LGG_THREADS, LGG_DYNAMIC, st>>>(G);
    SKP_LAUNCH_CHECK();
  }
  return SKP_OK;
}

}  // namespace skp
