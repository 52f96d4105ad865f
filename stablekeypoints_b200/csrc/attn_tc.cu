// Fused cross-attention-map forward on tcgen05 tensor cores (sm_100a).
//
//   maps[b, t, p] = sum_l w_l * mean_head softmax_t( scale_l * q'_l[b,head,p,:] . k_l[head,t,:] )
//
// replaces the captured branch of the reference's patched CrossAttention.forward
// (unsupervised_keypoints/ptp_utils.py:508-538) and collect_maps (optimize.py:27-79).
//
// Design (DESIGN.md section 4 A has the long version and the measurements):
//  * operands.  Default ("upsample after the logits", template LR, attn_tc_lr.cu): A = the bicubic interpolation table of
//    the pixel tile (resident in shared memory: it depends on the source size only), B = the tile's window of
//    NATIVE-resolution logits s_low = scale*log2(e) * q.k (bf16 hi + lo), so S_tile = W_tile . s_low^T needs K = 64 / 128
//    taps whatever dim_head is.  Staged form (wide windows): q bicubic-upsampled to R x R by prep kernels, K = dim_head.
//    bf16x3 (hi*hi + hi*lo + lo*hi, fp32 accumulate) is the fp32-grade "parity" arithmetic; plain bf16 is the fast mode.
//  * main kernel: one CTA per (image, 128 consecutive feature pixels, 256-token slice); a cluster of NSPLIT CTAs covers all
//    tokens of the same pixels.  Warp 0 = TMA producer (128B-swizzled tiles, mbarrier ring), warp 1 = tcgen05.mma issuer
//    (warp-uniform, elected issue; M=128, N=256, K=16, accumulators in TMEM, two 256-column buffers ping-pong across the
//    (layer, head) units), 16 softmax warps: thread (row, 64-token quarter) reads its logits in four pipelined 16-column
//    tcgen05.ld chunks, exponentiates (constant shift per (image, head) when the logit range is bounded, else row max),
//    parks the exponentials back in TMEM, exchanges (max, sum) with the other quarters and the other CTA of the cluster
//    through (distributed) shared memory + mbarriers, and one unit later accumulates weight * P into 64 fp32 registers.
//    SELECT (mode C): the few selected tokens of a thread's columns are a compact list, their exponentials live in registers,
//    nothing is parked.  The (h, R^2, T) probability tensor never exists; the output is written once.
#include <cstdlib>

#include "attn_tc.cuh"

#ifndef SKP_FW_XLAG
#define SKP_FW_XLAG 1          // 2: SELECT-mode statistics lag of two units (see XLAG in the kernel)
#endif

namespace skp {

struct TcParams {
  TcOperands ops;
  int T, R2, n_idx;
  int64_t idx_stride;
  const int64_t* token_idx;
  float* maps;
  float* row_stats;   // optional (B, total_heads, R2, 2): (row max, sum exp2) in log2 units
};

// ------------------------------------------------------------------------------------------
// prep: q (B, s*s, C) fp32 -> bicubic to R x R, * scale*log2e, split to bf16 hi (+lo), laid out
// [b*h + head][pixel][d]
// ------------------------------------------------------------------------------------------
template <int NPROD>
__global__ void tc_prep_q_kernel(const float* __restrict__ q, int s, int C, int heads, int R, float mult,
                                 __nv_bfloat16* __restrict__ hi, __nv_bfloat16* __restrict__ lo) {
  const int p = blockIdx.x, b = blockIdx.y;
  const int oy = p / R, ox = p % R;
  const int d = C / heads, R2 = R * R;
  int iy, ix;
  float wy[4], wx[4];
  int yy[4], xx[4];
  if (s != R) {
    const float ratio = static_cast<float>(s) / R;
    bicubic_coeffs(oy, ratio, &iy, wy);
    bicubic_coeffs(ox, ratio, &ix, wx);
#pragma unroll
    for (int a = 0; a < 4; ++a) {
      yy[a] = min(max(iy - 1 + a, 0), s - 1);
      xx[a] = min(max(ix - 1 + a, 0), s - 1);
    }
  }
  const float* src = q + static_cast<int64_t>(b) * s * s * C;
  for (int cp = threadIdx.x; cp < C / 2; cp += blockDim.x) {
    const int c = cp * 2;
    float2 v;
    if (s == R) {
      v = *reinterpret_cast<const float2*>(src + static_cast<int64_t>(p) * C + c);
    } else {
      v = make_float2(0.f, 0.f);
#pragma unroll
      for (int a = 0; a < 4; ++a) {
        float2 row = make_float2(0.f, 0.f);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float2 t = __ldg(reinterpret_cast<const float2*>(src + static_cast<int64_t>(yy[a] * s + xx[e]) * C + c));
          row.x += wx[e] * t.x;
          row.y += wx[e] * t.y;
        }
        v.x += wy[a] * row.x;
        v.y += wy[a] * row.y;
      }
    }
    v.x *= mult;
    v.y *= mult;
    const int head = c / d, dd = c % d;
    const int64_t o = ((static_cast<int64_t>(b) * heads + head) * R2 + p) * d + dd;
    const __nv_bfloat16 hx = __float2bfloat16_rn(v.x), hy = __float2bfloat16_rn(v.y);
    *reinterpret_cast<__nv_bfloat162*>(hi + o) = __halves2bfloat162(hx, hy);
    if (NPROD == 3) {
      *reinterpret_cast<__nv_bfloat162*>(lo + o) =
          __halves2bfloat162(__float2bfloat16_rn(v.x - __bfloat162float(hx)), __float2bfloat16_rn(v.y - __bfloat162float(hy)));
    }
  }
}

// Row-wise separable version of the above for s != R: one CTA = (output row y, image b, 256-channel
// chunk).  Phase 1: the 4 source rows are combined with the y weights into shared memory
// (s x 256 fp32); phase 2: every output pixel of the row is 4 shared-memory taps.  ~4.5 FMA per
// output instead of 20, all global reads and writes coalesced (bf16x2 stores, 128 B per warp).
constexpr int PREP_CH = 256;     // channels per CTA (128 threads x 2)
template <int NPROD>
__global__ void __launch_bounds__(128)
tc_prep_q_rows_kernel(const float* __restrict__ q, int s, int C, int heads, int R, float mult,
                      __nv_bfloat16* __restrict__ hi, __nv_bfloat16* __restrict__ lo) {
  extern __shared__ float yc[];                 // [s][PREP_CH]
  __shared__ float wxs[256 * 4];
  __shared__ int ixs[256];
  const int oy = blockIdx.x, b = blockIdx.y;
  const int c = blockIdx.z * PREP_CH + threadIdx.x * 2;
  const int d = C / heads, R2 = R * R;
  const float ratio = static_cast<float>(s) / R;
  for (int x = threadIdx.x; x < R; x += blockDim.x) {
    int ix;
    float w[4];
    bicubic_coeffs(x, ratio, &ix, w);
    ixs[x] = ix;
    wxs[4 * x] = w[0]; wxs[4 * x + 1] = w[1]; wxs[4 * x + 2] = w[2]; wxs[4 * x + 3] = w[3];
  }
  if (c < C) {
    int iy;
    float wy[4];
    bicubic_coeffs(oy, ratio, &iy, wy);
    const float* src = q + static_cast<int64_t>(b) * s * s * C + c;
    for (int sx = 0; sx < s; ++sx) {
      float2 acc = make_float2(0.f, 0.f);
#pragma unroll
      for (int a = 0; a < 4; ++a) {
        const int yy = min(max(iy - 1 + a, 0), s - 1);
        const float2 t = __ldg(reinterpret_cast<const float2*>(src + static_cast<int64_t>(yy * s + sx) * C));
        acc.x = fmaf(wy[a], t.x, acc.x);
        acc.y = fmaf(wy[a], t.y, acc.y);
      }
      *reinterpret_cast<float2*>(yc + sx * PREP_CH + threadIdx.x * 2) = make_float2(acc.x * mult, acc.y * mult);
    }
  }
  __syncthreads();
  if (c >= C) return;
  const int head = c / d, dd = c % d;
  const int64_t o0 = ((static_cast<int64_t>(b) * heads + head) * R2 + static_cast<int64_t>(oy) * R) * d + dd;
  for (int x = 0; x < R; ++x) {
    const int ix = ixs[x];
    float2 v = make_float2(0.f, 0.f);
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int xx = min(max(ix - 1 + e, 0), s - 1);
      const float w = wxs[4 * x + e];
      const float2 t = *reinterpret_cast<const float2*>(yc + xx * PREP_CH + threadIdx.x * 2);
      v.x = fmaf(w, t.x, v.x);
      v.y = fmaf(w, t.y, v.y);
    }
    const int64_t o = o0 + static_cast<int64_t>(x) * d;
    const __nv_bfloat16 hx = __float2bfloat16_rn(v.x), hy = __float2bfloat16_rn(v.y);
    *reinterpret_cast<__nv_bfloat162*>(hi + o) = __halves2bfloat162(hx, hy);
    if (NPROD == 3) {
      *reinterpret_cast<__nv_bfloat162*>(lo + o) =
          __halves2bfloat162(__float2bfloat16_rn(v.x - __bfloat162float(hx)), __float2bfloat16_rn(v.y - __bfloat162float(hy)));
    }
  }
}

// k (Bk, T, C) fp32 -> bf16 hi (+lo) laid out [bk*h + head][token][d]
template <int NPROD>
__global__ void tc_prep_k_kernel(const float* __restrict__ k, int T, int C, int heads, __nv_bfloat16* __restrict__ hi,
                                 __nv_bfloat16* __restrict__ lo) {
  const int t = blockIdx.x, bk = blockIdx.y;
  const int d = C / heads;
  const float* src = k + (static_cast<int64_t>(bk) * T + t) * C;
  for (int cp = threadIdx.x; cp < C / 2; cp += blockDim.x) {
    const int c = cp * 2;
    const float2 v = *reinterpret_cast<const float2*>(src + c);
    const int head = c / d, dd = c % d;
    const int64_t o = ((static_cast<int64_t>(bk) * heads + head) * T + t) * d + dd;
    const __nv_bfloat16 hx = __float2bfloat16_rn(v.x), hy = __float2bfloat16_rn(v.y);
    *reinterpret_cast<__nv_bfloat162*>(hi + o) = __halves2bfloat162(hx, hy);
    if (NPROD == 3) {
      *reinterpret_cast<__nv_bfloat162*>(lo + o) =
          __halves2bfloat162(__float2bfloat16_rn(v.x - __bfloat162float(hx)), __float2bfloat16_rn(v.y - __bfloat162float(hy)));
    }
  }
}

}  // namespace skp

namespace skp {

// shared-memory carve-up (offsets from a 1024-byte aligned base)
template <int NPROD, int BN = TC_BN, bool LR = false>
struct TcSmem {
  // BN = 256: one CTA per SM, 192 KB of operand stages.  BN = 128: TWO CTAs per SM (each 256 TMEM columns, <= 113 KB of
  // shared memory): one stage -- the other CTA of the SM fills the bubbles a single stage leaves.
  // LR ("upsample after the logits"): the A operand is the interpolation table of the CTA's pixel tile, which depends on
  // the layer's source size only: it stays RESIDENT (up to 128 taps = two 64-tap blocks, hi + lo = 64 KB) and is reloaded
  // when the source size changes (once per CTA for the SD capture sets) instead of once per (layer, head) unit; the ring
  // then only carries the B operand (the window of native-resolution logits).  The kernel is co-limited by the L2 -> SM
  // operand traffic (13.8 GB per 8-image call with A in the ring = 5.6 TB/s at 2.44 ms): 96 -> 64 KB per unit.
  static constexpr int kBBytes = BN * TC_BK * 2;
  static constexpr int kAResident = LR ? 4 * TC_A_BYTES : 0;           // [k-block 0: hi, lo][k-block 1: hi, lo]
  static constexpr int kStages = BN == 128 ? 1 : (NPROD == 3 ? 2 : 4);
  static constexpr int kStageBytes = (NPROD == 3 ? 2 : 1) * ((LR ? 0 : TC_A_BYTES) + kBBytes);
  static constexpr int kRingOff = kAResident;
  static constexpr int kTiles = kAResident + kStages * kStageBytes;    // 192 KB (BN = 256); 64 / 96 KB (BN = 128)
  static constexpr int kBarOff = kTiles;                               // full[], empty[], tfull[2][2], tempty[2][2], xfull[2], afull
  static constexpr int kStatBufs = (BN == 128 || SKP_FW_XLAG < 2) ? 2 : 4;                  // ring of row-statistics slots / xfull barriers (see FW lag below)
  static constexpr int kNumBars = 2 * kStages + 9 + kStatBufs;
  static constexpr int kTmemPtrOff = kBarOff + 8 * kNumBars;
  static constexpr int kTokOff = kTmemPtrOff + 8;                      // int tok[TC_MAX_SEL]
  static constexpr int kAmaxOff = kTokOff + 4 * TC_MAX_SEL;            // float amax[FW_MAX_UNITS]: max |s_low| of every unit's (image, head)
  static constexpr int kStatsOff = kAmaxOff + 4 * 128;                 // float2 stats[kStatBufs][8 part][128 row]
  static constexpr int kTotal = kStatsOff + kStatBufs * 8 * TC_BM * 8;
  // slack for the manual 1024 B alignment of the tiles.  Two CTAs per SM leave 115712 B each: the BN = 128 / LR layout
  // (115440 B) keeps 256 B of slack -- the dynamic window starts 1 KB-aligned on sm_100 and the kernel traps if not
  static constexpr int kSlack = (BN == 128 && LR) ? 256 : 1024;
  static constexpr int kDynamic = kTotal + kSlack;
  static_assert(kDynamic <= 232448, "over the 227 KB of dynamic shared memory per CTA");
};

// Phase profiler of the softmax warps (diagnostic builds only: nvcc -DSKP_FWD_PROF): cycles per phase averaged
// over the units of one CTA, printed by a few warps of blocks (0..1, 0).
#ifdef SKP_FWD_PROF
#define FW_PROF_DECL long long pf_t = clock64(), pf_acc[12] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
#define FW_PROF(i) { const long long pf_n = clock64(); pf_acc[i] += pf_n - pf_t; pf_t = pf_n; }
#else
#define FW_PROF_DECL
#define FW_PROF(i)
#endif

// 2^d for an integer-valued d <= 0 (or -inf), flushed to 0 below 2^-126, without MUFU: the integer lands in the
// low mantissa bits of d + 1.5 * 2^23 and is shifted into the exponent field.
__device__ __forceinline__ float pow2_int(float d) {
  const int bits = __float_as_int(fmaxf(d, -127.0f) + 12582912.0f);
  return __int_as_float((bits << 23) + 0x3f800000);
}

// (A degree-5 polynomial exp2 on the FMA pipe for every fourth exponential was measured neutral in round 1 and slower in round 2 --
// twelve instructions against three in an issue-bound loop -- and is gone; DESIGN.md section 4 A.)

// 20 warps = 5 warpgroups: warpgroup 0 = {TMA producer, MMA issuer, 2 idle warps} gives registers back
// (setmaxnreg.dec), warpgroups 1-4 = 16 softmax warps take them (setmaxnreg.inc): 64 fp32 accumulators + one
// 32-column TMEM chunk per thread.  The softmax side is latency bound (dependent FMNMX / FADD chains, MUFU and
// tcgen05.ld round trips), so it runs four warps per scheduler rather than two wide ones.
//
// BN = tokens per CTA.  256 (default): one CTA per SM, a cluster of two covers T <= 512.  128 (SKP_FWD_BN=128, A/B only):
// the same work as TWO CTAs per SM (cluster of four), written to test whether a second resident CTA, out of phase with
// the first, would fill the MUFU pipe while the first is in its latency-bound phases.  Measured on B200 (round 2): it does
// NOT -- 2.91 ms against 2.63 ms per 8-image call, and the same holds for fully independent CTAs (T = 128, no cluster:
// 0.77 ms = 3.1 ms per 512 tokens) with or without a half-unit start skew between the CTAs of an SM.  Each softmax warp
// needs ~5 k cycles per unit for its 64 columns whatever else is resident (a dependent-latency chain: 6 TMEM round trips,
// the statistics exchange, 64-element FMNMX / FADD / FFMA chains), so throughput = resident warps / that latency, and the
// resident warps are capped by the register file (128 x 256 fp32 accumulators per SM = half of it).  An M=128 K=16 SS-mode
// MMA takes 64 cycles at N = 128 and 128 at N = 256 (scripts/mma_rate_probe.cu): the tensor time per token is the same.
constexpr int FW_SOFTMAX_WARP0 = 4;
constexpr int FW_QCOLS = 64;                                  // tokens per softmax thread
constexpr int FW_MAX_PARTS = 8;                               // 64-token column parts of a cluster (2 x 4 or 4 x 2)
#ifndef SKP_FW_NH
#define SKP_FW_NH 1
#endif
constexpr int FW_NH = SKP_FW_NH;                                      // column groups a TMEM buffer is handed over in (1 or 2)
// LR: with [lo, hi] the range of the native-resolution logits of a unit's (image, head), every upsampled logit lies within
// c +- 0.945 (hi - lo), c = (hi + lo) / 2.  When that half-width is <= FW_FAST_HALFWIDTH (log2 units), exp2(S - c) can
// neither overflow (512 x 2^118 < 2^128) nor flush a single term (2^-118 is a normal float): the softmax is shift
// invariant, so the row-max pass (a TMEM round trip per chunk + 64 compares per thread + the exchange of maxima) is skipped
// and the unit's statistics are (m = round(c), l = sum exp2(S - m)).  Wider ranges take the max-subtracted path.
constexpr float FW_FAST_HALFWIDTH = 117.0f;
constexpr int FW_MAX_UNITS = 128;                             // units (heads of all captured layers) the flag table holds
template <int BN>
struct FwCfg {
  static constexpr int kParts = BN / FW_QCOLS;                // column parts per CTA: 4 or 2
  static constexpr int kSoftmaxThreads = 128 * kParts;        // 16 or 8 warps
  static constexpr int kThreads = 32 * FW_SOFTMAX_WARP0 + kSoftmaxThreads;   // 640 or 384
  static constexpr int kMinCtas = BN == 128 ? 2 : 1;
  // setmaxnreg.inc only draws from what .dec released: 128*32 + 512*112 == 640*96 (BN 256); 128*32 + 256*104 == 384*80 (BN 128)
  static constexpr int kSoftmaxRegs = BN == 128 ? 104 : 112;
};
template <int NPROD, int NSPLIT, bool SELECT, int BN, bool LR>
__global__ void __launch_bounds__(FwCfg<BN>::kThreads, FwCfg<BN>::kMinCtas)
attn_map_fwd_tc_kernel(const __grid_constant__ TcParams P) {
  using SM = TcSmem<NPROD, BN, LR>;
  static_assert(!LR || NPROD == 3, "the upsample-after-logits operands are hi + lo pairs");
  constexpr int FW_PARTS = FwCfg<BN>::kParts;
  constexpr int FW_SOFTMAX_THREADS = FwCfg<BN>::kSoftmaxThreads;
  constexpr int TC_BNk = BN;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = ptx::smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  uint8_t* base_ptr = smem_raw + (base - raw);
  if (base - raw > static_cast<uint32_t>(SM::kSlack)) __trap();

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int warp_u = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);   // the same value, warp-uniform for the compiler (MMA role)
  const uint32_t cta_rank = NSPLIT > 1 ? ptx::cluster_ctarank() : 0u;
  const int tile = blockIdx.x / NSPLIT;
  const int b = blockIdx.y;
  const int p0 = tile * TC_BM;
  const int tok0 = static_cast<int>(cta_rank) * TC_BNk;

  auto bar_full = [&](int s) { return base + SM::kBarOff + 8u * s; };
  auto bar_empty = [&](int s) { return base + SM::kBarOff + 8u * (SM::kStages + s); };
  // TMEM buffers are handed over in column halves: tfull/tempty[buffer][half]
  auto bar_tfull = [&](int i, int hf) { return base + SM::kBarOff + 8u * (2 * SM::kStages + 2 * i + hf); };
  auto bar_tempty = [&](int i, int hf) { return base + SM::kBarOff + 8u * (2 * SM::kStages + 4 + 2 * i + hf); };
  const uint32_t bar_afull = base + SM::kBarOff + 8u * (2 * SM::kStages + 8);
  auto bar_xfull = [&](int i) { return base + SM::kBarOff + 8u * (2 * SM::kStages + 9 + i); };
  volatile uint32_t* tmem_ptr = reinterpret_cast<volatile uint32_t*>(base_ptr + SM::kTmemPtrOff);
  int* tok_s = reinterpret_cast<int*>(base_ptr + SM::kTokOff);
  float* amax_s = reinterpret_cast<float*>(base_ptr + SM::kAmaxOff);
  float2* stats = reinterpret_cast<float2*>(base_ptr + SM::kStatsOff);

  // ---- one-time setup ---------------------------------------------------------------------
  if (threadIdx.x == 0) {
    for (int s = 0; s < SM::kStages; ++s) {
      ptx::mbar_init(bar_full(s), 1);
      ptx::mbar_init(bar_empty(s), 1);
    }
    for (int i = 0; i < 2; ++i) {
      for (int hf = 0; hf < 2; ++hf) {
        ptx::mbar_init(bar_tfull(i, hf), 1);
        ptx::mbar_init(bar_tempty(i, hf), FW_SOFTMAX_THREADS / FW_NH);
      }
    }
    for (int i = 0; i < SM::kStatBufs; ++i) ptx::mbar_init(bar_xfull(i), FW_SOFTMAX_THREADS);
    ptx::mbar_init(bar_afull, 1);
    ptx::fence_barrier_init();
  }
  if (LR && threadIdx.x < FW_MAX_UNITS) {
    // unit u = global head index -> (layer, head): the constant shift m of the bounded path, or +inf (max-subtracted
    // path) for units past the table, without a recorded range, or whose range is too wide
    float v = INFINITY;
    const int u = threadIdx.x;
    if (u < P.ops.total_heads && P.ops.total_heads <= FW_MAX_UNITS) {
      int l = 0;
      while (l + 1 < P.ops.n_layers && u >= P.ops.head_base[l + 1]) ++l;
      if (P.ops.absmax[l]) {
        const int* key = P.ops.absmax[l] + 2 * (blockIdx.y * P.ops.heads[l] + (u - P.ops.head_base[l]));
        const float hi = lr_range_value(key[0]), lo = -lr_range_value(key[1]);
        if (hi >= lo && 0.945f * (hi - lo) + 1.0f <= FW_FAST_HALFWIDTH) v = rintf(0.5f * (hi + lo));   // integer-valued like the row maxima
      }
    }
    amax_s[u] = v;
  }
  if (SELECT && threadIdx.x < TC_MAX_SEL) tok_s[threadIdx.x] = threadIdx.x < P.n_idx ? static_cast<int>(P.token_idx[blockIdx.y * P.idx_stride + threadIdx.x]) : -1;
  if (warp == 1) {
    ptx::tmem_alloc(ptx::smem_u32(const_cast<uint32_t*>(tmem_ptr)), 2 * BN);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (NSPLIT > 1) ptx::cluster_sync();   // peers' barriers are initialised before any remote arrive
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp < FW_SOFTMAX_WARP0) {
    ptx::setmaxnreg_dec<32>();
    if (warp == 0) {
    // =========================== TMA producer (one lane) ===================================
    if (lane == 0) {
      for (int l = 0; l < P.ops.n_layers; ++l) {
        ptx::prefetch_tensormap(&P.ops.q_map[l][0]);
        ptx::prefetch_tensormap(&P.ops.k_map[l][0]);
        if (NPROD == 3) {
          ptx::prefetch_tensormap(&P.ops.q_map[l][1]);
          ptx::prefetch_tensormap(&P.ops.k_map[l][1]);
        }
      }
      uint32_t it = 0;
      int a_key = -1;                     // LR: source size of the interpolation table resident in the A region
      for (int l = 0; l < P.ops.n_layers; ++l) {
        const int h = P.ops.heads[l], d = P.ops.d[l];
        const int nkb = (d + TC_BK - 1) / TC_BK;
        // LR: the window of native-resolution logits the tile interpolates from starts at this source pixel
        const int b_k0 = LR ? lr_window_row0(tile, P.ops.win_s[l], P.ops.R, P.ops.win_nr[l]) * P.ops.win_s[l] : 0;
        if (LR && P.ops.win_s[l] != a_key) {
          // new interpolation table: every MMA issued so far read the old one -- wait for the last ring slot's release
          // (tcgen05.commit covers all earlier MMAs of the issuing thread), then load all k-blocks on their own barrier
          if (it > 0) ptx::mbar_wait_hint(bar_empty((it - 1) % SM::kStages), ((it - 1) / SM::kStages) & 1u, 2000u);
          ptx::mbar_arrive_expect_tx(bar_afull, nkb * 2 * TC_A_BYTES);
          for (int kb = 0; kb < nkb; ++kb) {
            ptx::tma_load_3d(base + kb * 2 * TC_A_BYTES, &P.ops.q_map[l][0], bar_afull, kb * TC_BK, 0, tile);
            ptx::tma_load_3d(base + kb * 2 * TC_A_BYTES + TC_A_BYTES, &P.ops.q_map[l][1], bar_afull, kb * TC_BK, 0, tile);
          }
          a_key = P.ops.win_s[l];
        }
        for (int head = 0; head < h; ++head) {
          const int qz = b * h + head;
          const int kz = (P.ops.k_shared[l] ? 0 : b * h) + head;
          // operand coordinates: (k offset, row, batch).  Staged: A = q' tile, B = k; LR: B = the logit window of (image, head)
          const int b_z = LR ? qz : kz;
          for (int kb = 0; kb < nkb; ++kb, ++it) {
            const int s = it % SM::kStages;
            const uint32_t ph = (it / SM::kStages) & 1u;
            ptx::mbar_wait_hint(bar_empty(s), ph ^ 1u, 2000u);
            ptx::mbar_arrive_expect_tx(bar_full(s), SM::kStageBytes);
            const uint32_t st = base + SM::kRingOff + s * SM::kStageBytes;
            if (LR) {
              ptx::tma_load_3d(st, &P.ops.k_map[l][0], bar_full(s), b_k0 + kb * TC_BK, tok0, b_z);
              ptx::tma_load_3d(st + SM::kBBytes, &P.ops.k_map[l][1], bar_full(s), b_k0 + kb * TC_BK, tok0, b_z);
            } else {
              ptx::tma_load_3d(st, &P.ops.q_map[l][0], bar_full(s), kb * TC_BK, p0, qz);
              ptx::tma_load_3d(st + TC_A_BYTES, &P.ops.k_map[l][0], bar_full(s), kb * TC_BK, tok0, b_z);
              if (NPROD == 3) {
                ptx::tma_load_3d(st + TC_A_BYTES + SM::kBBytes, &P.ops.q_map[l][1], bar_full(s), kb * TC_BK, p0, qz);
                ptx::tma_load_3d(st + 2 * TC_A_BYTES + SM::kBBytes, &P.ops.k_map[l][1], bar_full(s), kb * TC_BK, tok0, b_z);
              }
            }
          }
        }
      }
    }
  } else if (warp_u == 1) {
    // =========================== MMA issuer (whole warp, elected issue: ptx::umma_bf16_elect) ==
    {
      // FW_NH = 2 hands each buffer over as two N = 128 column halves with their own barriers (two softmax
      // groups half a unit out of phase).  Round 1 measured ~130 cycles per MMA whether N was 128 or 256 and blamed
      // the tensor pipe; scripts/mma_small_probe.cu says 64 vs 128 cycles -- the 130 were the single-lane issuer's
      // ELECT / R2UR waterfall around every instruction, gone with the warp-uniform issue below.
      constexpr uint32_t idesc = ptx::umma_idesc_bf16(TC_BM, TC_BNk / FW_NH);
      constexpr uint32_t kHalfB = (TC_BNk / FW_NH) * 128;     // byte offset of the second group's first token row in a K tile
      uint32_t it = 0, u = 0, a_cnt = 0;
      int a_key = -1;
#ifdef SKP_FWD_PROF
      long long mp_full = 0, mp_tempty = 0, mp_issue = 0, mp_t0 = clock64(), mp_t;
#endif
      for (int l = 0; l < P.ops.n_layers; ++l) {
        const int h = P.ops.heads[l], d = P.ops.d[l];
        const int nkb = (d + TC_BK - 1) / TC_BK;
        if (LR && P.ops.win_s[l] != a_key) {      // the resident interpolation table of this source size has landed
          ptx::mbar_wait_hint(bar_afull, a_cnt & 1u, 2000u);
          ptx::tc_fence_after();
          ++a_cnt;
          a_key = P.ops.win_s[l];
        }
        for (int head = 0; head < h; ++head, ++u) {
          const uint32_t buf = u & 1u;
          for (int kb = 0; kb < nkb; ++kb, ++it) {
            const int s = it % SM::kStages;
#ifdef SKP_FWD_PROF
            mp_t = clock64();
#endif
            ptx::mbar_wait_hint(bar_full(s), (it / SM::kStages) & 1u, 2000u);
            ptx::tc_fence_after();
#ifdef SKP_FWD_PROF
            mp_full += clock64() - mp_t;
#endif
            const uint32_t st = base + SM::kRingOff + s * SM::kStageBytes;
            // operand tiles of this k-block: LR: A resident, ring = [B hi | B lo]; staged: ring = [A hi | B hi | A lo | B lo]
            const uint32_t a_hi_s = LR ? base + kb * 2 * TC_A_BYTES : st;
            const uint32_t a_lo_s = LR ? a_hi_s + TC_A_BYTES : st + TC_A_BYTES + SM::kBBytes;
            const uint32_t b_hi_s = LR ? st : st + TC_A_BYTES;
            const uint32_t b_lo_s = LR ? st + SM::kBBytes : st + 2 * TC_A_BYTES + SM::kBBytes;
            const int nk16 = (min(TC_BK, d - kb * TC_BK) + 15) / 16;
            for (int hf = 0; hf < FW_NH; ++hf) {
              if (kb == 0) {
#ifdef SKP_FWD_PROF
                mp_t = clock64();
#endif
                ptx::mbar_wait_hint(bar_tempty(buf, hf), ((u >> 1) & 1u) ^ 1u, 2000u);   // softmax drained this half buffer
                ptx::tc_fence_after();
#ifdef SKP_FWD_PROF
                mp_tempty += clock64() - mp_t;
                mp_t = clock64();
#endif
              }
              const uint32_t d_tmem = tmem_base + buf * TC_BNk + hf * (TC_BNk / FW_NH);
              for (int kk = 0; kk < nk16; ++kk) {
                const uint64_t a_hi = ptx::umma_desc_sw128(a_hi_s + kk * 32);
                const uint64_t b_hi = ptx::umma_desc_sw128(b_hi_s + hf * kHalfB + kk * 32);
                const uint32_t first = (kb | kk) != 0;
                if (NPROD == 3) {
                  const uint64_t a_lo = ptx::umma_desc_sw128(a_lo_s + kk * 32);
                  const uint64_t b_lo = ptx::umma_desc_sw128(b_lo_s + hf * kHalfB + kk * 32);
                  ptx::umma_bf16_elect(d_tmem, a_hi, b_lo, idesc, first);
                  ptx::umma_bf16_elect(d_tmem, a_lo, b_hi, idesc, 1u);
                  ptx::umma_bf16_elect(d_tmem, a_hi, b_hi, idesc, 1u);
                } else {
                  ptx::umma_bf16_elect(d_tmem, a_hi, b_hi, idesc, first);
                }
              }
              if (kb == nkb - 1) ptx::umma_commit_elect(bar_tfull(buf, hf));   // this half's logits are in TMEM
            }
            ptx::umma_commit_elect(bar_empty(s));      // smem slot free once these MMAs retire
#ifdef SKP_FWD_PROF
            mp_issue += clock64() - mp_t;
#endif
          }
        }
      }
#ifdef SKP_FWD_PROF
      if (blockIdx.x < 2 && blockIdx.y == 0 && lane == 0)
        printf("cta %d mma thread: per unit: wait operands %lld, wait tempty %lld, issue %lld, total %lld\n", blockIdx.x,
               mp_full / u, mp_tempty / u, mp_issue / u, (clock64() - mp_t0) / u);
#endif
    }
    }
  } else {
    ptx::setmaxnreg_inc<FwCfg<BN>::kSoftmaxRegs>();
    // =========================== softmax / accumulate warps =================================
    // Thread = (pixel row, 64-token quarter of the CTA's 256 tokens): four warps per scheduler hide the
    // MUFU / TMEM latencies of each other.  Per unit u = (layer, head) the order is
    //   A(u): row max of my columns  ->  C(u-1): normalise + accumulate the PREVIOUS unit  ->  B(u): exponentials,
    // so the statistics exchange of unit u-1 (posted at the end of the previous iteration) has pass A to
    // land, and the MMAs of unit u+1 (which need the buffer C(u-1) frees) run under pass B.
    const int quad = warp & 3;                 // TMEM lane quadrant this warp may access
    const int qt = (warp - FW_SOFTMAX_WARP0) >> 2;   // column quarter
    const int row = quad * 32 + lane;
    const int col0 = qt * FW_QCOLS;
    const int ncol = max(0, min(FW_QCOLS, P.T - (tok0 + col0)));
    const uint32_t t_lane = tmem_base + (static_cast<uint32_t>(quad * 32) << 16);
    const int hf = qt / (FW_PARTS / FW_NH);    // hand-over group of the TMEM buffers
    const int part = static_cast<int>(cta_rank) * FW_PARTS + qt;
    constexpr int NPART = FW_PARTS * NSPLIT;
    constexpr int NCH = FW_QCOLS / 32;
    float acc[SELECT ? TC_MAX_SEL : FW_QCOLS];
#pragma unroll
    for (int j = 0; j < (SELECT ? TC_MAX_SEL : FW_QCOLS); ++j) acc[j] = 0.0f;

    // SELECT (extraction: only the K selected tokens are accumulated): the selected tokens of MY 64 columns as a compact,
    // warp-uniform list (at most 16, ~1.25 on average: 10 tokens over 8 column parts) -- TMEM column and output plane of
    // entry k packed 8 / 4 bits apiece, so every per-unit loop below runs over static k with a predicate `k < n_mine`
    // instead of sixteen range tests on shared-memory token ids.  Their exponentials are computed from single-column loads
    // of the LOGITS during pass B and stay in registers until the unit's statistics arrive: nothing is parked in TMEM (no
    // tcgen05.st, no pass-C load) and the buffer is handed back right after pass B's loads.  (Mode C used to be SLOWER than
    // mode A: 2.95 vs 2.45 ms per 8 images.)
    uint32_t n_mine = 0;
    unsigned long long my_cols[2] = {0ull, 0ull}, my_planes = 0ull;
    if (SELECT) {
#pragma unroll
      for (int j = 0; j < TC_MAX_SEL; ++j) {
        const int t = tok_s[j] - tok0 - col0;
        if (j < P.n_idx && t >= 0 && t < ncol) {
          if (n_mine < 8) my_cols[0] |= static_cast<unsigned long long>(t) << (8 * n_mine);
          else my_cols[1] |= static_cast<unsigned long long>(t) << (8 * (n_mine - 8));
          my_planes |= static_cast<unsigned long long>(j) << (4 * n_mine);
          ++n_mine;
        }
      }
    }
    auto my_col = [&](int k) { return static_cast<uint32_t>(my_cols[k >> 3] >> (8 * (k & 7))) & 0xffu; };
    float e_sel[SELECT ? TC_MAX_SEL : 1];      // of the unit whose pass B ran last = the unit the next pass C normalises
#pragma unroll
    for (int k = 0; k < (SELECT ? TC_MAX_SEL : 1); ++k) e_sel[k] = 0.0f;

    // Units between a unit's pass B and its normalisation (pass C).  1: the exponentials wait in TMEM, two buffers.  SELECT keeps
    // them in registers, so it can afford 2: a unit's statistics are then one whole unit old when they are needed and the 32
    // softmax warps of the cluster stop meeting at that barrier every unit (with lag 1 every warp's pass C waits for the
    // slowest warp's pass B of the same unit: lockstep, the MUFU pipe idle outside pass B).  A statistics slot may be
    // overwritten (remote st.async of unit u) once every reader is past unit u - XB: XB >= 2 * lag.
    constexpr int XLAG = (SELECT && BN == TC_BN) ? SKP_FW_XLAG : 1;
    constexpr int XB = 2 * XLAG;
    static_assert(XB <= SM::kStatBufs, "statistics ring");
    // normalise + accumulate unit uu, whose exponentials this thread parked in TMEM one iteration ago (SELECT: holds in e_u)
    FW_PROF_DECL
    auto pass_c = [&](uint32_t uu, float m_own, float wgt_u, bool fast_uu, const float (&e_u)[SELECT ? TC_MAX_SEL : 1]) {
      const uint32_t buf = uu & 1u;
      const uint32_t xb = uu % XB, par = (uu / XB) & 1u;
      const uint32_t taddr = t_lane + buf * TC_BNk + col0;
      ptx::mbar_wait_hint(bar_xfull(xb), par, 500u);
      FW_PROF(3)
      float mg = m_own, total = 0.0f;
      if (fast_uu) {
        // every part posted (the unit's constant m, sum exp2(S - m)): the row's statistics are the plain sum
        float tq[NPART];
#pragma unroll
        for (int q = 0; q < NPART; ++q) tq[q] = stats[(xb * FW_MAX_PARTS + q) * TC_BM + row].y;
#pragma unroll
        for (int st = 1; st < NPART; st <<= 1)
#pragma unroll
          for (int q = 0; q + st < NPART; q += 2 * st) tq[q] += tq[q + st];
        total = tq[0];
      } else {
        mg = -INFINITY;
#pragma unroll
        for (int q = 0; q < NPART; ++q) mg = fmaxf(mg, stats[(xb * FW_MAX_PARTS + q) * TC_BM + row].x);
#pragma unroll
        for (int q = 0; q < NPART; ++q) {
          const float2 sq = stats[(xb * FW_MAX_PARTS + q) * TC_BM + row];
          total += sq.y * pow2_int(sq.x - mg);
        }
      }
      if (P.row_stats != nullptr && part == 0)
        reinterpret_cast<float2*>(P.row_stats)[(static_cast<int64_t>(b) * P.ops.total_heads + uu) * P.R2 + p0 + row] =
            make_float2(mg, total);
      const float cf = ncol > 0 ? __fdividef(wgt_u * pow2_int(m_own - mg), total) : 0.0f;
      FW_PROF(4)
      if (!SELECT) {
        // four 16-column chunks, the load of chunk c + 1 in flight under the FMAs of chunk c (unconditional loads, like pass B;
        // cf is 0 for a thread without columns and the write-out skips columns past T)
        auto fma16 = [&](int c, const uint32_t (&w)[16]) {
#pragma unroll
          for (int j = 0; j < 16; ++j) acc[16 * c + j] = fmaf(cf, __uint_as_float(w[j]), acc[16 * c + j]);
        };
        {
          uint32_t a0[16];
          ptx::tmem_ld16(taddr, a0);
          ptx::tmem_ld_wait();
          FW_PROF(5)
          {
            uint32_t a1[16];
            ptx::tmem_ld16(taddr + 16, a1);
            fma16(0, a0);
            ptx::tmem_ld_wait();
            {
              uint32_t a2[16];
              ptx::tmem_ld16(taddr + 32, a2);
              fma16(1, a1);
              ptx::tmem_ld_wait();
              {
                uint32_t a3[16];
                ptx::tmem_ld16(taddr + 48, a3);
                fma16(2, a2);
                ptx::tmem_ld_wait();
                fma16(3, a3);
              }
            }
          }
          FW_PROF(6)
        }
      } else {
        // the selected tokens' exponentials of unit uu are in registers (pass B): no TMEM access, no buffer to release
#pragma unroll
        for (int g = 0; g < TC_MAX_SEL; g += 4) {
          if (static_cast<uint32_t>(g) < n_mine) {
#pragma unroll
            for (int k = g; k < g + 4; ++k) acc[k] = fmaf(cf, e_u[k], acc[k]);
          }
        }
      }
      if (!SELECT) {
        ptx::tc_fence_before();
        ptx::mbar_arrive(bar_tempty(buf, hf));
      }
    };

    uint32_t u = 0;
    float m_h[XLAG], wgt_h[XLAG], e_h[XLAG][SELECT ? TC_MAX_SEL : 1];      // [0] = the previous unit, [1] = the one before
    bool fast_h[XLAG];
#pragma unroll
    for (int d = 0; d < XLAG; ++d) {
      m_h[d] = wgt_h[d] = 0.0f;
      fast_h[d] = false;
#pragma unroll
      for (int k = 0; k < (SELECT ? TC_MAX_SEL : 1); ++k) e_h[d][k] = 0.0f;
    }
    for (int l = 0; l < P.ops.n_layers; ++l) {
      const int h = P.ops.heads[l];
      const float wgt = P.ops.weight[l];
      for (int head = 0; head < h; ++head, ++u) {
        const uint32_t buf = u & 1u, par = (u >> 1) & 1u;
        ptx::mbar_wait_hint(bar_tfull(buf, hf), par, 500u);
        ptx::tc_fence_after();
        FW_PROF(0)
        const uint32_t taddr = t_lane + buf * TC_BNk + col0;
        // bounded logits (see FW_FAST_HALFWIDTH): no row-max pass, m = the unit's constant.  Uniform over the cluster: one
        // value per (image, head)
        const float m_unit = LR && u < FW_MAX_UNITS ? amax_s[u] : INFINITY;
        const bool fast = LR && m_unit < INFINITY;
        // ---- pass A: row max over my columns (four independent chains)
        float m0 = -INFINITY, m1 = -INFINITY, m2 = -INFINITY, m3 = -INFINITY;
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
          if (!fast && 32 * c < ncol) {
            uint32_t v[32];
            ptx::tmem_ld32(taddr + 32 * c, v);
            ptx::tmem_ld_wait();
            FW_PROF(1)
            if (ncol >= 32 * (c + 1)) {
#pragma unroll
              for (int j = 0; j < 32; j += 8) {
                m0 = ptx::max3(m0, __uint_as_float(v[j]), __uint_as_float(v[j + 1]));
                m1 = ptx::max3(m1, __uint_as_float(v[j + 2]), __uint_as_float(v[j + 3]));
                m2 = ptx::max3(m2, __uint_as_float(v[j + 4]), __uint_as_float(v[j + 5]));
                m3 = ptx::max3(m3, __uint_as_float(v[j + 6]), __uint_as_float(v[j + 7]));
              }
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (32 * c + j < ncol) m0 = fmaxf(m0, __uint_as_float(v[j]));
            }
          }
        }
        // Rounded up to an integer: any m >= the true max stabilises the exponentials equally well, and
        // integer-valued maxima make every rescale 2^(m_part - m_row) an exponent-field add, not a MUFU.
        const float m = fast ? m_unit : ceilf(fmaxf(fmaxf(m0, m1), fmaxf(m2, m3)));
        FW_PROF(2)
        // ---- pass C of the previous unit
        if (u >= static_cast<uint32_t>(XLAG)) pass_c(u - XLAG, m_h[XLAG - 1], wgt_h[XLAG - 1], fast_h[XLAG - 1], e_h[XLAG - 1]);
        // ---- pass B: exponentials (parked back in TMEM) and their sum
        float s0 = 0.0f, s1 = 0.0f, s2 = 0.0f, s3 = 0.0f;
        uint32_t sv_sel[SELECT ? TC_MAX_SEL : 1];
        if (SELECT) {
          // logits of my selected tokens: predicated single-column loads, in flight together with the first chunk's load
          // (groups of four under one warp-uniform branch: a thread rarely owns more than four of the tokens)
#pragma unroll
          for (int g = 0; g < TC_MAX_SEL; g += 4) {
#pragma unroll
            for (int k = g; k < g + 4; ++k) sv_sel[k] = 0xff800000u;             // -inf: exp2 = 0 for the entries past n_mine
            if (static_cast<uint32_t>(g) < n_mine) {
#pragma unroll
              for (int k = g; k < g + 4; ++k) ptx::tmem_ld1_if(sv_sel[k], taddr + my_col(k), static_cast<uint32_t>(k) < n_mine);
            }
          }
        }
        // 16-column chunks, software-pipelined: the load of chunk c + 1 is in flight under the exponentials of chunk c (with
        // 32-column chunks loaded and waited for one after the other every unit exposed two TMEM round trips per pass, all
        // four warps of a scheduler at the same time, the MUFU pipe idle; two 16-column buffers cost the registers of one
        // 32-column chunk)
        {
          constexpr int NC16 = FW_QCOLS / 16;
          // loads are unconditional (columns past T's end are inside the buffer; their chunks are skipped below): straight-line
          // dataflow for the register allocator
          auto chunk = [&](int c, uint32_t (&w)[16]) {
            if (16 * c < ncol) {
              if (ncol >= 16 * (c + 1)) {
#pragma unroll
                for (int j = 0; j < 16; j += 4) {
                  const float e0 = ptx::ex2(__uint_as_float(w[j]) - m), e1 = ptx::ex2(__uint_as_float(w[j + 1]) - m);
                  const float e2 = ptx::ex2(__uint_as_float(w[j + 2]) - m), e3 = ptx::ex2(__uint_as_float(w[j + 3]) - m);
                  s0 += e0; s1 += e1; s2 += e2; s3 += e3;
                  w[j] = __float_as_uint(e0); w[j + 1] = __float_as_uint(e1);
                  w[j + 2] = __float_as_uint(e2); w[j + 3] = __float_as_uint(e3);
                }
              } else {
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                  const float e = (16 * c + j < ncol) ? ptx::ex2(__uint_as_float(w[j]) - m) : 0.0f;
                  s0 += e;
                  w[j] = __float_as_uint(e);
                }
              }
              if (!SELECT) ptx::tmem_st16(taddr + 16 * c, w);
            }
          };
          static_assert(NC16 == 4, "four 16-column chunks per thread");
          {
            uint32_t a0[16];
            ptx::tmem_ld16(taddr, a0);
            ptx::tmem_ld_wait();
            {
              uint32_t a1[16];
              ptx::tmem_ld16(taddr + 16, a1);
              chunk(0, a0);
              ptx::tmem_ld_wait();
              {
                uint32_t a2[16];
                ptx::tmem_ld16(taddr + 32, a2);
                chunk(1, a1);
                ptx::tmem_ld_wait();
                {
                  uint32_t a3[16];
                  ptx::tmem_ld16(taddr + 48, a3);
                  chunk(2, a2);
                  ptx::tmem_ld_wait();
                  if (SELECT) {
                    // my columns of the logits are in registers: the tensor pipe may reuse the buffer (unit u + 2)
                    ptx::tc_fence_before();
                    ptx::mbar_arrive(bar_tempty(buf, hf));
                  }
                  chunk(3, a3);
                }
              }
            }
          }
        }
        const float sum = (s0 + s1) + (s2 + s3);
        if (!SELECT) ptx::tmem_st_wait();
        if (SELECT) {
#pragma unroll
          for (int g = 0; g < TC_MAX_SEL; g += 4) {
            if (static_cast<uint32_t>(g) < n_mine) {
#pragma unroll
              for (int k = g; k < g + 4; ++k) e_sel[k] = ptx::ex2(__uint_as_float(sv_sel[k]) - m);
            }
          }
        }
        FW_PROF(9)
        // ---- post (max, sum) to every column part of the cluster.  Local slot: plain store + arrive.
        // Remote slot: st.async, which completes bytes on the peer's mbarrier itself (no cluster-scope
        // release fence, no L1 invalidate on the waiter).
        {
          const uint32_t xb = u % XB;
          const int sidx = (xb * FW_MAX_PARTS + part) * TC_BM + row;
          stats[sidx] = make_float2(m, sum);
          if (NSPLIT > 1) {
            const uint32_t slot = base + SM::kStatsOff + sidx * 8;
#pragma unroll
            for (uint32_t dst = 0; dst < static_cast<uint32_t>(NSPLIT); ++dst)
              if (dst != cta_rank) ptx::st_async_v2(ptx::mapa(slot, dst), m, sum, ptx::mapa(bar_xfull(xb), dst));
            if (warp == FW_SOFTMAX_WARP0 && lane == 0) ptx::mbar_arrive_expect_tx(bar_xfull(xb), (NSPLIT - 1) * FW_SOFTMAX_THREADS * 8);
            else ptx::mbar_arrive(bar_xfull(xb));
          } else {
            ptx::mbar_arrive(bar_xfull(xb));
          }
        }
#pragma unroll
        for (int d = XLAG - 1; d > 0; --d) {
          m_h[d] = m_h[d - 1]; wgt_h[d] = wgt_h[d - 1]; fast_h[d] = fast_h[d - 1];
#pragma unroll
          for (int k = 0; k < (SELECT ? TC_MAX_SEL : 1); ++k) e_h[d][k] = e_h[d - 1][k];
        }
        m_h[0] = m; wgt_h[0] = wgt; fast_h[0] = fast;
#pragma unroll
        for (int k = 0; k < (SELECT ? TC_MAX_SEL : 1); ++k) e_h[0][k] = e_sel[k];
        FW_PROF(10)
      }
    }
#ifdef SKP_FWD_PROF
    if (blockIdx.x < 2 && blockIdx.y == 0 && lane == 0 && (warp & 3) == 0)
      printf("cta %d warp %2d: wait_mma %lld | A ld %lld cmp %lld | wait_x %lld comb %lld | C ld %lld fma %lld (+arrive in next) | B ld %lld cmp %lld st+wait %lld | post %lld\n",
             blockIdx.x, warp, pf_acc[0] / u, pf_acc[1] / u, pf_acc[2] / u, pf_acc[3] / u, pf_acc[4] / u, pf_acc[5] / u, pf_acc[6] / u,
             pf_acc[7] / u, pf_acc[8] / u, pf_acc[9] / u, pf_acc[10] / u);
#endif
#pragma unroll
    for (int d = XLAG; d >= 1; --d)
      if (u >= static_cast<uint32_t>(d)) pass_c(u - d, m_h[d - 1], wgt_h[d - 1], fast_h[d - 1], e_h[d - 1]);
    // ---- write-out: lanes hold consecutive pixels -> 128 B coalesced stores per token
    if (!SELECT) {
      float* out = P.maps + (static_cast<int64_t>(b) * P.T + tok0 + col0) * P.R2 + p0 + row;
#pragma unroll
      for (int j = 0; j < FW_QCOLS; ++j)
        if (j < ncol) out[static_cast<int64_t>(j) * P.R2] = acc[j];
    } else {
#pragma unroll
      for (int k = 0; k < TC_MAX_SEL; ++k) {
        const int j = static_cast<int>(my_planes >> (4 * k)) & 15;
        if (static_cast<uint32_t>(k) < n_mine) P.maps[(static_cast<int64_t>(b) * P.n_idx + j) * P.R2 + p0 + row] = acc[k];
      }
    }
  }

  // ---- teardown -------------------------------------------------------------------------------
  __syncwarp();
  ptx::tc_fence_before();
  __syncthreads();
  if (NSPLIT > 1) ptx::cluster_sync();   // nobody exits while a peer may still store into its smem
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, 2 * BN);
  }
}

}  // namespace skp

// ==========================================================================================
// host side
// ==========================================================================================
namespace skp {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || !p) return nullptr;
    fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// bf16 tensor [batch][rows][d] (d contiguous) -> 3-D map with a (64, box_rows, 1) box, 128B swizzle
int tc_encode_map(CUtensorMap* m, void* ptr, int d, int rows, int batch, int box_rows) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled entry point not available");
    return SKP_ERR_CUDA;
  }
  cuuint64_t dims[3] = {static_cast<cuuint64_t>(d), static_cast<cuuint64_t>(rows), static_cast<cuuint64_t>(batch)};
  cuuint64_t strides[2] = {static_cast<cuuint64_t>(d) * 2, static_cast<cuuint64_t>(d) * 2 * rows};
  cuuint32_t box[3] = {static_cast<cuuint32_t>(TC_BK), static_cast<cuuint32_t>(box_rows), 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, ptr, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d (d=%d rows=%d batch=%d)", static_cast<int>(r), d, rows, batch);
    return SKP_ERR_CUDA;
  }
  return SKP_OK;
}

// fp32 (inner, rows, batch) array, box (32 floats = one 128 B swizzle row, box_rows, 1): the backward's logit-gradient flush
int tc_encode_map_f32(CUtensorMap* m, void* ptr, int inner, int rows, int batch, int box_rows) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled entry point not available");
    return SKP_ERR_CUDA;
  }
  cuuint64_t dims[3] = {static_cast<cuuint64_t>(inner), static_cast<cuuint64_t>(rows), static_cast<cuuint64_t>(batch)};
  cuuint64_t strides[2] = {static_cast<cuuint64_t>(inner) * 4, static_cast<cuuint64_t>(inner) * 4 * rows};
  cuuint32_t box[3] = {32, static_cast<cuuint32_t>(box_rows), 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, ptr, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled (fp32) failed with CUresult %d (inner=%d rows=%d batch=%d)", static_cast<int>(r), inner, rows, batch);
    return SKP_ERR_CUDA;
  }
  return SKP_OK;
}

bool attn_map_tc_supported(const skp_attn_layer* h_layers, int n_layers, int B, int T, int R, int n_idx) {
  if (!h_layers || n_layers < 1 || n_layers > TC_MAX_LAYERS || B < 1 || B > 65535) return false;
  if (T < 1 || T > 2 * TC_BN || (static_cast<int64_t>(R) * R) % TC_BM != 0) return false;
  if (n_idx > TC_MAX_SEL) return false;
  for (int l = 0; l < n_layers; ++l) {
    const skp_attn_layer& a = h_layers[l];
    if (a.heads < 1 || a.C % a.heads != 0) return false;
    const int d = a.C / a.heads;
    if (d % 16 != 0 || d > 512 || a.s < 1 || R % a.s != 0) return false;
  }
  return true;
}


// staged operands: upsampled q' and k, bf16 hi (+ lo)
size_t attn_map_staged_workspace(const skp_attn_layer* h_layers, int n_layers, int B, int T, int R, int precision) {
  const int nbuf = precision == SKP_PREC_TC_BF16 ? 1 : 2;
  size_t total = 0;
  for (int l = 0; l < n_layers; ++l) {
    const skp_attn_layer& a = h_layers[l];
    total += nbuf * tc_align_up(static_cast<size_t>(B) * R * R * a.C * 2, 1024);
    total += nbuf * tc_align_up(static_cast<size_t>(a.k_shared ? 1 : B) * T * a.C * 2, 1024);
  }
  return total + 1024;
}

// parity mode on shapes whose interpolation windows are narrow: upsample after the logits (attn_tc_lr.cu), no staged
// q'.  SKP_FWD_LR=0 keeps the staged-operand path (A/B measurements, cross-checks); read per call.
static bool fwd_uses_lr(const skp_attn_layer* h_layers, int n_layers, int B, int T, int R, int n_idx, int precision) {
  const char* e = getenv("SKP_FWD_LR");
  return precision != SKP_PREC_TC_BF16 && !(e && atoi(e) == 0) && attn_map_lr_supported(h_layers, n_layers, B, T, R, n_idx);
}

// does skp_attn_map_fwd (all tokens, this precision) stage upsample-after-logits operands?  The backward may then reuse them
bool attn_map_fwd_stages_lr(const skp_attn_layer* h_layers, int n_layers, int B, int T, int R, int precision) {
  return fwd_uses_lr(h_layers, n_layers, B, T, R, 0, precision);
}

// what the forward call will need: 0.45 GB for 8 SD-2.1 images on the upsample-after-logits path, 2.35 GB staged
size_t attn_map_tc_workspace(const skp_attn_layer* h_layers, int n_layers, int B, int T, int R, int precision) {
  if (fwd_uses_lr(h_layers, n_layers, B, T, R, 0, precision)) return attn_map_lr_workspace(h_layers, n_layers, B, T, R) + 1024;
  return attn_map_staged_workspace(h_layers, n_layers, B, T, R, precision);
}

template <int NPROD, int NSPLIT, bool SELECT, int BN = TC_BN, bool LR = false>
static int launch_tc(const TcParams& P, int B, int R2, cudaStream_t st) {
  using SM = TcSmem<NPROD, BN, LR>;
  auto kern = attn_map_fwd_tc_kernel<NPROD, NSPLIT, SELECT, BN, LR>;
  SKP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SM::kDynamic));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(static_cast<unsigned>(R2 / TC_BM * NSPLIT), static_cast<unsigned>(B), 1);
  cfg.blockDim = dim3(FwCfg<BN>::kThreads, 1, 1);
  cfg.dynamicSmemBytes = SM::kDynamic;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = NSPLIT;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = NSPLIT > 1 ? 1 : 0;
  note_launch();
  SKP_CUDA(cudaLaunchKernelEx(&cfg, kern, P));
  return SKP_OK;
}

int tc_prepare(const skp_attn_layer* h_layers, int n_layers, int B, int T, int R, bool x3, void* workspace, cudaStream_t st,
               TcOperands* O, char** ws_end) {
  const int R2 = R * R;
  char* ws = reinterpret_cast<char*>(tc_align_up(reinterpret_cast<size_t>(workspace), 1024));
  int base = 0;
  for (int l = 0; l < n_layers; ++l) {
    const skp_attn_layer& a = h_layers[l];
    SKP_REQUIRE(a.q && a.k, "skp_attn_map: layer %d has null q/k", l);
    const int d = a.C / a.heads;
    const int Bk = a.k_shared ? 1 : B;
    const size_t qbytes = tc_align_up(static_cast<size_t>(B) * R2 * a.C * 2, 1024);
    const size_t kbytes = tc_align_up(static_cast<size_t>(Bk) * T * a.C * 2, 1024);
    __nv_bfloat16* q_hi = reinterpret_cast<__nv_bfloat16*>(ws); ws += qbytes;
    __nv_bfloat16* q_lo = nullptr;
    if (x3) { q_lo = reinterpret_cast<__nv_bfloat16*>(ws); ws += qbytes; }
    __nv_bfloat16* k_hi = reinterpret_cast<__nv_bfloat16*>(ws); ws += kbytes;
    __nv_bfloat16* k_lo = nullptr;
    if (x3) { k_lo = reinterpret_cast<__nv_bfloat16*>(ws); ws += kbytes; }
    const float mult = a.scale * 1.4426950408889634f;
    dim3 gq(R2, B), gk(T, Bk);
    const bool rows = a.s != R && R <= 256 && a.s <= 64 && (a.C / a.heads) % 2 == 0;
    dim3 gr(R, B, (a.C + PREP_CH - 1) / PREP_CH);
    const size_t rsm = static_cast<size_t>(a.s) * PREP_CH * sizeof(float);
    if (x3) {
      if (rows) {
        SKP_CUDA(cudaFuncSetAttribute(tc_prep_q_rows_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * PREP_CH * 4));
        tc_prep_q_rows_kernel<3><<<(::skp::note_launch(), gr), 128, rsm, st>>>(a.q, a.s, a.C, a.heads, R, mult, q_hi, q_lo);
      } else {
        tc_prep_q_kernel<3><<<(::skp::note_launch(), gq), 256, 0, st>>>(a.q, a.s, a.C, a.heads, R, mult, q_hi, q_lo);
      }
      tc_prep_k_kernel<3><<<(::skp::note_launch(), gk), 256, 0, st>>>(a.k, T, a.C, a.heads, k_hi, k_lo);
    } else {
      if (rows) {
        SKP_CUDA(cudaFuncSetAttribute(tc_prep_q_rows_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * PREP_CH * 4));
        tc_prep_q_rows_kernel<1><<<(::skp::note_launch(), gr), 128, rsm, st>>>(a.q, a.s, a.C, a.heads, R, mult, q_hi, q_lo);
      } else {
        tc_prep_q_kernel<1><<<(::skp::note_launch(), gq), 256, 0, st>>>(a.q, a.s, a.C, a.heads, R, mult, q_hi, q_lo);
      }
      tc_prep_k_kernel<1><<<(::skp::note_launch(), gk), 256, 0, st>>>(a.k, T, a.C, a.heads, k_hi, k_lo);
    }
    SKP_LAUNCH_CHECK();
    int rc = tc_encode_map(&O->q_map[l][0], q_hi, d, R2, B * a.heads, TC_BM);
    if (rc == SKP_OK) rc = tc_encode_map(&O->k_map[l][0], k_hi, d, T, Bk * a.heads, TC_BN);
    if (rc == SKP_OK && x3) rc = tc_encode_map(&O->q_map[l][1], q_lo, d, R2, B * a.heads, TC_BM);
    if (rc == SKP_OK && x3) rc = tc_encode_map(&O->k_map[l][1], k_lo, d, T, Bk * a.heads, TC_BN);
    if (rc != SKP_OK) return rc;
    O->k_ptr[l][0] = k_hi; O->k_ptr[l][1] = k_lo;
    O->heads[l] = a.heads; O->d[l] = d; O->k_shared[l] = a.k_shared; O->head_base[l] = base;
    O->weight[l] = a.weight / a.heads; O->scale[l] = a.scale;
    base += a.heads;
  }
  O->n_layers = n_layers; O->total_heads = base;
  if (ws_end) *ws_end = ws;
  return SKP_OK;
}

int attn_map_fwd_tc(const skp_attn_layer* h_layers, int n_layers, int B, int T, int R, const int64_t* token_idx, int n_idx,
                    int64_t idx_stride, float* maps, float* row_stats, int precision, void* workspace, size_t workspace_bytes, cudaStream_t st) {
  const size_t need = attn_map_tc_workspace(h_layers, n_layers, B, T, R, precision);
  if (!workspace || workspace_bytes < need) {
    set_error("skp_attn_map_fwd: workspace %zu B < required %zu B", workspace_bytes, need);
    return SKP_ERR_WORKSPACE;
  }
  const bool x3 = precision != SKP_PREC_TC_BF16;
  const int R2 = R * R;
  TcParams P;
  memset(&P, 0, sizeof(P));
  const bool lr = fwd_uses_lr(h_layers, n_layers, B, T, R, token_idx ? n_idx : 0, precision);
  // SKP_FWD_BN=128: 128-token CTAs, two per SM (cluster of four) -- measured slower, kept for A/B runs (see FwCfg)
  const char* ebn = getenv("SKP_FWD_BN");
  const bool bn128 = lr && x3 && T > TC_BN && ebn && atoi(ebn) == 128;
  int rc = lr ? tc_prepare_lr(h_layers, n_layers, B, T, R, workspace, st, &P.ops, nullptr, bn128 ? 128 : TC_BN)
              : tc_prepare(h_layers, n_layers, B, T, R, x3, workspace, st, &P.ops, nullptr);
  if (rc != SKP_OK) return rc;
  if (const char* ef = getenv("SKP_FWD_FAST"))      // SKP_FWD_FAST=0: always the max-subtracted path (A/B runs, tests)
    if (atoi(ef) == 0)
      for (int l = 0; l < TC_MAX_LAYERS; ++l) P.ops.absmax[l] = nullptr;
  P.T = T; P.R2 = R2;
  P.n_idx = token_idx ? n_idx : 0;
  P.token_idx = token_idx; P.maps = maps; P.idx_stride = idx_stride; P.row_stats = row_stats;
  const bool sel = token_idx != nullptr;
  const bool split = T > TC_BN;
  if (bn128) return sel ? launch_tc<3, 4, true, 128, true>(P, B, R2, st) : launch_tc<3, 4, false, 128, true>(P, B, R2, st);
  if (lr) {
    if (split) return sel ? launch_tc<3, 2, true, TC_BN, true>(P, B, R2, st) : launch_tc<3, 2, false, TC_BN, true>(P, B, R2, st);
    return sel ? launch_tc<3, 1, true, TC_BN, true>(P, B, R2, st) : launch_tc<3, 1, false, TC_BN, true>(P, B, R2, st);
  }
  if (x3) {
    if (split) return sel ? launch_tc<3, 2, true>(P, B, R2, st) : launch_tc<3, 2, false>(P, B, R2, st);
    return sel ? launch_tc<3, 1, true>(P, B, R2, st) : launch_tc<3, 1, false>(P, B, R2, st);
  }
  if (split) return sel ? launch_tc<1, 2, true>(P, B, R2, st) : launch_tc<1, 2, false>(P, B, R2, st);
  return sel ? launch_tc<1, 1, true>(P, B, R2, st) : launch_tc<1, 1, false>(P, B, R2, st);
}

}  // namespace skp
