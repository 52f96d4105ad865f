// Backward of the fused cross-attention map on tcgen05 tensor cores (sm_100a): gradient of
//   maps[b, t, p] = sum_l w_l * mean_head softmax_t( scale_l * q'_l[b,head,p,:] . k_l[head,t,:] )
// w.r.t. the module's low-resolution q and k, for an upstream gradient that is non-zero on the
// n_idx selected token planes only (optimize.py:420-424).  It is what
// autograd does in the reference through ptp_utils.py:513-538 + optimize.py:27-79, without the
// stored (h, R^2, T) probabilities.
//
// Per unit = (layer, head, image, 128-pixel tile), on one persistent CTA per (token half, unit range):
//   S     = Q' K^T              (tcgen05, bf16x3 or bf16; logits of this CTA's 256 tokens, TMEM)
//   S_sel = Q' K_sel^T          (the n_idx selected tokens' logits, N = 16, TMEM)
//   P     = exp2(S - m) / l     with the row max / sum (m, l) SAVED BY THE FORWARD (row_stats)
//   delta = sum_j P_sel[j] dP[j],  dS = P * (dP - delta)      -> bf16 tile in shared memory
// staged layers (template LR = false):
//   dQ'   = dS K                (tcgen05; A = dS K-major, B = K hi as an MN-major operand)
//   dK   += dS^T Q'             (tcgen05; A = dS and B = Q' hi both as MN-major operands; TMEM-
//                                resident across all units of a head, flushed once per head)
//   dq   += U^T dQ'             (adjoint of the bicubic upsample, separable: x-gather through
//                                shared memory, y taps as red.global.add.v4.f32)
// "upsample after the logits" layers (LR = true; Q' = the tile's interpolation table W, K = its window of native-resolution logits):
//   dz   += dS^T W              (tcgen05; TMEM-resident over the tiles that share a window, flushed through shared memory
//                                and a TMA reduce-add into a bf16 dz; attn_tc_bwd_lr.cu turns dz into dq / dk)
// Roles (12 warps): TMA producer (3 hi / 2 lo operand slots, windows fetched a unit ahead), S issuer and dS-product issuer
// (whole warps, elected issue), 8 softmax warps (S preloaded to registers, the accumulator handed back at once; the next
// unit's scalar prep and row prologue inside the dense pass).  DESIGN.md section 4 A' has the measurement behind each choice.
// No cluster, no cross-CTA exchange: the two token halves only meet in the global reductions.
#include <cstdlib>

#include "attn_tc.cuh"

namespace skp {

// Diagnostic builds only (nvcc -DSKP_DIAG): the SKP_BWD_DBG environment switches (bits 1 / 2 / 4 skip the reductions, the
// adjoint, the drain -- WRONG gradients by design, for phase timing; 8 prints the in-kernel phase profile, which allocates
// and synchronises; 32 / 128 / 256 change the MMA issue order) and the clock reads behind them.  The product library is
// built without them: no debug switch can change a result, nothing allocates or synchronises (header contract).
#ifdef SKP_DIAG
#define BW_CLK() clock64()
#define BW_DBG(P) ((P).dbg)
#else
#define BW_CLK() 0ll
#define BW_DBG(P) 0
#endif

// Timestamp trace (diagnostic builds only: nvcc -DSKP_BWD_TRACE): lane 0 of every warp of one CTA stamps clock64 at the
// marked events of units BW_TR_U0 .. +BW_TR_N and prints them at the end of the kernel (scripts/bwd_trace.py reads the lines).
#ifdef SKP_BWD_TRACE
constexpr int BW_TR_U0 = 40, BW_TR_N = 12, BW_TR_EV = 12, BW_TR_CTA = 2;
#ifndef SKP_BWD_TRACE_STAGED
constexpr bool BW_TR_LR = true;      // -DSKP_BWD_TRACE_STAGED: trace the staged kernel instead (event 10 = adjoint done)
#else
constexpr bool BW_TR_LR = false;
#endif
__device__ long long g_bw_tr[12][BW_TR_N][BW_TR_EV];
#define BW_TR(ev) { if (LR == BW_TR_LR && !WIDE && blockIdx.x == BW_TR_CTA && lane == 0 && it >= BW_TR_U0 && it < BW_TR_U0 + BW_TR_N) g_bw_tr[warp][it - BW_TR_U0][ev] = clock64(); }
#else
#define BW_TR(ev)
#endif

constexpr int BW_SEL = 16;            // selected tokens per launch (UMMA N of the S_sel product)
constexpr int BW_KSEL_BYTES = BW_SEL * 128;   // 16 rows x 128 B
constexpr int BW_DS_BYTES = TC_BM * TC_BN * 2;   // 64 KB: dS tile, 4 slabs of [128 px][64 tok] bf16
constexpr int BW_WXT_FLOATS = 8 * 256;        // x-adjoint table: s rows of 5 R/s taps padded to a multiple of 4 (<= 8 R/s... R <= 256)
constexpr float kLog2eBw = 1.4426950408889634f;
constexpr int BW_STG_LD = 68;          // fp32 staging row pitch (64 + 4): conflict-free float4 stores with lane = row

struct BwdParams {
  TcOperands ops;
  CUtensorMap ksel_map[TC_MAX_LAYERS][2];   // [layer][hi, lo]: dims (d, 16, B*h), box (64, 16, 1)
  float* dq[TC_MAX_LAYERS];                 // (B, s*s, C), accumulated into
  float* dk[TC_MAX_LAYERS];                 // (Bk, T, C), accumulated into
  int s[TC_MAX_LAYERS], C[TC_MAX_LAYERS];
  // "upsample after the logits" launches (template LR; ops then holds W tables / native-resolution logits, ksel_map the
  // selected tokens' logit rows): fp32 accumulator of the logit gradient dz[b*h + head][token][s*s] of every layer
  float* dz[TC_MAX_LAYERS];
  CUtensorMap dz_map[TC_MAX_LAYERS];        // dz as (s*s, T, B*h) fp32, box (32, 128, 1): target of the window flush (TMA reduce-add)
  const unsigned char* wr0[TC_MAX_LAYERS];  // first source row of every tile's window (lr_window_row0 costs two divisions
                                            // and a float->int chain, ~700 cycles of dependent latency per call)
  int ghead[TC_MAX_LAYERS];                 // first head of the layer among ALL captured layers (row_stats index)
  int heads_all;                            // heads of all captured layers (row_stats stride)
  int B, T, R, R2, n_idx, tiles, n_units, n_chunks;
  int64_t idx_stride;
  const int64_t* token_idx;
  const float* dmaps_sel;                   // (B, n_idx, R2)
  const float* dpT;                         // pixel-major copy (B, R2, 16), planes j >= n_idx zero
  const float* row_stats;                   // (B, total_heads, R2, 2)
  unsigned long long* prof;                 // SKP_BWD_DBG & 8: per-phase cycle counters of CTA 0
  int dbg;                                  // SKP_DIAG builds: SKP_BWD_DBG bits (profiling experiments only): 1 no reductions, 2 no adjoint, 4 no drain
};

// k_sel[(b*h + head)][j][dd] = k[bk][token_idx[b][j]][head*d + dd]  (rows j >= n_idx are zero)
template <int NPROD>
__global__ void bw_prep_ksel_kernel(const float* __restrict__ k, int T, int C, int heads, int k_shared,
                                    const int64_t* __restrict__ token_idx, int n_idx, int64_t idx_stride,
                                    __nv_bfloat16* __restrict__ hi, __nv_bfloat16* __restrict__ lo) {
  const int j = blockIdx.x, b = blockIdx.y;
  const int d = C / heads;
  const int t = j < n_idx ? static_cast<int>(token_idx[b * idx_stride + j]) : -1;
  const float* src = k + (static_cast<int64_t>(k_shared ? 0 : b) * T + (t < 0 ? 0 : t)) * C;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const float v = (t >= 0 && t < T) ? src[c] : 0.0f;
    const int head = c / d, dd = c % d;
    const int64_t o = ((static_cast<int64_t>(b) * heads + head) * BW_SEL + j) * d + dd;
    const __nv_bfloat16 h = __float2bfloat16_rn(v);
    hi[o] = h;
    if (NPROD == 3) lo[o] = __float2bfloat16_rn(v - __bfloat162float(h));
  }
}

// LR: s_sel[(b*h + head)][j][sp] = s_low[(b*h + head)][token_idx[b][j]][sp]  (rows j >= n_idx are zero)
__global__ void bw_gather_ssel_kernel(const __nv_bfloat16* __restrict__ s_hi, const __nv_bfloat16* __restrict__ s_lo, int T, int s2,
                                      int heads, const int64_t* __restrict__ token_idx, int n_idx, int64_t idx_stride,
                                      __nv_bfloat16* __restrict__ hi, __nv_bfloat16* __restrict__ lo) {
  const int j = blockIdx.x, z = blockIdx.y, b = z / heads;
  const int t = j < n_idx ? static_cast<int>(token_idx[b * idx_stride + j]) : -1;
  const bool ok = t >= 0 && t < T;
  const int64_t src = (static_cast<int64_t>(z) * T + (ok ? t : 0)) * s2, dst = (static_cast<int64_t>(z) * BW_SEL + j) * s2;
  for (int c = threadIdx.x; c < s2; c += blockDim.x) {
    hi[dst + c] = ok ? s_hi[src + c] : __float2bfloat16_rn(0.0f);
    if (lo) lo[dst + c] = ok ? s_lo[src + c] : __float2bfloat16_rn(0.0f);
  }
}

// WIDE (LR only): 128-tap interpolation windows (32^2 sources at R = 128) as TWO 64-tap k-blocks.  The logit gradient of a
// window then needs 128 TMEM columns per 128 tokens, which only fits next to S when the CTA owns 128 tokens (S 128 + dz
// 128 + S_sel 32 columns) instead of 256: four token quarters per tile, ONE operand stage (64 KB: both k-blocks, hi + lo),
// strictly serial S -> dS -> dz per unit.  Slower per exponential than the staged dQ'/dK kernel, but it does not see
// dim_head at all: it is the tensor-core path of wide-window layers with dim_head > 64 (SD-1.5: 80 at 32^2).
template <int NPROD, bool WIDE = false>
struct BwSmem {
  static constexpr int kNkb = WIDE ? 2 : 1;                         // 64-wide k-blocks of the S product
  static constexpr int kBn = WIDE ? 128 : TC_BN;                    // tokens per CTA
  // A operand (Q' tile / interpolation table of the pixel tile): the hi half is also the B operand of the products that
  // consume dS, so its slot is busy until those retire; the lo half only until the unit's logit product has.  With two
  // combined stages the tile of unit u+1 could only be fetched once the dz product of unit u-1 had retired, and S(u+1)
  // arrived ~1 k cycles late every other unit (SKP_BWD_TRACE): three hi slots + two lo slots instead (+16 KB).
  static constexpr int kHiStages = WIDE ? 1 : 3;
  static constexpr int kLoStages = WIDE ? 1 : 2;
  static constexpr int kAHalf = kNkb * TC_A_BYTES;                  // one slot: k-block 0 (, 1)
  static constexpr int kQStage = (NPROD == 3 ? 2 : 1) * kAHalf;     // bytes of a unit's A operand (hi + lo)
  static constexpr int kQOff = 0;
  static constexpr int kLoOff = kQOff + kHiStages * kAHalf;
  static constexpr int kKBlock = kBn * TC_BK * 2;                   // one k-block of the k / logit-window tile
  static constexpr int kKHalf = kNkb * kKBlock;
  static constexpr int kKOff = kLoOff + (NPROD == 3 ? kLoStages * kAHalf : 0);          // k hi (+ lo)
  static constexpr int kKBytes = (NPROD == 3 ? 2 : 1) * kKHalf;
  static constexpr int kKselOff = kKOff + kKBytes;
  static constexpr int kKselHalf = kNkb * BW_KSEL_BYTES;
  static constexpr int kKselBytes = (NPROD == 3 ? 2 : 1) * kKselHalf;
  static constexpr int kDsOff = kKselOff + kKselBytes;              // dS tile; aliased by the fp32 dQ' staging
  static constexpr int kDsBytes = TC_BM * kBn * 2;
  static constexpr int kWxtOff = kDsOff + (WIDE ? kDsBytes : BW_DS_BYTES);
  static constexpr int kBarOff = kWxtOff + BW_WXT_FLOATS * 4;       // q_full[3], hi_empty[3], lo_empty[2], s_full, ds_full, mma_done, sel_full[2], s_free
  static constexpr int kTmemPtrOff = kBarOff + 8 * 14;
  static constexpr int kTokOff = kTmemPtrOff + 8;                   // int tok[16]
  static constexpr int kTotal = kTokOff + 4 * BW_SEL;
  static constexpr int kDynamic = kTotal + 1024;
  static_assert(kDynamic <= 232448, "over the 227 KB of dynamic shared memory per CTA");
};

__device__ __forceinline__ uint64_t bw_desc(uint32_t addr, uint32_t lbo, uint32_t sbo) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((addr & 0x3FFFFu) >> 4);
  d |= static_cast<uint64_t>(lbo >> 4) << 16;
  d |= static_cast<uint64_t>(sbo >> 4) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(2) << 61;      // SWIZZLE_128B
  return d;
}
__device__ __host__ constexpr uint32_t bw_idesc(int M, int N, int a_mn, int b_mn) {
  return (1u << 4) | (1u << 7) | (1u << 10) | (static_cast<uint32_t>(a_mn) << 15) | (static_cast<uint32_t>(b_mn) << 16) |
         (static_cast<uint32_t>(N >> 3) << 17) | (static_cast<uint32_t>(M >> 4) << 24);
}
__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
// dmaps_sel (B, n_idx, R2) -> (B, R2, 16): the softmax threads read the n_idx upstream gradients of their pixel row
__global__ void bw_transpose_dp_kernel(const float* __restrict__ in, int n_idx, int R2, float* __restrict__ out) {
  const int b = blockIdx.y, p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= R2) return;
  float v[BW_SEL];
#pragma unroll
  for (int j = 0; j < BW_SEL; ++j) v[j] = j < n_idx ? in[(static_cast<int64_t>(b) * n_idx + j) * R2 + p] : 0.0f;
  float4* o = reinterpret_cast<float4*>(out + (static_cast<int64_t>(b) * R2 + p) * BW_SEL);
#pragma unroll
  for (int q = 0; q < BW_SEL / 4; ++q) o[q] = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
}
__global__ void bw_window_table_kernel(int s, int R, int nr, int tiles, unsigned char* tab) {
  for (int t = threadIdx.x; t < tiles; t += blockDim.x) tab[t] = static_cast<unsigned char>(lr_window_row0(t, s, R, nr));
}
__device__ __forceinline__ void named_bar_sync(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }
struct BwUnit {
  int l, head, hg, b, tile;
  int r0;      // LR launches: first source row of the tile's window (table load issued when the unit is decoded, so the
               // window tests at the end of a unit compare registers instead of waiting on two dependent loads)
};
__device__ __forceinline__ int bw_r0(const BwdParams& P, int l, int tile) { return P.wr0[l] ? static_cast<int>(P.wr0[l][tile]) : 0; }
__device__ __forceinline__ BwUnit bw_decode(const BwdParams& P, int u) {
  BwUnit x;
  const int per_head = P.B * P.tiles;
  x.hg = u / per_head;
  const int rem = u - x.hg * per_head;
  x.b = rem / P.tiles;
  x.tile = rem - x.b * P.tiles;
  int l = 0;
  while (l + 1 < P.ops.n_layers && x.hg >= P.ops.head_base[l + 1]) ++l;
  x.l = l;
  x.head = x.hg - P.ops.head_base[l];
  x.r0 = bw_r0(P, x.l, x.tile);
  return x;
}
// does the dK accumulator have to be flushed after unit u (last of its head / image / range)?
__device__ __forceinline__ bool bw_flush_after(const BwdParams& P, int u, int u_end) {
  if (u + 1 >= u_end) return true;
  const int per_head = P.B * P.tiles;
  const int hg = u / per_head, hg1 = (u + 1) / per_head;
  if (hg != hg1) return true;
  int l = 0;
  while (l + 1 < P.ops.n_layers && hg >= P.ops.head_base[l + 1]) ++l;
  if (P.ops.k_shared[l]) return false;
  return (u - hg * per_head) / P.tiles != (u + 1 - hg * per_head) / P.tiles;
}

// Units are walked in order, so after the first decode the (layer, head, image, tile) tuple is advanced
// incrementally: the integer divisions of bw_decode were ~1.5k cycles of dependent latency per unit.
__device__ __forceinline__ BwUnit bw_next(const BwdParams& P, BwUnit x) {
  if (++x.tile == P.tiles) {
    x.tile = 0;
    if (++x.b == P.B) {
      x.b = 0;
      ++x.hg;
      if (x.l + 1 < P.ops.n_layers && x.hg >= P.ops.head_base[x.l + 1]) ++x.l;
      x.head = x.hg - P.ops.head_base[x.l];
    }
  }
  x.r0 = bw_r0(P, x.l, x.tile);
  return x;
}
// same predicate as bw_flush_after, from the current and the next tuple.  LR: the accumulator holds the logit gradient
// of one window of source rows of one (image, head): flush when the next tile interpolates from another window.
template <bool LR>
__device__ __forceinline__ bool bw_flush_between(const BwdParams& P, const BwUnit& x, const BwUnit& nx, bool last) {
  if (LR)
    return last || nx.hg != x.hg || nx.b != x.b || nx.r0 != x.r0;
  return last || nx.hg != x.hg || (!P.ops.k_shared[x.l] && nx.b != x.b);
}

// 12 warps = 3 warpgroups: {TMA, S issuer, dz issuer, idle} shrink to 64 registers, the 8 softmax warps grow to 216
constexpr int BW_THREADS = 384;
// setmaxnreg moves registers inside the CTA's launch allocation (384 threads x 168): what the four small warps give back
// must cover what the eight softmax warps take, or setmaxnreg.inc waits forever
constexpr int BW_REGS_LAUNCH = 168, BW_REGS_SMALL = 64, BW_REGS_SOFTMAX = 216;
static_assert(4 * (BW_REGS_LAUNCH - BW_REGS_SMALL) >= 8 * (BW_REGS_SOFTMAX - BW_REGS_LAUNCH), "setmaxnreg budget");
#ifndef BW_PREP_CHUNK
#define BW_PREP_CHUNK 0     // dense chunk behind which the next unit's scalar prep and row prefetch sit (1: 3.14 vs 2.98 ms; under the TMEM round trip of the S preload, ahead of s_free: 3.06)
#endif
#ifndef BW_PROLOGUE_LEAD
#define BW_PROLOGUE_LEAD 1  // chunks between the next unit's S_sel load and the last chunk, where its row prologue runs (2: same time)
#endif
#ifndef BW_LR_HOLD
#define BW_LR_HOLD 3      // chunks (32 tokens each) of a row's dS kept in registers across the wait for the previous dz product
                          // (measured on the final kernel, per 8-image call: 1: 3.04 ms, 2: 3.04, 3: 2.98, 4: 3.50)
#endif
constexpr int BW_SOFTMAX_WARP0 = 4;
// LR: "upsample after the logits" gradient.  The dQ' / dK products and the bicubic adjoint are replaced by ONE product
// dz += dS^T W (tokens x 64 taps, K = 128 pixels; same operand forms as the dK product with the W tile of this pixel
// tile in place of Q'), accumulated in TMEM over the tiles that share a window of source rows and flushed with vector
// reductions into dz[b*h + head][token][window]; attn_tc_bwd_lr.cu turns dz into dq / dk at native resolution.
template <int NPROD, int NH, bool LR, bool WIDE = false>
__global__ void __launch_bounds__(BW_THREADS, 1)
attn_map_bwd_tc_kernel(const __grid_constant__ BwdParams P) {
  using SM = BwSmem<NPROD, WIDE>;
  static_assert(!WIDE || (LR && NPROD == 3), "wide windows exist on the upsample-after-logits path only");
  constexpr int NKB = SM::kNkb, BNT = SM::kBn, HS = SM::kHiStages, LS = SM::kLoStages;
  constexpr int CPT = BNT / 2;            // logit columns per softmax thread (row, column half): 128, WIDE: 64
  constexpr int NCHUNK = CPT / 32;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = ptx::smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  uint8_t* base_ptr = smem_raw + (base - raw);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;   // warp-uniform for the compiler
  const int half = blockIdx.x % NH, chunk = blockIdx.x / NH;
  const int u_begin = static_cast<int>(static_cast<int64_t>(chunk) * P.n_units / P.n_chunks);
  const int u_end = static_cast<int>(static_cast<int64_t>(chunk + 1) * P.n_units / P.n_chunks);
  const int tok0 = half * BNT;
  const int ntok = max(0, min(BNT, P.T - tok0));            // tokens this CTA owns
  const int n_mh = ntok > 128 ? 2 : 1;

  auto bar_qfull = [&](int s) { return base + SM::kBarOff + 8u * s; };               // unit's A operand landed (slot it % HS)
  auto bar_hiempty = [&](int s) { return base + SM::kBarOff + 8u * (3 + s); };        // hi slot: every MMA of the unit retired
  auto bar_loempty = [&](int s) { return base + SM::kBarOff + 8u * (6 + s); };        // lo slot: the unit's logit products retired
  const uint32_t bar_sfull = base + SM::kBarOff + 8u * 8;
  const uint32_t bar_dsfull = base + SM::kBarOff + 8u * 9;
  const uint32_t bar_mmadone = base + SM::kBarOff + 8u * 10;
  const uint32_t bar_sfree = base + SM::kBarOff + 8u * 13;
  auto bar_selfull = [&](uint32_t i) { return base + SM::kBarOff + 8u * (11 + (i & 1u)); };
  auto a_hi_slot = [&](uint32_t i) { return base + SM::kQOff + (i % HS) * SM::kAHalf; };
  auto a_lo_slot = [&](uint32_t i) { return base + SM::kLoOff + (i % LS) * SM::kAHalf; };
  volatile uint32_t* tmem_ptr = reinterpret_cast<volatile uint32_t*>(base_ptr + SM::kTmemPtrOff);
  volatile int* tok_s = reinterpret_cast<volatile int*>(base_ptr + SM::kTokOff);
  float* wxt = reinterpret_cast<float*>(base_ptr + SM::kWxtOff);
  float* stg = reinterpret_cast<float*>(base_ptr + SM::kDsOff);       // [128][64] fp32, chunk-swizzled

  if (threadIdx.x == 0) {
    for (int s = 0; s < 3; ++s) {
      ptx::mbar_init(bar_qfull(s), 1);
      ptx::mbar_init(bar_hiempty(s), 1);
    }
    for (int s = 0; s < 2; ++s) ptx::mbar_init(bar_loempty(s), 1);
    ptx::mbar_init(bar_sfull, 1);
    ptx::mbar_init(bar_dsfull, TC_SOFTMAX_THREADS);
    ptx::mbar_init(bar_mmadone, 1);
    ptx::mbar_init(bar_selfull(0), 1);
    ptx::mbar_init(bar_selfull(1), 1);
    ptx::mbar_init(bar_sfree, TC_SOFTMAX_THREADS);
    ptx::fence_barrier_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc(ptx::smem_u32(const_cast<uint32_t*>(tmem_ptr)), 512);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  const uint32_t tm_S = tmem_base, tm_dQ = tmem_base + 256, tm_dK = tmem_base + 320;
  auto tm_sel = [&](uint32_t i) { return tmem_base + 448u + 16u * (i & 1u); };     // S_sel, double-buffered by unit parity

  // ---- the products that consume the dS tile of unit (x, it), issued by one thread (LR: lane 0 of warp 2; staged: the S issuer)
  auto issue_post = [&](const BwUnit& x, uint32_t it, bool dk_fresh) {
    const int d = P.ops.d[x.l];
    const uint32_t ks = base + SM::kKOff, dsb = base + SM::kDsOff;
    const uint32_t qs = a_hi_slot(it);
    if (!LR) {
      // dQ' (128 px x d) = dS (K-major A, K = 256 tokens) * K_hi (MN-major B: rows = tokens)
      const uint32_t id_dq = bw_idesc(TC_BM, d, 0, 1);
      const int nkt = (ntok + 15) / 16;
      for (int k = 0; k < nkt; ++k)
        ptx::umma_bf16_elect(tm_dQ, bw_desc(dsb + (k >> 2) * 16384 + (k & 3) * 32, 16, 1024), bw_desc(ks + k * 2048, 16384, 1024),
                       id_dq, k != 0);
      // dK (128 tok x d per half) += dS^T (MN-major A: rows = pixels) * Q'_hi (MN-major B: rows = pixels)
      const uint32_t id_dk = bw_idesc(128, d, 1, 1);
      for (int mh = 0; mh < n_mh; ++mh)
        for (int k = 0; k < TC_BM / 16; ++k)
          ptx::umma_bf16_elect(tm_dK + mh * 64, bw_desc(dsb + mh * 2 * 16384 + k * 2048, 16384, 1024),
                         bw_desc(qs + k * 2048, 16384, 1024), id_dk, (dk_fresh && k == 0) ? 0u : 1u);
    } else {
      // dz (128 tok x 64 taps per half) += dS^T (MN-major A: rows = pixels) * W_hi (MN-major B: rows = pixels; the
      // A stage of this unit's S product).  Descriptors = the unit's base descriptor + a constant per k-step.
      const uint32_t id_dz = bw_idesc(128, WIDE ? 128 : TC_BK, 1, 1);    // WIDE: both 64-tap blocks of W_hi, 16 KB apart (LBO)
      const uint64_t b0 = bw_desc(qs, 16384, 1024);
      for (int mh = 0; mh < n_mh; ++mh) {
        const uint64_t a0 = bw_desc(dsb + mh * 2 * 16384, 16384, 1024);
#pragma unroll
        for (int k = 0; k < TC_BM / 16; ++k)
          ptx::umma_bf16_elect(tm_dK + mh * 64, a0 + static_cast<uint64_t>(k * (2048 >> 4)), b0 + static_cast<uint64_t>(k * (2048 >> 4)), id_dz,
                         (dk_fresh && k == 0) ? 0u : 1u);
      }
    }
    ptx::umma_commit_elect(bar_mmadone);
    ptx::umma_commit_elect(bar_hiempty(it % HS));
  };

  if (warp < BW_SOFTMAX_WARP0) {
  ptx::setmaxnreg_dec<BW_REGS_SMALL>();
  if (warp == 0) {
    // =========================== TMA producer (one lane) ===================================
    if (lane == 0) {
      // LR: the A stage holds the interpolation table W of the tile (it also is the B operand of the dz product, as Q'
      // is of dK), the "k" / "k_sel" buffers hold the tile's window of native-resolution logits of the CTA's 256 tokens
      // / of the selected tokens: reloaded when the window moves (every R/s tiles) instead of once per head.
      int prev_hg = -1, prev_b = -1, prev_r0 = -1;
      uint32_t it = 0;
      BwUnit x = bw_decode(P, u_begin);
      for (int u = u_begin; u < u_end; ++u, ++it, x = bw_next(P, x)) {
        const int h = P.ops.heads[x.l];
        const int st = it % HS;
        ptx::mbar_wait_hint(bar_hiempty(st), ((it / HS) & 1u) ^ 1u, 2000u);
        if (NPROD == 3) ptx::mbar_wait_hint(bar_loempty(it % LS), ((it / LS) & 1u) ^ 1u, 2000u);
        const int r0 = LR ? x.r0 : 0;
        const bool need_k = LR ? (x.hg != prev_hg || x.b != prev_b || r0 != prev_r0)
                               : (x.hg != prev_hg || (!P.ops.k_shared[x.l] && x.b != prev_b));
        const bool need_ksel = need_k || x.b != prev_b;
        // previous unit done with k / k_sel.  LR: only the logit products read them (lo_empty: S and S_sel of the previous unit
        // have retired, early in its dS pass), so a new window is fetched a unit ahead like the tables; staged path: the dQ'
        // product reads k too (hi_empty: all MMAs of the previous unit)
        if (need_ksel && it > 0) {
          if (LR && !WIDE) ptx::mbar_wait_hint(bar_loempty((it - 1) % LS), ((it - 1) / LS) & 1u, 2000u);
          else ptx::mbar_wait_hint(bar_hiempty((it - 1) % HS), ((it - 1) / HS) & 1u, 2000u);
        }
        const uint32_t bytes = SM::kQStage + (need_k ? SM::kKBytes : 0) + (need_ksel ? SM::kKselBytes : 0);
        ptx::mbar_arrive_expect_tx(bar_qfull(st), bytes);
        const int qz = x.b * h + x.head;
        const uint32_t qs = a_hi_slot(it), qs_lo = a_lo_slot(it);
        const int a_row = LR ? 0 : x.tile * TC_BM, a_z = LR ? x.tile : qz;     // (k, row, batch) of the A tile
        const int b_k0 = LR ? r0 * P.s[x.l] : 0;                              // first tap of the window in the s*s row
#pragma unroll
        for (int kb = 0; kb < NKB; ++kb) {
          ptx::tma_load_3d(qs + kb * TC_A_BYTES, &P.ops.q_map[x.l][0], bar_qfull(st), kb * TC_BK, a_row, a_z);
          if (NPROD == 3) ptx::tma_load_3d(qs_lo + kb * TC_A_BYTES, &P.ops.q_map[x.l][1], bar_qfull(st), kb * TC_BK, a_row, a_z);
        }
        if (need_k) {
          const int kz = LR ? qz : (P.ops.k_shared[x.l] ? 0 : x.b * h) + x.head;
#pragma unroll
          for (int kb = 0; kb < NKB; ++kb) {
            ptx::tma_load_3d(base + SM::kKOff + kb * SM::kKBlock, &P.ops.k_map[x.l][0], bar_qfull(st), b_k0 + kb * TC_BK, tok0, kz);
            if (NPROD == 3)
              ptx::tma_load_3d(base + SM::kKOff + SM::kKHalf + kb * SM::kKBlock, &P.ops.k_map[x.l][1], bar_qfull(st), b_k0 + kb * TC_BK, tok0, kz);
          }
        }
        if (need_ksel) {
#pragma unroll
          for (int kb = 0; kb < NKB; ++kb) {
            ptx::tma_load_3d(base + SM::kKselOff + kb * BW_KSEL_BYTES, &P.ksel_map[x.l][0], bar_qfull(st), b_k0 + kb * TC_BK, 0, qz);
            if (NPROD == 3)
              ptx::tma_load_3d(base + SM::kKselOff + SM::kKselHalf + kb * BW_KSEL_BYTES, &P.ksel_map[x.l][1], bar_qfull(st), b_k0 + kb * TC_BK, 0, qz);
          }
        }
        prev_hg = x.hg;
        prev_b = x.b;
        prev_r0 = r0;
      }
    }
  } else if (warp == 1) {
    // =========================== MMA issuer: logit products (whole warp, elected issue) ========
    {
      uint32_t it = 0;
      long long pf_q = 0, pf_ds = 0, pf_is = 0;      // profiler (SKP_BWD_DBG & 8), kept in registers
      BwUnit x = bw_decode(P, u_begin);
      const uint32_t ks = base + SM::kKOff, ksel = base + SM::kKselOff;
      // ---- S = Q' K^T (256 tokens) and S_sel = Q' K_sel^T (16 tokens) of the unit in stage (jt & 1)
      // mode 0: both, S_sel interleaved into the first half of S; 1: S_sel only; 2: S only
      auto issue_S = [&](int l, uint32_t jt, int mode) {
        const uint32_t qs = a_hi_slot(jt), ql = a_lo_slot(jt);
        const uint32_t tsel = tm_sel(jt), selbar = bar_selfull(jt);
        long long c0 = BW_CLK();
        ptx::mbar_wait_hint(bar_qfull(jt % HS), (jt / HS) & 1u, 2000u);
        ptx::tc_fence_after();
        pf_q += BW_CLK() - c0;
        c0 = BW_CLK();
        const uint32_t id_S = bw_idesc(TC_BM, BNT, 0, 0), id_sel = bw_idesc(TC_BM, BW_SEL, 0, 0);
        // byte offset of 16-wide k-step kk inside an operand whose 64-wide k-blocks are `blk` bytes apart
        auto ko = [](int kk, uint32_t blk) { return static_cast<uint32_t>(kk >> 2) * blk + static_cast<uint32_t>(kk & 3) * 32u; };
        // LR: the K dimension is the 64-tap window (the launcher only sends such layers here): fixed trip count, the
        // descriptors of the unrolled loops are the stage's base descriptors plus constants
        const int nk16 = LR ? NKB * (TC_BK / 16) : P.ops.d[l] / 16;
        if (mode == 1) {
#pragma unroll
          for (int kk = 0; kk < nk16; ++kk) {
            const uint64_t a_hi = bw_desc(qs + ko(kk, TC_A_BYTES), 16, 1024), s_hi = bw_desc(ksel + ko(kk, BW_KSEL_BYTES), 16, 1024);
            if (NPROD == 3) {
              const uint64_t a_lo = bw_desc(ql + ko(kk, TC_A_BYTES), 16, 1024);
              const uint64_t s_lo = bw_desc(ksel + SM::kKselHalf + ko(kk, BW_KSEL_BYTES), 16, 1024);
              ptx::umma_bf16_elect(tsel, a_hi, s_lo, id_sel, kk != 0);
              ptx::umma_bf16_elect(tsel, a_lo, s_hi, id_sel, 1u);
              ptx::umma_bf16_elect(tsel, a_hi, s_hi, id_sel, 1u);
            } else {
              ptx::umma_bf16_elect(tsel, a_hi, s_hi, id_sel, kk != 0);
            }
          }
          ptx::umma_commit_elect(selbar);
          pf_is += BW_CLK() - c0;
          return;
        }
        if (mode == 2) {
#pragma unroll
          for (int kk = 0; kk < nk16; ++kk) {
            const uint64_t a_hi = bw_desc(qs + ko(kk, TC_A_BYTES), 16, 1024), b_hi = bw_desc(ks + ko(kk, SM::kKBlock), 16, 1024);
            if (NPROD == 3) {
              const uint64_t a_lo = bw_desc(ql + ko(kk, TC_A_BYTES), 16, 1024);
              const uint64_t b_lo = bw_desc(ks + SM::kKHalf + ko(kk, SM::kKBlock), 16, 1024);
              ptx::umma_bf16_elect(tm_S, a_hi, b_lo, id_S, kk != 0);
              ptx::umma_bf16_elect(tm_S, a_lo, b_hi, id_S, 1u);
              ptx::umma_bf16_elect(tm_S, a_hi, b_hi, id_S, 1u);
            } else {
              ptx::umma_bf16_elect(tm_S, a_hi, b_hi, id_S, kk != 0);
            }
          }
          ptx::umma_commit_elect(bar_sfull);
          ptx::umma_commit_elect(bar_loempty(jt % LS));
          pf_is += BW_CLK() - c0;
          return;
        }
        // The selected tokens' logits go out two per 256-token MMA during the first half of the product and get their
        // own barrier: the softmax warps turn them into the row's delta (16 serial exponentials) while the second half
        // still runs, and the N = 16 MMAs (latency-bound on one accumulator) hide under the N = 256 ones.
        const int n_selblk = (nk16 + 1) >> 1;       // S k-blocks that carry two S_sel k-blocks each
#pragma unroll
        for (int kk = 0; kk < nk16; ++kk) {
          const uint64_t a_hi = bw_desc(qs + kk * 32, 16, 1024), b_hi = bw_desc(ks + kk * 32, 16, 1024);
          const int k0 = 2 * kk, k1 = 2 * kk + 1;
          const bool sel0 = kk < n_selblk, sel1 = sel0 && k1 < nk16;
          const uint64_t p_hi = bw_desc(qs + k0 * 32, 16, 1024), r_hi = bw_desc(qs + k1 * 32, 16, 1024);
          const uint64_t s0_hi = bw_desc(ksel + k0 * 32, 16, 1024), s1_hi = bw_desc(ksel + k1 * 32, 16, 1024);
          if (NPROD == 3) {
            const uint64_t a_lo = bw_desc(ql + kk * 32, 16, 1024);
            const uint64_t b_lo = bw_desc(ks + SM::kKHalf + kk * 32, 16, 1024);
            const uint64_t p_lo = bw_desc(ql + k0 * 32, 16, 1024), r_lo = bw_desc(ql + k1 * 32, 16, 1024);
            const uint64_t s0_lo = bw_desc(ksel + SM::kKselHalf + k0 * 32, 16, 1024);
            const uint64_t s1_lo = bw_desc(ksel + SM::kKselHalf + k1 * 32, 16, 1024);
            ptx::umma_bf16_elect(tm_S, a_hi, b_lo, id_S, kk != 0);
            if (sel0) {
              ptx::umma_bf16_elect(tsel, p_hi, s0_lo, id_sel, kk != 0);
              ptx::umma_bf16_elect(tsel, p_lo, s0_hi, id_sel, 1u);
            }
            ptx::umma_bf16_elect(tm_S, a_lo, b_hi, id_S, 1u);
            if (sel0) ptx::umma_bf16_elect(tsel, p_hi, s0_hi, id_sel, 1u);
            if (sel1) ptx::umma_bf16_elect(tsel, r_hi, s1_lo, id_sel, 1u);
            ptx::umma_bf16_elect(tm_S, a_hi, b_hi, id_S, 1u);
            if (sel1) {
              ptx::umma_bf16_elect(tsel, r_lo, s1_hi, id_sel, 1u);
              ptx::umma_bf16_elect(tsel, r_hi, s1_hi, id_sel, 1u);
            }
          } else {
            ptx::umma_bf16_elect(tm_S, a_hi, b_hi, id_S, kk != 0);
            if (sel0) ptx::umma_bf16_elect(tsel, p_hi, s0_hi, id_sel, kk != 0);
            if (sel1) ptx::umma_bf16_elect(tsel, r_hi, s1_hi, id_sel, 1u);
          }
          if (kk + 1 == n_selblk) ptx::umma_commit_elect(selbar);
        }
        ptx::umma_commit_elect(bar_sfull);
        ptx::umma_commit_elect(bar_loempty(jt % LS));
        pf_is += BW_CLK() - c0;
      };
      // LR: two issuing threads.  This one: the logit products S / S_sel; lane 0 of warp 2: the dz product that consumes dS.
      // One thread issuing everything spent ~100 cycles per small MMA (N = 16 / 64: 48 on the tensor pipe,
      // scripts/mma_small_probe.cu) computing descriptors and moving them to uniform registers between the instructions,
      // with S(u+1) queued behind that.  Staged path: this thread issues dQ' / dK too, AHEAD of S(u+1) (the softmax warps
      // drain dQ' right after their dS store; S(u+1) first was measured 7 % slower on SDXL's four 32^2 layers).
      bool s_issued = false;       // S of this unit went out early, in the previous iteration
      bool dk_fresh = true;
      for (int u = u_begin; u < u_end; ++u, ++it) {
        const BwUnit nxu = bw_next(P, x);
        BW_TR(0)
        if (!s_issued) {
          if (LR && it > 0) {
            // the accumulator is free once every softmax thread holds S(u-1) in registers
            ptx::mbar_wait_hint(bar_sfree, (it - 1) & 1u, 1000u);
            ptx::tc_fence_after();
          }
          if (WIDE) {                 // two k-blocks: S_sel first (its barrier frees the softmax warps' delta), then S
            issue_S(x.l, it, 1);
            issue_S(x.l, it, 2);
          } else {
            issue_S(x.l, it, 0);
          }
        }
        s_issued = false;
        // LR: the next unit's S_sel runs on the idle tensor pipe while the softmax warps are in this unit's dS pass -- also when
        // it interpolates from another window (the producer refills k / k_sel as soon as this unit's logit products have
        // retired; waiting for the dz product first cost ~5 k cycles per window change).  WIDE: one operand stage -- the next
        // unit's interpolation table can only be loaded once this unit's dz product has read the current one.
        const bool early = LR && !WIDE && u + 1 < u_end;
        if (early) {
          issue_S(nxu.l, it + 1, 1);
          BW_TR(1)
          // ... and its S as soon as every softmax thread holds S(u) in registers (s_free): under the whole dS pass of unit u
          long long c9 = BW_CLK();
          ptx::mbar_wait_hint(bar_sfree, it & 1u, 1000u);
          ptx::tc_fence_after();
          BW_TR(2)
          pf_ds += BW_CLK() - c9;
          issue_S(nxu.l, it + 1, 2);
          BW_TR(3)
          s_issued = true;
        }
        if (!LR) {
          ptx::mbar_wait_hint(bar_dsfull, it & 1u, 2000u);
          ptx::tc_fence_after();
          issue_post(x, it, dk_fresh);
          dk_fresh = bw_flush_between<false>(P, x, nxu, u + 1 >= u_end);
        }
        x = nxu;
      }
      if (P.prof && lane == 0 && blockIdx.x == ((BW_DBG(P) >> 6) & 1)) {
        P.prof[0] = pf_q; P.prof[1] = pf_ds; P.prof[8] = pf_is;
      }
    }
  } else if (warp == 2) {
    // =========================== LR: MMA issuer of the dz product (whole warp, elected issue) ==
    if (LR) {
      uint32_t it = 0;
      bool dk_fresh = true;
      BwUnit x = bw_decode(P, u_begin);
      for (int u = u_begin; u < u_end; ++u, ++it) {
        const BwUnit nxu = bw_next(P, x);
        // dS (bf16, [px][tok]) in shared memory.  The operand stage it is multiplied with was waited for by the S issuer
        // (q_full): S(u) read it before the softmax warps could produce dS(u).
        ptx::mbar_wait_hint(bar_dsfull, it & 1u, 2000u);
        ptx::tc_fence_after();
        BW_TR(4)
        issue_post(x, it, dk_fresh);
        BW_TR(5)
        dk_fresh = bw_flush_between<true>(P, x, nxu, u + 1 >= u_end);
        x = nxu;
      }
    }
  }
  } else {
    ptx::setmaxnreg_inc<BW_REGS_SOFTMAX>();
    // =========================== softmax-recompute / dS / drain warps ========================
    const int quad = warp & 3;                 // TMEM lane quadrant this warp may access
    const int ch = (warp - BW_SOFTMAX_WARP0) >> 2;   // column half of the CTA's 256 tokens
    const int row = quad * 32 + lane;
    const int stid = threadIdx.x - BW_SOFTMAX_WARP0 * 32;   // 0..255
    const int col0 = ch * CPT;
    const int ncol = max(0, min(CPT, ntok - col0));
    const uint32_t lane_off = static_cast<uint32_t>(quad * 32) << 16;
    const uint32_t dsb = base + SM::kDsOff;
    uint8_t* ds_ptr = base_ptr + SM::kDsOff;
    const int R = P.R;
    uint32_t it = 0;
    int prev_l = -1, prev_b = -1;
    uint32_t poff[BW_SEL];                     // byte offset of (my row, selected plane j's token) in the dS tile (planes of my columns)
    uint32_t valid_mask = 0, mine_mask = 0;    // bit j: plane j names a token (< T) / a token of this thread's columns
    float2 nx_ml = make_float2(0.f, 1.f);
    float nx_dp[BW_SEL];
    auto load_row = [&](const BwUnit& y, float2& ml, float (&dpr)[BW_SEL]) {
      const int pp = y.tile * TC_BM + row;
      ml = __ldg(reinterpret_cast<const float2*>(P.row_stats) + (static_cast<int64_t>(y.b) * P.heads_all + P.ghead[y.l] + y.head) * P.R2 + pp);
      // upstream gradients of the row: one 64-byte row of the pixel-major copy (4 x 16-byte loads instead of 16 scalar
      // loads from 16 planes)
      const float4* g4 = reinterpret_cast<const float4*>(P.dpT) + (static_cast<int64_t>(y.b) * P.R2 + pp) * (BW_SEL / 4);
#pragma unroll
      for (int q = 0; q < BW_SEL / 4; ++q) {
        const float4 g = __ldg(g4 + q);
        dpr[4 * q] = g.x; dpr[4 * q + 1] = g.y; dpr[4 * q + 2] = g.z; dpr[4 * q + 3] = g.w;
      }
    };
    long long pf[6] = {0, 0, 0, 0, 0, 0}, c_end = BW_CLK();   // profiler (SKP_BWD_DBG & 8), kept in registers
    bool mma_pending = false;      // LR: the wait for the previous unit's dz product is deferred to the first dS store
    bool flush_pending = false;    // LR: a window flush (TMA reduce out of the dS tile's bytes) may still be reading shared memory
    BwUnit x = bw_decode(P, u_begin), nxu = x;
    float wgt = 0.0f;                            // P.ops.weight[x.l], re-read when the layer changes
    float m = 0.0f, nd = 0.0f;                   // row prologue of the current unit (row_prologue below)
    uint32_t pv[BW_SEL / 2];
    bool have_row = false;                       // ... already computed during the previous unit
    if (u_begin < u_end) load_row(x, nx_ml, nx_dp);
    // The scalar work of a unit's top (next tuple, its window row, the addresses of the next row's statistics and upstream
    // gradients) is a ~1.2 k-cycle chain of dependent parameter-bank reads, 64-bit multiplies and reconvergence scopes
    // that every softmax warp repeats in lockstep with the MUFU pipe idle: it now sits behind the first chunk of the dense
    // pass, where the exponentials leave half the issue slots free (PREP below).
    for (int u = u_begin; u < u_end; ++u, ++it, x = nxu) {
      BW_TR(0)
      const int d = P.ops.d[x.l], s = P.s[x.l], C = P.C[x.l];
      const int up = ((R & (R - 1)) | (s & (s - 1))) == 0 ? R >> (__ffs(s) - 1) : R / s, nj = 5 * up;
      const int njp = (nj + 3) & ~3;            // table row pitch: the gather loops run in groups of four taps
      const float scale = P.ops.scale[x.l];
      if (x.l != prev_l || x.b != prev_b) {
        wgt = P.ops.weight[x.l];
        named_bar_sync(1, TC_SOFTMAX_THREADS);
        if (!LR && x.l != prev_l) {
          const float ratio = static_cast<float>(s) / R;
          for (int e = stid; e < s * njp; e += TC_SOFTMAX_THREADS) {
            const int sx = e / njp, jj = e % njp, xx = max(0, (sx - 2) * up) + jj;
            float wsum = 0.0f;
            if (jj < nj && xx < R) {
              int ix;
              float w[4];
              bicubic_coeffs(xx, ratio, &ix, w);
#pragma unroll
              for (int a = 0; a < 4; ++a) wsum += (min(max(ix - 1 + a, 0), s - 1) == sx) ? w[a] : 0.0f;
            }
            wxt[e] = wsum;
          }
        }
        if (stid < BW_SEL) tok_s[stid] = stid < P.n_idx ? static_cast<int>(P.token_idx[x.b * P.idx_stride + stid]) : -1;
        named_bar_sync(1, TC_SOFTMAX_THREADS);
        // the range tests and the dS-tile addresses of the two per-unit loops over the selected planes, once per image
        // instead of once per unit (registers: the per-unit loops read them 32 times)
        valid_mask = mine_mask = 0;
#pragma unroll
        for (int j = 0; j < BW_SEL; ++j) {
          const int t = tok_s[j], tl = t - tok0;
          if (t >= 0 && t < P.T) valid_mask |= 1u << j;
          if (j < P.n_idx && tl >= col0 && tl < col0 + ncol) mine_mask |= 1u << j;
          poff[j] = dsb + static_cast<uint32_t>((tl >> 6) * 16384 + row * 128 + ((((tl & 63) >> 3) ^ (row & 7)) << 4) + (tl & 7) * 2);
        }
        prev_l = x.l;
        prev_b = x.b;
      }
      // (m, l) and the n_idx upstream gradients of this row were fetched one unit ahead (nx_*): a unit
      // is too short to hide a DRAM round trip behind itself.
      // Row prologue of unit `jt` from nx_*: m, the row's delta = sum_j P_sel[j] dP[j] (as nd = -delta / l) and dS of the selected
      // planes' tokens = P (dP - delta), packed: 8 registers live across the dense pass instead of 32.  LR: for the next unit
      // it runs inside the last chunk of this unit's dense pass (its S_sel went out a unit ahead: loaded before chunk 2), where
      // the TMEM round trip, 16 exponentials and dependent sums (~0.6 k cycles at the top of a unit with the MUFU pipe idle)
      // fill issue slots; at the top only for the first unit and after an image / layer change (masks and weight change:
      // what was computed ahead is dropped).
      auto prologue_load = [&](uint32_t jt, uint32_t (&sv)[BW_SEL]) {
        ptx::mbar_wait(bar_selfull(jt), (jt >> 1) & 1u);
        ptx::tc_fence_after();
        ptx::tmem_ld16(tm_sel(jt) + lane_off, sv);
      };
      // plain arithmetic (no volatile asm): the compiler may interleave it with the exponentials of a dense chunk
      auto prologue_math = [&](const uint32_t (&sv)[BW_SEL], float w, float& m_o, float& nd_o, uint32_t (&pv_o)[BW_SEL / 2]) {
        const float mm = nx_ml.x, inv_l = ptx::rcp_approx(nx_ml.y);      // l is a normal number on both forward paths (max-subtracted: l >= 1; constant shift: l > 2^-118): no special cases
        float psel[BW_SEL], dp[BW_SEL];           // P and weighted upstream gradient of the selected planes' tokens (this row)
        float dl[4] = {0.0f, 0.0f, 0.0f, 0.0f};     // four chains instead of one 16-deep dependent FFMA chain
#pragma unroll
        for (int j = 0; j < BW_SEL; ++j) {
          dp[j] = w * nx_dp[j];
          psel[j] = ((valid_mask >> j) & 1u) ? ptx::ex2(__uint_as_float(sv[j]) - mm) * inv_l : 0.0f;
          dl[j & 3] = fmaf(psel[j], dp[j], dl[j & 3]);
        }
        const float delta = (dl[0] + dl[1]) + (dl[2] + dl[3]);
#pragma unroll
        for (int j = 0; j < BW_SEL / 2; ++j) {
          const __nv_bfloat162 h2 = __floats2bfloat162_rn(psel[2 * j] * (dp[2 * j] - delta), psel[2 * j + 1] * (dp[2 * j + 1] - delta));
          pv_o[j] = *reinterpret_cast<const uint32_t*>(&h2);
        }
        m_o = mm;
        nd_o = -delta * inv_l;
      };
      long long c1 = BW_CLK();
      pf[5] += c1 - c_end;                      // loop top: decode, prefetch issue, table rebuilds
      BW_TR(1)
      if (!have_row) {
        uint32_t sv[BW_SEL];
        prologue_load(it, sv);
        ptx::tmem_ld_wait();
        prologue_math(sv, wgt, m, nd, pv);
      }
      BW_TR(2)
      pf[0] += BW_CLK() - c1;
      c1 = BW_CLK();
      // ---- dS = P * (dP - delta): dense part -P*delta for my 128 columns, then the n_idx corrections
      BW_TR(3)
      {
        long long c2 = BW_CLK();
        ptx::mbar_wait(bar_sfull, it & 1u);
        ptx::tc_fence_after();
        BW_TR(4)
        const long long c3 = BW_CLK();
        pf[0] += c3 - c2;
        c1 += c3 - c2;                          // the dS slot of the profile excludes this wait
      }
      constexpr bool AHEAD = LR && !WIDE;       // the next unit's row prologue runs inside this unit's last dense chunk
      float m_n = 0.0f, nd_n = 0.0f;
      uint32_t pv_n[BW_SEL / 2];
      {
        // LR: the thread's whole row segment of S goes to registers at once and the accumulator is handed back (s_free): the
        // S product of the next unit then runs under this unit's entire dS pass instead of under its last chunk (with two
        // 32-column buffers the softmax warps waited 0.6 - 2 k cycles per unit for S once their loop top was out of the way)
        constexpr int NBUF = LR ? NCHUNK : 2;
        uint32_t v[NBUF][32];
        uint32_t sv_n[BW_SEL];
#pragma unroll
        for (int j = 0; j < BW_SEL; ++j) sv_n[j] = 0u;
        // LR: the first HOLD chunks of dS stay packed in registers until the previous unit's dz product has retired
        // (it still reads the single dS tile): the exponentials of those chunks run under that product
        constexpr int HOLD = LR ? (BW_LR_HOLD < NCHUNK ? BW_LR_HOLD : NCHUNK) : 1;
        uint32_t pk[NCHUNK][16];
        auto store_chunk = [&](int c) {
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const int cc = col0 + 32 * c + 8 * q;
            const uint32_t off = (cc >> 6) * 16384 + row * 128 + ((((cc & 63) >> 3) ^ (row & 7)) << 4);
            *reinterpret_cast<uint4*>(ds_ptr + off) = make_uint4(pk[c][4 * q], pk[c][4 * q + 1], pk[c][4 * q + 2], pk[c][4 * q + 3]);
          }
        };
        if (LR) {
#pragma unroll
          for (int c = 0; c < NCHUNK; ++c) ptx::tmem_ld32(tm_S + lane_off + col0 + 32 * c, v[c]);
          ptx::tmem_ld_wait();
          // S(u) is in registers: the tensor pipe may overwrite the accumulator with S(u+1)
          ptx::tc_fence_before();
          ptx::mbar_arrive(bar_sfree);
          BW_TR(7)
        } else {
          ptx::tmem_ld32(tm_S + lane_off + col0, v[0]);
        }
#pragma unroll
        for (int c = 0; c < NCHUNK; ++c) {
          if (!LR) {
            ptx::tmem_ld_wait();
            if (c < NCHUNK - 1) ptx::tmem_ld32(tm_S + lane_off + col0 + 32 * (c + 1), v[(c + 1) & 1]);
          }
          if (AHEAD && c == NCHUNK - 1 - BW_PROLOGUE_LEAD && u + 1 < u_end) prologue_load(it + 1, sv_n);
          if (AHEAD && c == NCHUNK - 1) {
            ptx::tmem_ld_wait();
            prologue_math(sv_n, wgt, m_n, nd_n, pv_n);
          }
          if ((LR && !WIDE) || 32 * (c + 1) <= ncol) {
            // all 32 columns of the chunk are tokens (all but the chunk holding T's end): no per-column range test.
            // LR: none at all -- columns past T hold logits of zero-filled window rows, their dS only reaches the dz rows of
            // those non-existent tokens (M dimension of the product), which the flush's tensor map clips
            // (the staged path sums over tokens in dQ', the WIDE flush writes rows itself: both need the zeros)
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const float e0 = ptx::ex2(__uint_as_float(v[c % NBUF][2 * j]) - m) * nd;
              const float e1 = ptx::ex2(__uint_as_float(v[c % NBUF][2 * j + 1]) - m) * nd;
              const __nv_bfloat162 h2 = __floats2bfloat162_rn(e0, e1);
              pk[c][j] = *reinterpret_cast<const uint32_t*>(&h2);
            }
          } else {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              float e0 = ptx::ex2(__uint_as_float(v[c % NBUF][2 * j]) - m) * nd;
              float e1 = ptx::ex2(__uint_as_float(v[c % NBUF][2 * j + 1]) - m) * nd;
              e0 = (32 * c + 2 * j < ncol) ? e0 : 0.0f;
              e1 = (32 * c + 2 * j + 1 < ncol) ? e1 : 0.0f;
              const __nv_bfloat162 h2 = __floats2bfloat162_rn(e0, e1);
              pk[c][j] = *reinterpret_cast<const uint32_t*>(&h2);
            }
          }
          if (c == BW_PREP_CHUNK) {
            // PREP: the next unit's tuple; LR: its row statistics / upstream gradients go out now and land under the rest
            // of this unit (there is no long MMA wait at the end of an LR unit to hide them).  Unconditional (the last unit
            // re-reads its own row): a load under a branch made the compiler copy the loaded registers right behind it,
            // i.e. wait for the whole memory round trip (780 cycles per unit).
            nxu = bw_next(P, x);
            if (LR) load_row(u + 1 < u_end ? nxu : x, nx_ml, nx_dp);
            BW_TR(5)
          }
          if (c + 1 == HOLD) {
            if (LR && mma_pending) {
              ptx::mbar_wait(bar_mmadone, (it - 1) & 1u);
              ptx::tc_fence_after();
              mma_pending = false;
            }
            if (LR && !WIDE && flush_pending) {
              // the previous unit's window flush was staged in the dS tile: the TMA engine must have read it
              if (stid == 0) ptx::bulk_wait_group_read0();
              named_bar_sync(1, TC_SOFTMAX_THREADS);
              flush_pending = false;
            }
#pragma unroll
            for (int cc = 0; cc <= c; ++cc) store_chunk(cc);
            BW_TR(6)
          } else if (c + 1 > HOLD) {
            store_chunk(c);
          }
        }
      }
      BW_TR(8)
      // selected planes' tokens of my columns: dS = P (dP - delta), from the fp32 P of the S_sel product.  A predicated 16-bit
      // STORE over the dense pass's -P delta (same thread, program order), not a read-modify-write of the bf16 tile: sixteen
      // reconvergence scopes with a dependent LDS -> add -> pack -> STS chain each were ~19 % of the softmax warps' samples
#pragma unroll
      for (int j = 0; j < BW_SEL; ++j) {
        const unsigned short hv = static_cast<unsigned short>((j & 1) ? pv[j >> 1] >> 16 : pv[j >> 1] & 0xffffu);
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %0, 0;\n\t@p st.shared.b16 [%1], %2;\n\t}" ::"r"((mine_mask >> j) & 1u),
                     "r"(poff[j]), "h"(hv) : "memory");
      }
      pf[1] += BW_CLK() - c1;
      c1 = BW_CLK();
      ptx::fence_proxy_async();
      ptx::tc_fence_before();
      ptx::mbar_arrive(bar_dsfull);
      BW_TR(9)
      pf[2] += BW_CLK() - c1;
      c1 = BW_CLK();

      // next unit's row statistics / upstream gradients: issued here, where this thread would only spin
      // on the dQ/dK MMAs, they are back long before the next loop top
      if (!LR) load_row(u + 1 < u_end ? nxu : x, nx_ml, nx_dp);
      // ---- dQ' tile: TMEM -> fp32 staging (aliases the dS tile, free once the MMAs are done)
      const bool lr_flush = LR && bw_flush_between<true>(P, x, nxu, u + 1 >= u_end);
      if (!LR || lr_flush) {
        ptx::mbar_wait(bar_mmadone, it & 1u);
        ptx::tc_fence_after();
      } else {
        mma_pending = true;
      }
      BW_TR(11)
      pf[3] += BW_CLK() - c1;
      c1 = BW_CLK();
      if (!LR && ch * 32 < d && !(BW_DBG(P) & 4)) {
        uint32_t v[32];
        ptx::tmem_ld32(tm_dQ + lane_off + ch * 32, v);
        ptx::tmem_ld_wait();
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const int cidx = ch * 8 + q;
          *reinterpret_cast<float4*>(stg + row * BW_STG_LD + (cidx << 2)) =
              make_float4(__uint_as_float(v[4 * q]) * scale, __uint_as_float(v[4 * q + 1]) * scale,
                          __uint_as_float(v[4 * q + 2]) * scale, __uint_as_float(v[4 * q + 3]) * scale);
        }
      }
      if (!LR) {
      ptx::tc_fence_before();
      named_bar_sync(2, TC_SOFTMAX_THREADS);
      // ---- adjoint of the bicubic upsample: x-gather from staging, y taps as vector reductions
      {
        const int span = R <= TC_BM ? R : TC_BM;
        const int rows_per_tile = TC_BM / span;
        const int p0 = x.tile * TC_BM;
        // R, s and d/4 are powers of two for every supported model: shifts instead of ~150-cycle divisions
        const bool p2 = ((R & (R - 1)) | (s & (s - 1)) | ((d >> 2) & ((d >> 2) - 1))) == 0;
        const int lg_c4 = __ffs(d >> 2) - 1, lg_s = __ffs(s) - 1;
        const int y0 = p2 ? p0 >> (__ffs(R) - 1) : p0 / R, x0 = p2 ? p0 & (R - 1) : p0 % R;
        const int n_c4 = d >> 2;
        const int n_out = rows_per_tile * s * n_c4;
        const float ratio = static_cast<float>(s) / R;
        float* dq = P.dq[x.l] + static_cast<int64_t>(x.b) * s * s * C + x.head * d;
        for (int o = stid; o < ((BW_DBG(P) & 2) ? 0 : n_out); o += TC_SOFTMAX_THREADS) {
          const int c4 = p2 ? o & (n_c4 - 1) : o % n_c4;
          const int sx = p2 ? (o >> lg_c4) & (s - 1) : (o / n_c4) % s;
          const int yy = p2 ? o >> (lg_c4 + lg_s) : o / (n_c4 * s);
          const int xlo = max(0, (sx - 2) * up);
          // taps outside the tile's x span get weight 0 and a clamped (valid) row: fixed trip count,
          // four independent accumulators so the shared-memory loads pipeline
          float4 a4[4];
#pragma unroll
          for (int g = 0; g < 4; ++g) a4[g] = make_float4(0.f, 0.f, 0.f, 0.f);
          const int j0 = max(0, x0 - xlo), j1 = min(nj, x0 + span - xlo);
          const float* wrow = wxt + sx * njp;
          if (span == R && R == TC_BM) {
            // one whole image row per tile: every tap of the table is inside the tile (taps past the row
            // end have weight 0 and read finite bytes of the dS area), addresses are base + immediates
            const float4* src = reinterpret_cast<const float4*>(stg) + xlo * (BW_STG_LD / 4) + c4;
#pragma unroll 2
            for (int jb = 0; jb < nj; jb += 4) {
#pragma unroll
              for (int g = 0; g < 4; ++g) {
                const float w = wrow[jb + g];
                const float4 t4 = src[(jb + g) * (BW_STG_LD / 4)];
                a4[g].x = fmaf(w, t4.x, a4[g].x);
                a4[g].y = fmaf(w, t4.y, a4[g].y);
                a4[g].z = fmaf(w, t4.z, a4[g].z);
                a4[g].w = fmaf(w, t4.w, a4[g].w);
              }
            }
          } else {
            for (int jb = 0; jb < nj; jb += 4) {
#pragma unroll
              for (int g = 0; g < 4; ++g) {
                const int j = jb + g;
                const bool ok = j >= j0 && j < j1;
                const float w = ok ? wrow[j] : 0.0f;
                const int r = yy * span + min(max(xlo + j - x0, 0), span - 1);
                const float4 t4 = *reinterpret_cast<const float4*>(stg + r * BW_STG_LD + (c4 << 2));
                a4[g].x = fmaf(w, t4.x, a4[g].x);
                a4[g].y = fmaf(w, t4.y, a4[g].y);
                a4[g].z = fmaf(w, t4.z, a4[g].z);
                a4[g].w = fmaf(w, t4.w, a4[g].w);
              }
            }
          }
          float4 acc;
          acc.x = (a4[0].x + a4[1].x) + (a4[2].x + a4[3].x);
          acc.y = (a4[0].y + a4[1].y) + (a4[2].y + a4[3].y);
          acc.z = (a4[0].z + a4[1].z) + (a4[2].z + a4[3].z);
          acc.w = (a4[0].w + a4[1].w) + (a4[2].w + a4[3].w);
          int iy;
          float wy[4];
          bicubic_coeffs(y0 + yy, ratio, &iy, wy);
#pragma unroll
          for (int a = 0; a < 4; ++a) {
            const int sy = min(max(iy - 1 + a, 0), s - 1);
            if (!(BW_DBG(P) & 1)) red_add_v4(dq + static_cast<int64_t>(sy * s + sx) * C + c4 * 4, wy[a] * acc.x, wy[a] * acc.y, wy[a] * acc.z, wy[a] * acc.w);
          }
        }
      }
      named_bar_sync(2, TC_SOFTMAX_THREADS);
      BW_TR(10)
      }
      pf[4] += BW_CLK() - c1;
      c_end = BW_CLK();
      if (LR) {
        // ---- dz: flush the window accumulator (token rows x 64 taps) when the next tile leaves this window
        // thread (row, ch): token row of the accumulator's lane; 64 of its tap columns (WIDE: tap half ch of the one
        // 128-token accumulator; else all 64 taps of token half ch)
        if (!WIDE && lr_flush) {
          // The accumulator goes through shared memory to the TMA engine, which adds it into dz at L2 (cp.reduce.async.bulk):
          // token half ch = two [128 tokens][32 taps] fp32 tiles (128 B rows, 128B swizzle) in the bytes of the dS tile, free
          // once the dz product has retired (waited for above).  As red.global.add.v4.f32 from the registers (each warp
          // instruction touching 32 half-used sectors) a flush kept the softmax warps for ~5.4 k cycles (SKP_BWD_TRACE).
          if (ch < n_mh) {
#pragma unroll
            for (int c = 0; c < 2; ++c) {
              uint32_t v[32];
              ptx::tmem_ld32(tm_dK + lane_off + ch * 64 + c * 32, v);
              ptx::tmem_ld_wait();
#if SKP_DZ_BF16
              // bf16 accumulation: one [128 tokens][64 taps] bf16 tile (128 B rows) per token half
              uint8_t* tile = ds_ptr + ch * 16384 + row * 128;
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                uint32_t w4[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                  const __nv_bfloat162 h2 = __floats2bfloat162_rn(__uint_as_float(v[8 * q + 2 * e]), __uint_as_float(v[8 * q + 2 * e + 1]));
                  w4[e] = *reinterpret_cast<const uint32_t*>(&h2);
                }
                *reinterpret_cast<uint4*>(tile + (((4 * c + q) ^ (row & 7)) << 4)) = make_uint4(w4[0], w4[1], w4[2], w4[3]);
              }
#else
              uint8_t* tile = ds_ptr + (ch * 2 + c) * 16384 + row * 128;
#pragma unroll
              for (int q = 0; q < 8; ++q)
                *reinterpret_cast<uint4*>(tile + ((q ^ (row & 7)) << 4)) = make_uint4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
#endif
            }
            ptx::tc_fence_before();
          }
          ptx::fence_proxy_async();
          named_bar_sync(1, TC_SOFTMAX_THREADS);
          if (stid == 0) {
            const int z = x.b * P.ops.heads[x.l] + x.head;
#if SKP_DZ_BF16
            for (int mh = 0; mh < n_mh; ++mh) ptx::tma_reduce_add_3d(&P.dz_map[x.l], dsb + mh * 16384, x.r0 * s, tok0 + mh * 128, z);
#else
            for (int t4 = 0; t4 < 2 * n_mh; ++t4)
              ptx::tma_reduce_add_3d(&P.dz_map[x.l], dsb + t4 * 16384, x.r0 * s + (t4 & 1) * 32, tok0 + (t4 >> 1) * 128, z);
#endif
            ptx::bulk_commit_group();
          }
          flush_pending = true;
      BW_TR(10)
        }
        if (WIDE && lr_flush) {
          const int t = tok0 + (WIDE ? 0 : ch * 128) + row;
          const int h = P.ops.heads[x.l];
          const int r0 = x.r0;
          const int nvalid = min(WIDE ? 128 : TC_BK, s * s - r0 * s) - (WIDE ? ch * 64 : 0);      // taps of the window inside the source image
          float* dz = P.dz[x.l] + ((static_cast<int64_t>(x.b) * h + x.head) * P.T + t) * (s * s) + r0 * s + (WIDE ? ch * 64 : 0);
#pragma unroll
          for (int c = 0; c < 2; ++c) {
            uint32_t v[32];
            ptx::tmem_ld32(tm_dK + lane_off + ch * 64 + c * 32, v);
            ptx::tmem_ld_wait();
            if (t < P.T && !(BW_DBG(P) & 1)) {
#pragma unroll
              for (int q = 0; q < 8; ++q)
                if (c * 32 + q * 4 < nvalid)
                  red_add_v4(dz + c * 32 + q * 4, __uint_as_float(v[4 * q]), __uint_as_float(v[4 * q + 1]),
                             __uint_as_float(v[4 * q + 2]), __uint_as_float(v[4 * q + 3]));
            }
          }
      BW_TR(10)
          ptx::tc_fence_before();
        }
      } else
      // ---- dK: flush the TMEM accumulator when this head (or image, for a per-image k) is finished
      if (bw_flush_between<false>(P, x, nxu, u + 1 >= u_end) && ch < n_mh) {
        const int t = tok0 + ch * 128 + row;
        float* dk = P.dk[x.l] + (static_cast<int64_t>(P.ops.k_shared[x.l] ? 0 : x.b) * P.T + t) * C + x.head * d;
        const float inv = 1.0f / kLog2eBw;
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          if (c * 32 < d) {
            uint32_t v[32];
            ptx::tmem_ld32(tm_dK + lane_off + ch * 64 + c * 32, v);
            ptx::tmem_ld_wait();
            if (t < P.T) {
#pragma unroll
              for (int q = 0; q < 8; ++q)
                if (c * 32 + q * 4 < d)
                  red_add_v4(dk + c * 32 + q * 4, __uint_as_float(v[4 * q]) * inv, __uint_as_float(v[4 * q + 1]) * inv,
                             __uint_as_float(v[4 * q + 2]) * inv, __uint_as_float(v[4 * q + 3]) * inv);
            }
          }
        }
        ptx::tc_fence_before();
      }
      // hand over the row prologue computed ahead (dropped when the next unit changes image or layer: its masks / weight differ)
      have_row = AHEAD && u + 1 < u_end && nxu.b == x.b && nxu.l == x.l;
      if (have_row) {
        m = m_n;
        nd = nd_n;
#pragma unroll
        for (int j = 0; j < BW_SEL / 2; ++j) pv[j] = pv_n[j];
      }
    }
    if (P.prof && blockIdx.x == ((BW_DBG(P) >> 6) & 1) && stid == 0) {
      for (int i = 0; i < 5; ++i) P.prof[2 + i] = pf[i];
      P.prof[7] = it;
      P.prof[10] = pf[5];
    }
  }

  // ---- teardown -------------------------------------------------------------------------------
  if (LR && !WIDE && threadIdx.x == BW_SOFTMAX_WARP0 * 32) ptx::bulk_wait_group0();      // the last window flush (issued by this thread)
  __syncwarp();
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, 512);
  }
#ifdef SKP_BWD_TRACE
  if (LR == BW_TR_LR && !WIDE && blockIdx.x == BW_TR_CTA && lane == 0 && (warp == 1 || warp >= BW_SOFTMAX_WARP0))
    for (int i = 0; i < BW_TR_N; ++i) {
      const long long* e = g_bw_tr[warp][i];
      const long long t0 = g_bw_tr[1][0][0];
      printf("TR w%02d u%02d %lld %lld %lld %lld %lld %lld %lld %lld %lld %lld %lld %lld\n", warp, i, e[0] - t0, e[1] - t0, e[2] - t0, e[3] - t0,
             e[4] - t0, e[5] - t0, e[6] - t0, e[7] - t0, e[8] - t0, e[9] - t0, e[10] - t0, e[11] - t0);
    }
#endif
}

// ==========================================================================================
// host side
// ==========================================================================================
// attn_tc_bwd_lr.cu / attn_tc_lr.cu
size_t lr_bwd_post_workspace(const skp_attn_layer& a, int B, int T);
int lr_bwd_post(const skp_attn_layer& a, int B, int T, const float* dz, char* ws, cudaStream_t st, bool dz_bf16);
int lr_launch_weights(int s, int R, const LrGeom& g, __nv_bfloat16* hi, __nv_bfloat16* lo, cudaStream_t st);

// Which main kernel a layer's gradient takes.  The "upsample after the logits" kernels do not see dim_head (their K
// dimension is the interpolation window; only the native-resolution GEMMs before and after see it):
//   BWD_LR64   windows of <= 64 taps (16^2 sources at R = 128): 256 tokens per CTA, the fast path
//   BWD_LR128  windows of <= 128 taps (32^2 sources at R = 128): 128 tokens per CTA (BwSmem WIDE), slower per exponential
//              than the staged kernel -- taken when the staged kernel cannot run (dim_head > 64: SD-1.5's 80 at 32^2) or
//              when SKP_BWD_LR=2 asks for it (tests, A/B)
//   BWD_STAGED staged q' with dQ'/dK products + bicubic adjoint: dim_head <= 64 (64-column TMEM accumulators)
//   BWD_NONE   none of them (e.g. 64^2 sources with dim_head > 64): the call falls back to the CUDA-core kernel
// SKP_BWD_LR=0 sends every layer down the staged path (A/B measurements); read per call: the tests switch it.
enum { BWD_NONE = 0, BWD_STAGED = 1, BWD_LR64 = 2, BWD_LR128 = 3 };
static int bwd_layer_kind(const skp_attn_layer& a, int R) {
  const char* e = getenv("SKP_BWD_LR");
  const int mode = e ? atoi(e) : 1;
  const int d = a.C / a.heads;
  const bool staged_ok = d <= 64 && d % 16 == 0;
  const bool lr_shape = (a.s * a.s) % 8 == 0 && a.s != R && R % TC_BM == 0;
  const int kw = lr_shape ? lr_geometry(a.s, R).kw : 0;
  if (mode != 0 && kw == TC_BK) return BWD_LR64;
  if (mode != 0 && kw == 2 * TC_BK && (!staged_ok || mode == 2)) return BWD_LR128;
  return staged_ok ? BWD_STAGED : BWD_NONE;
}
static bool bwd_layer_lr(const skp_attn_layer& a, int R) { return bwd_layer_kind(a, R) >= BWD_LR64; }

bool attn_map_bwd_tc_supported(const skp_attn_layer* h_layers, int n_layers, int B, int T, int R, int n_idx) {
  if (!attn_map_tc_supported(h_layers, n_layers, B, T, R, n_idx)) return false;
  if (n_idx < 1 || n_idx > BW_SEL || R > 256) return false;
  if (!((R <= TC_BM && TC_BM % R == 0) || (R % TC_BM == 0))) return false;
  for (int l = 0; l < n_layers; ++l)
    if (bwd_layer_kind(h_layers[l], R) == BWD_NONE) return false;
  return true;
}

size_t attn_map_bwd_tc_workspace(const skp_attn_layer* h_layers, int n_layers, int B, int T, int R, int precision) {
  size_t total = 8192 + tc_align_up(static_cast<size_t>(B) * R * R * BW_SEL * 4, 1024);
  const int nbuf = precision == SKP_PREC_TC_BF16 ? 1 : 2;
  for (int l = 0; l < n_layers; ++l) {
    const skp_attn_layer& a = h_layers[l];
    const bool lr = bwd_layer_lr(a, R);
    const size_t kd = lr ? static_cast<size_t>(a.s) * a.s : a.C / a.heads;     // selected-token rows: k rows, or (LR) logit rows
    total += nbuf * tc_align_up(static_cast<size_t>(B) * a.heads * BW_SEL * kd * 2, 1024);
    if (lr) total += attn_map_lr_workspace(&a, 1, B, T, R) + lr_bwd_post_workspace(a, B, T) + 2048;
    else total += attn_map_staged_workspace(&a, 1, B, T, R, precision);
  }
  return total + 1024;
}

template <int NPROD, int NH, bool LR, bool WIDE = false>
static int launch_bwd(const BwdParams& P, cudaStream_t st) {
  using SM = BwSmem<NPROD, WIDE>;
  auto kern = attn_map_bwd_tc_kernel<NPROD, NH, LR, WIDE>;
  SKP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SM::kDynamic));
  kern<<<(::skp::note_launch(), P.n_chunks * NH), BW_THREADS, SM::kDynamic, st>>>(P);
  SKP_LAUNCH_CHECK();
  return SKP_OK;
}

// one launch of the main kernel over a subset of the captured layers (`sub`, with their global head offsets)
static int bwd_launch_group(const skp_attn_layer* sub, const int* ghead, int n_sub, int heads_all, int kind, int B, int T, int R,
                            const int64_t* token_idx, int n_idx, int64_t idx_stride, const float* dmaps_sel, const float* dpT,
                            const float* row_stats, bool x3, char** ws_io, cudaStream_t st, const LrLayerBufs* ready) {
  BwdParams P;
  memset(&P, 0, sizeof(P));
  char* ws = nullptr;
  const bool lr = kind >= BWD_LR64, wide = kind == BWD_LR128;
  SKP_REQUIRE(!lr || x3, "skp_attn_map_bwd: the upsample-after-logits path is bf16x3 only");
  int rc = lr ? tc_prepare_lr(sub, n_sub, B, T, R, *ws_io, st, &P.ops, &ws, wide ? 128 : TC_BN, ready)
              : tc_prepare(sub, n_sub, B, T, R, x3, *ws_io, st, &P.ops, &ws);
  if (rc != SKP_OK) return rc;
  const int tiles = R * R / TC_BM;
  char* post_ws[TC_MAX_LAYERS];
  for (int l = 0; l < n_sub; ++l) {
    const skp_attn_layer& a = sub[l];
    SKP_REQUIRE(a.dq && a.dk, "skp_attn_map_bwd: layer has null dq/dk");
    const int d = a.C / a.heads;
    const int kd = lr ? a.s * a.s : d;       // inner length of the selected-token rows: dim_head, or (LR) the s*s logits
    const size_t bytes = tc_align_up(static_cast<size_t>(B) * a.heads * BW_SEL * kd * 2, 1024);
    __nv_bfloat16* hi = reinterpret_cast<__nv_bfloat16*>(ws); ws += bytes;
    __nv_bfloat16* lo = nullptr;
    if (x3) { lo = reinterpret_cast<__nv_bfloat16*>(ws); ws += bytes; }
    dim3 g(BW_SEL, B);
    if (lr) bw_gather_ssel_kernel<<<(::skp::note_launch(), dim3(BW_SEL, B * a.heads)), 128, 0, st>>>(P.ops.s_ptr[l][0], P.ops.s_ptr[l][1], T, kd, a.heads, token_idx, n_idx, idx_stride, hi, lo);
    else if (x3) bw_prep_ksel_kernel<3><<<(::skp::note_launch(), g), 256, 0, st>>>(a.k, T, a.C, a.heads, a.k_shared, token_idx, n_idx, idx_stride, hi, lo);
    else bw_prep_ksel_kernel<1><<<(::skp::note_launch(), g), 256, 0, st>>>(a.k, T, a.C, a.heads, a.k_shared, token_idx, n_idx, idx_stride, hi, lo);
    SKP_LAUNCH_CHECK();
    rc = tc_encode_map(&P.ksel_map[l][0], hi, kd, BW_SEL, B * a.heads, BW_SEL);
    if (rc == SKP_OK && x3) rc = tc_encode_map(&P.ksel_map[l][1], lo, kd, BW_SEL, B * a.heads, BW_SEL);
    if (rc != SKP_OK) return rc;
    P.dq[l] = a.dq; P.dk[l] = a.dk; P.s[l] = a.s; P.C[l] = a.C; P.ghead[l] = ghead[l];
    if (lr) {
      const LrGeom geo = lr_geometry(a.s, R);
      P.wr0[l] = reinterpret_cast<unsigned char*>(ws);
      bw_window_table_kernel<<<(::skp::note_launch(), 1), 256, 0, st>>>(a.s, R, geo.nr, tiles, reinterpret_cast<unsigned char*>(ws));
      SKP_LAUNCH_CHECK();
      ws += 1024;
      const size_t dzb = tc_align_up(static_cast<size_t>(B) * a.heads * T * a.s * a.s * 4, 1024);
      P.dz[l] = reinterpret_cast<float*>(ws);
      SKP_CUDA(cudaMemsetAsync(ws, 0, dzb, st));
      // (the wide-window kernel adds into fp32 dz with vector reductions and does not use the map)
      rc = (SKP_DZ_BF16 && kind == BWD_LR64) ? tc_encode_map(&P.dz_map[l], ws, a.s * a.s, T, B * a.heads, 128)
                                                 : tc_encode_map_f32(&P.dz_map[l], ws, a.s * a.s, T, B * a.heads, 128);
      if (rc != SKP_OK) return rc;
      ws += dzb;
      post_ws[l] = ws;
      ws += lr_bwd_post_workspace(a, B, T) - dzb;
    }
  }
  P.heads_all = heads_all;
  P.B = B; P.T = T; P.R = R; P.R2 = R * R; P.n_idx = n_idx; P.tiles = tiles;
  P.idx_stride = idx_stride; P.token_idx = token_idx; P.dmaps_sel = dmaps_sel; P.dpT = dpT; P.row_stats = row_stats;
  P.n_units = P.ops.total_heads * B * P.tiles;
#ifdef SKP_DIAG
  if (const char* e = getenv("SKP_BWD_DBG")) P.dbg = atoi(e);
  static unsigned long long* d_prof = nullptr;
  if (P.dbg & 8) {
    if (!d_prof) cudaMalloc(&d_prof, 128);
    cudaMemsetAsync(d_prof, 0, 128, st);
    P.prof = d_prof;
  }
#endif
  const int nh = wide ? (T > 384 ? 4 : T > 256 ? 3 : T > 128 ? 2 : 1) : (T > TC_BN ? 2 : 1);    // token parts per tile
  const int sms = num_sms();      // persistent grid: one CTA per SM of the CURRENT device
  P.n_chunks = sms / nh < P.n_units ? sms / nh : P.n_units;
  if (wide) {
    rc = nh == 4 ? launch_bwd<3, 4, true, true>(P, st) : nh == 3 ? launch_bwd<3, 3, true, true>(P, st)
       : nh == 2 ? launch_bwd<3, 2, true, true>(P, st) : launch_bwd<3, 1, true, true>(P, st);
  } else if (lr) {
    if (x3) rc = nh == 2 ? launch_bwd<3, 2, true>(P, st) : launch_bwd<3, 1, true>(P, st);
    else rc = nh == 2 ? launch_bwd<1, 2, true>(P, st) : launch_bwd<1, 1, true>(P, st);
  } else {
    if (x3) rc = nh == 2 ? launch_bwd<3, 2, false>(P, st) : launch_bwd<3, 1, false>(P, st);
    else rc = nh == 2 ? launch_bwd<1, 2, false>(P, st) : launch_bwd<1, 1, false>(P, st);
  }
  if (rc != SKP_OK) return rc;
#ifdef SKP_DIAG
  if (P.dbg & 8) {
    unsigned long long h[16];
    cudaStreamSynchronize(st);
    cudaMemcpy(h, d_prof, 128, cudaMemcpyDeviceToHost);
    const double n = h[7] ? static_cast<double>(h[7]) : 1.0;
    fprintf(stderr, "[bwd prof%s, cycles/unit over %llu units] mma: wait q_full %.0f, issue S %.0f, wait ds_full %.0f, issue dQ/dK %.0f | softmax: loop top %.0f, "
            "wait s_full %.0f, dS %.0f, fence+arrive %.0f, wait mma_done %.0f, drain+adjoint %.0f\n", lr ? " LR" : "", h[7], h[0] / n, h[8] / n, h[1] / n, h[9] / n,
            h[10] / n, h[2] / n, h[3] / n, h[4] / n, h[5] / n, h[6] / n);
  }
#endif
  if (lr)
    for (int l = 0; l < n_sub; ++l) {
      rc = lr_bwd_post(sub[l], B, T, P.dz[l], post_ws[l], st, SKP_DZ_BF16 && kind == BWD_LR64);
      if (rc != SKP_OK) return rc;
    }
  *ws_io = ws;
  return SKP_OK;
}

bool attn_map_fwd_stages_lr(const skp_attn_layer* h_layers, int n_layers, int B, int T, int R, int precision);

// fwd_workspace (optional): the workspace skp_attn_map_fwd staged its operands in for the SAME layers / q / k / B / T / R /
// precision.  Layers on the upsample-after-logits path then take their native-resolution logits and interpolation tables
// from it instead of recomputing them (two head splits, the logit GEMM and the table per layer).
int attn_map_bwd_tc(const skp_attn_layer* h_layers, int n_layers, int B, int T, int R, const int64_t* token_idx, int n_idx,
                    int64_t idx_stride, const float* dmaps_sel, const float* row_stats, int precision, void* workspace,
                    size_t workspace_bytes, cudaStream_t st, const void* fwd_workspace, size_t fwd_workspace_bytes) {
  SKP_REQUIRE(token_idx && dmaps_sel && row_stats, "skp_attn_map_bwd (tcgen05): token_idx, dmaps_sel and row_stats are required");
  const size_t need = attn_map_bwd_tc_workspace(h_layers, n_layers, B, T, R, precision);
  if (!workspace || workspace_bytes < need) {
    set_error("skp_attn_map_bwd: workspace %zu B < required %zu B", workspace_bytes, need);
    return SKP_ERR_WORKSPACE;
  }
  const bool x3 = precision != SKP_PREC_TC_BF16;
  // up to three launches: staged dQ' / dK + bicubic adjoint, upsample-after-logits with 64-tap and with 128-tap windows
  skp_attn_layer group[4][TC_MAX_LAYERS];
  int ghead[4][TC_MAX_LAYERS], n_group[4] = {0, 0, 0, 0}, heads_all = 0;
  LrLayerBufs fwd_bufs[TC_MAX_LAYERS], ready[4][TC_MAX_LAYERS];
  bool reuse = false;
  if (fwd_workspace && attn_map_fwd_stages_lr(h_layers, n_layers, B, T, R, precision)) {
    const char* end = lr_carve(h_layers, n_layers, B, T, R, const_cast<void*>(fwd_workspace), fwd_bufs);
    SKP_REQUIRE(static_cast<size_t>(end - static_cast<const char*>(fwd_workspace)) <= fwd_workspace_bytes,
                "skp_attn_map_bwd: the forward workspace handed back is smaller than the forward's operand layout");
    reuse = true;
  }
  for (int l = 0; l < n_layers; ++l) {
    const int gi = bwd_layer_kind(h_layers[l], R);
    SKP_REQUIRE(gi != BWD_NONE, "skp_attn_map_bwd: layer %d has no tensor-core backward", l);
    ready[gi][n_group[gi]] = fwd_bufs[l];
    group[gi][n_group[gi]] = h_layers[l];
    ghead[gi][n_group[gi]++] = heads_all;
    heads_all += h_layers[l].heads;
  }
  char* ws = reinterpret_cast<char*>(tc_align_up(reinterpret_cast<size_t>(workspace), 1024));
  float* dpT = reinterpret_cast<float*>(ws);
  ws += tc_align_up(static_cast<size_t>(B) * R * R * BW_SEL * 4, 1024);
  bw_transpose_dp_kernel<<<(::skp::note_launch(), dim3((R * R + 255) / 256, B)), 256, 0, st>>>(dmaps_sel, n_idx, R * R, dpT);
  SKP_LAUNCH_CHECK();
  for (int gi = BWD_STAGED; gi <= BWD_LR128; ++gi) {
    if (n_group[gi] == 0) continue;
    int rc = bwd_launch_group(group[gi], ghead[gi], n_group[gi], heads_all, gi, B, T, R, token_idx, n_idx, idx_stride, dmaps_sel,
                              dpT, row_stats, x3, &ws, st, reuse && gi >= BWD_LR64 ? ready[gi] : nullptr);
    if (rc != SKP_OK) return rc;
  }
  return SKP_OK;
}

}  // namespace skp
