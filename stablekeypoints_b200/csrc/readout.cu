// Bandwidth-bound readout / selection / loss kernels.
// One pass (or a few L2-resident passes) over the map planes; coalesced 128-bit loads where
// the plane size allows, warp-shuffle + shared-memory block reductions, fp64 accumulators for
// the long sums that decide integer rankings.
#include "common.cuh"

namespace skp {

// =========================================================================================
// per-plane argmax (eval.py:39-60).  grid = (chunks, N); packed 64-bit atomicMax keeps the
// largest value and, among equals, the smallest flat index (first-index tie rule).
// =========================================================================================
__global__ void plane_argmax_kernel(const float* __restrict__ maps, int64_t plane, unsigned long long* keys) {
  __shared__ unsigned long long scratch[32];
  const float* p = maps + static_cast<int64_t>(blockIdx.y) * plane;
  const int64_t per = (plane + gridDim.x - 1) / gridDim.x;
  const int64_t lo = blockIdx.x * per;
  const int64_t hi = lo + per < plane ? lo + per : plane;
  unsigned long long best = 0ull;
  if (((reinterpret_cast<uintptr_t>(p + lo) & 15) == 0)) {
    const int64_t n4 = (hi - lo) >> 2;
    const float4* p4 = reinterpret_cast<const float4*>(p + lo);
    for (int64_t i = threadIdx.x; i < n4; i += blockDim.x) {
      float4 v = __ldg(p4 + i);
      uint32_t base = static_cast<uint32_t>(lo + (i << 2));
      unsigned long long k0 = argmax_key(v.x, base), k1 = argmax_key(v.y, base + 1);
      unsigned long long k2 = argmax_key(v.z, base + 2), k3 = argmax_key(v.w, base + 3);
      k0 = k0 > k1 ? k0 : k1;
      k2 = k2 > k3 ? k2 : k3;
      k0 = k0 > k2 ? k0 : k2;
      best = best > k0 ? best : k0;
    }
    for (int64_t i = lo + (n4 << 2) + threadIdx.x; i < hi; i += blockDim.x) {
      unsigned long long k = argmax_key(p[i], static_cast<uint32_t>(i));
      best = best > k ? best : k;
    }
  } else {
    for (int64_t i = lo + threadIdx.x; i < hi; i += blockDim.x) {
      unsigned long long k = argmax_key(p[i], static_cast<uint32_t>(i));
      best = best > k ? best : k;
    }
  }
  best = block_max_u64(best, scratch);
  if (threadIdx.x == 0 && hi > lo) atomicMax(keys + blockIdx.y, best);
}

__global__ void argmax_decode_kernel(unsigned long long* keys, int N, float* maxval) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  unsigned long long k = keys[i];
  if (maxval) maxval[i] = argmax_key_value(k);
  reinterpret_cast<int64_t*>(keys)[i] = static_cast<int64_t>(argmax_key_index(k));
}

// masked argmax used by find_k_max_pixels / plane_stats: pixels inside any previous disc read
// as 0 (the reference multiplies the map by the mask, eval.py:108-109).
struct Discs {
  float r[8], c[8];
  int n;
  float rad2;
};
__device__ __forceinline__ float disc_masked(float v, int row, int col, const Discs& d) {
  for (int j = 0; j < d.n; ++j) {
    float dc = static_cast<float>(col) - d.c[j];
    float dr = static_cast<float>(row) - d.r[j];
    float d2 = __fadd_rn(__fmul_rn(dc, dc), __fmul_rn(dr, dr));
    if (!(d2 > d.rad2)) v = v * 0.0f;
  }
  return v;
}
__device__ unsigned long long block_masked_argmax(const float* __restrict__ p, int H, int W, const Discs& d,
                                                  unsigned long long* scratch) {
  unsigned long long best = 0ull;
  const int n = H * W;
  if (d.n == 0 && (n & 3) == 0 && (reinterpret_cast<uintptr_t>(p) & 15) == 0) {
    // no disc yet (the first, usually only, subject): 16-byte loads, no per-pixel row/column arithmetic
    const float4* p4 = reinterpret_cast<const float4*>(p);
    for (int i = threadIdx.x; i < (n >> 2); i += blockDim.x) {
      const float4 v = __ldg(p4 + i);
      const uint32_t b = static_cast<uint32_t>(i) << 2;
      unsigned long long k0 = argmax_key(v.x, b), k1 = argmax_key(v.y, b + 1);
      unsigned long long k2 = argmax_key(v.z, b + 2), k3 = argmax_key(v.w, b + 3);
      k0 = k0 > k1 ? k0 : k1;
      k2 = k2 > k3 ? k2 : k3;
      k0 = k0 > k2 ? k0 : k2;
      best = best > k0 ? best : k0;
    }
    return block_max_u64(best, scratch);
  }
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    float v = p[i];
    if (d.n) v = disc_masked(v, i / W, i % W, d);
    unsigned long long k = argmax_key(v, static_cast<uint32_t>(i));
    best = best > k ? best : k;
  }
  return block_max_u64(best, scratch);
}

__global__ void k_max_pixels_kernel(const float* __restrict__ maps, int N, int H, int W, int num, float* out) {
  __shared__ unsigned long long scratch[32];
  const float* p = maps + static_cast<int64_t>(blockIdx.x) * H * W;
  Discs d;
  d.n = 0;
  // python: rad = 0.05*h in double, squared and compared in fp32
  d.rad2 = static_cast<float>((0.05 * static_cast<double>(H)) * (0.05 * static_cast<double>(H)));
  for (int s = 0; s < num; ++s) {
    unsigned long long key = block_masked_argmax(p, H, W, d, scratch);
    uint32_t idx = argmax_key_index(key);
    float r = static_cast<float>(idx / W) + 0.5f, c = static_cast<float>(idx % W) + 0.5f;
    if (threadIdx.x == 0) {
      out[(static_cast<int64_t>(s) * N + blockIdx.x) * 2 + 0] = r;
      out[(static_cast<int64_t>(s) * N + blockIdx.x) * 2 + 1] = c;
    }
    d.r[d.n] = r;
    d.c[d.n] = c;
    d.n++;
    __syncthreads();
  }
}

// =========================================================================================
// windowed soft-argmax (eval.py:113-155)
// =========================================================================================
__global__ void weighted_avg_kernel(float* maps, int H, int W, float distance, int mutate, float* out) {
  __shared__ unsigned long long scratch[32];
  __shared__ double dscratch[32];
  float* p = maps + static_cast<int64_t>(blockIdx.x) * H * W;
  Discs none;
  none.n = 0;
  int r0 = 0, c0 = 0;
  const bool window = distance >= 0.0f;
  if (window) {
    unsigned long long key = block_masked_argmax(p, H, W, none, scratch);
    uint32_t idx = argmax_key_index(key);
    r0 = idx / W;
    c0 = idx % W;
  }
  double s0 = 0.0, sr = 0.0, sc = 0.0;
  int rlo = 0, rhi = H, clo = 0, chi = W;
  if (window && !mutate) {   // only the bounding box of the disc can contribute
    int rad = static_cast<int>(ceilf(distance)) + 1;
    rlo = max(0, r0 - rad); rhi = min(H, r0 + rad + 1);
    clo = max(0, c0 - rad); chi = min(W, c0 + rad + 1);
  }
  const int bw = chi - clo, bn = (rhi - rlo) * bw;
  for (int i = threadIdx.x; i < bn; i += blockDim.x) {
    int r = rlo + i / bw, c = clo + i % bw;
    float v = p[r * W + c];
    if (window) {
      float dr = static_cast<float>(r) - static_cast<float>(r0), dc = static_cast<float>(c) - static_cast<float>(c0);
      float dist = __fsqrt_rn(__fadd_rn(__fmul_rn(dr, dr), __fmul_rn(dc, dc)));
      if (dist > distance) {
        v = 0.0f;
        if (mutate) p[r * W + c] = 0.0f;
      }
    }
    s0 += v;
    sr += static_cast<double>(r) * v;
    sc += static_cast<double>(c) * v;
  }
  s0 = block_sum(s0, dscratch);
  sr = block_sum(sr, dscratch);
  sc = block_sum(sc, dscratch);
  if (threadIdx.x == 0) {
    double den = static_cast<double>(static_cast<float>(s0) + 1e-6f);
    out[blockIdx.x * 2 + 0] = static_cast<float>(sr / den) + 0.5f;
    out[blockIdx.x * 2 + 1] = static_cast<float>(sc / den) + 0.5f;
  }
}

// =========================================================================================
// per-plane Gaussian-KL and pixel-entropy (ptp_utils.py:86-112, 165-187)
// =========================================================================================
__global__ void plane_stats_kernel(const float* __restrict__ maps, int H, int W, float sigma, float eps,
                                   int num_subjects, float* kl_out, float* ent_out, int64_t* argmax_out) {
  __shared__ unsigned long long scratch[32];
  __shared__ double dscratch[32];
  const int n = H * W;
  const float* p = maps + static_cast<int64_t>(blockIdx.x) * n;
  // subjects: successive masked maxima (find_k_max_pixels); centres in pixel units as the
  // reference forms them: ((k+0.5)/H)*size with size == H, in fp32.
  Discs d;
  d.n = 0;
  d.rad2 = static_cast<float>((0.05 * static_cast<double>(H)) * (0.05 * static_cast<double>(H)));
  float cr[8], cc[8];
  uint32_t first_idx = 0;
  for (int s = 0; s < num_subjects; ++s) {
    unsigned long long key = block_masked_argmax(p, H, W, d, scratch);
    uint32_t idx = argmax_key_index(key);
    if (s == 0) first_idx = idx;
    if (s == 0 && argmax_out && threadIdx.x == 0) argmax_out[blockIdx.x] = idx;
    float r = static_cast<float>(idx / W) + 0.5f, c = static_cast<float>(idx % W) + 0.5f;
    d.r[d.n] = r;
    d.c[d.n] = c;
    d.n++;
    cr[s] = __fmul_rn(__fdiv_rn(r, static_cast<float>(H)), static_cast<float>(H));
    cc[s] = __fmul_rn(__fdiv_rn(c, static_cast<float>(H)), static_cast<float>(H));
    __syncthreads();
  }
  // Plane max for the pixel softmax = the value at the first (unmasked) argmax.  Everything else is ONE more pass
  // over the plane (L1/L2-resident after the argmax pass), fp32 per element, fp64 accumulators:
  //   s0 = sum e^(x-mx), s1 = sum e^(x-mx) (x-mx)                      -> entropy = log s0 - s1/s0
  //   a = gaussian/S + eps, z = sum a, A = sum a log a, Bv = sum a (x-mx)
  //   KL(target || softmax) = sum (a/z) (log(a/z) - (x - mx - log s0)) = (A - Bv)/z - log z + log s0
  // (the per-pixel double-precision log and division of the textbook form were 17x slower than the loads).
  const float mx = p[first_idx];
  const float inv2s2 = 1.0f / (2.0f * sigma * sigma);
  const float invS = 1.0f / static_cast<float>(num_subjects);
  const bool want_kl = kl_out != nullptr;
  const bool wp2 = (W & (W - 1)) == 0;
  const int lgw = __ffs(W) - 1;
  double s0 = 0.0, s1 = 0.0, z = 0.0, A = 0.0, Bv = 0.0;
  // per-pixel terms in fp32; the four pixels of a 16-byte load are summed in fp32 (a 3-add tree) and enter the fp64
  // accumulators once per vector: the float->double conversions (XU pipe, 16/clk/SM) were the bottleneck
  struct Terms { float e, ev, a, alog, av; };
  auto pixel = [&](float x, float rr, float cx) -> Terms {      // rr, cx: pixel centre (row + 0.5, col + 0.5)
    Terms t;
    const float v = x - mx;
    t.e = expf(v);
    t.ev = t.e * v;
    t.a = t.alog = t.av = 0.0f;
    if (want_kl) {
      float g = 0.0f;
      for (int s = 0; s < num_subjects; ++s) {
        const float dc = cx - cc[s], dr = rr - cr[s];
        g += expf(-(dc * dc + dr * dr) * inv2s2);
      }
      t.a = g * invS + eps;
      t.alog = t.a * logf(t.a);
      t.av = t.a * v;
    }
    return t;
  };
  if ((n & 3) == 0 && (W & 3) == 0 && (reinterpret_cast<uintptr_t>(p) & 15) == 0) {
    const float4* p4 = reinterpret_cast<const float4*>(p);
    for (int i4 = threadIdx.x; i4 < (n >> 2); i4 += blockDim.x) {
      const float4 x = __ldg(p4 + i4);
      const int i = i4 << 2;
      const int row = wp2 ? i >> lgw : i / W, col = wp2 ? i & (W - 1) : i - row * W;   // W % 4 == 0: one row per vector
      const float rr = static_cast<float>(row) + 0.5f, cx = static_cast<float>(col) + 0.5f;   // integers + 0.5: exact in fp32
      const Terms t0 = pixel(x.x, rr, cx), t1 = pixel(x.y, rr, cx + 1.0f), t2 = pixel(x.z, rr, cx + 2.0f), t3 = pixel(x.w, rr, cx + 3.0f);
      s0 += static_cast<double>((t0.e + t1.e) + (t2.e + t3.e));
      s1 += static_cast<double>((t0.ev + t1.ev) + (t2.ev + t3.ev));
      if (want_kl) {
        z += static_cast<double>((t0.a + t1.a) + (t2.a + t3.a));
        A += static_cast<double>((t0.alog + t1.alog) + (t2.alog + t3.alog));
        Bv += static_cast<double>((t0.av + t1.av) + (t2.av + t3.av));
      }
    }
  } else {
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      const Terms t = pixel(p[i], static_cast<float>(i / W) + 0.5f, static_cast<float>(i % W) + 0.5f);
      s0 += t.e; s1 += t.ev; z += t.a; A += t.alog; Bv += t.av;
    }
  }
  s0 = block_sum(s0, dscratch);
  s1 = block_sum(s1, dscratch);
  if (ent_out && threadIdx.x == 0) ent_out[blockIdx.x] = static_cast<float>(log(s0) - s1 / s0);
  if (!want_kl) return;
  z = block_sum(z, dscratch);
  A = block_sum(A, dscratch);
  Bv = block_sum(Bv, dscratch);
  if (threadIdx.x == 0) kl_out[blockIdx.x] = static_cast<float>((A - Bv) / z - log(z) + log(s0));
}

// Register-resident single-subject version for planes of exactly 512 * 4 * VPT pixels (64^2, 128^2, ...): every thread
// loads its VPT 16-byte vectors ONCE (all loads in flight before the first use: the plane crosses HBM once), takes the
// argmax and the softmax sums from registers (fp32 inside a thread, fp64 across threads), and all sums share one block
// reduction.  The Gaussian target only needs evaluating near the argmax: once exp(-d^2 / 2 sigma^2) < eps * 2^-25 the
// fp32 sum g + eps rounds to eps exactly, so everything outside a small window contributes closed-form terms.
constexpr int PS_THREADS = 512;
__device__ __forceinline__ float ps_ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
template <int VPT, bool KL>
__global__ void __launch_bounds__(PS_THREADS, 2)
plane_stats_reg_kernel(const float* __restrict__ maps, int W, float sigma, float eps, float* kl_out, float* ent_out,
                       int64_t* argmax_out) {
  __shared__ unsigned long long scratch[32];
  __shared__ double dsum[6][PS_THREADS / 32];
  constexpr int n = PS_THREADS * 4 * VPT;
  const float4* p4 = reinterpret_cast<const float4*>(maps + static_cast<int64_t>(blockIdx.x) * n);
  float4 x[VPT];
#pragma unroll
  for (int k = 0; k < VPT; ++k) x[k] = __ldg(p4 + threadIdx.x + k * PS_THREADS);
  // per-thread argmax on plain float compares (elements are visited in increasing flat index, so a strict '>' keeps
  // the first maximum; a NaN beats everything that is not already a NaN), one 64-bit key per thread for the block
  float bv = x[0].x;
  uint32_t bi = static_cast<uint32_t>(threadIdx.x) << 2;
#pragma unroll
  for (int k = 0; k < VPT; ++k) {
    const uint32_t b = static_cast<uint32_t>(threadIdx.x + k * PS_THREADS) << 2;
    const float xv[4] = {x[k].x, x[k].y, x[k].z, x[k].w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const bool upd = (xv[j] > bv) || (xv[j] != xv[j] && bv == bv);
      bv = upd ? xv[j] : bv;
      bi = upd ? b + j : bi;
    }
  }
  unsigned long long best = argmax_key(bv, bi);
  best = block_max_u64(best, scratch);
  const uint32_t idx = argmax_key_index(best);
  if (argmax_out && threadIdx.x == 0) argmax_out[blockIdx.x] = idx;
  if (!KL && !ent_out) return;
  const float mx = argmax_key_value(best);
  const int lgw = __ffs(W) - 1;             // W is a power of two on this path
  const int H = n >> lgw;
  // centre in pixel units exactly as the reference forms it: ((k + 0.5) / H) * size with size == H
  const float cr = __fmul_rn(__fdiv_rn(static_cast<float>(idx >> lgw) + 0.5f, static_cast<float>(H)), static_cast<float>(H));
  const float cc = __fmul_rn(__fdiv_rn(static_cast<float>(idx & (W - 1)) + 0.5f, static_cast<float>(H)), static_cast<float>(H));
  const float inv2s2 = 1.0f / (2.0f * sigma * sigma);
  // all pixels: s0 = sum e^v, s1 = sum e^v v, sv = sum v (v = x - max).  fp32 inside a thread (32 * VPT/8 terms), fp64 across
  float s0f = 0.0f, s1f = 0.0f, svf = 0.0f;
#pragma unroll
  for (int k = 0; k < VPT; ++k) {
    const float xv[4] = {x[k].x, x[k].y, x[k].z, x[k].w};
    float v[4], e[4], ev[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      v[j] = xv[j] - mx;
      // exp(v) as ex2 of a two-term product v * log2(e) (hi + lo): the argument carries ~2^-30 relative error,
      // the result the 2 ulp of ex2.approx -- the same class as expf at a third of the instructions
      e[j] = ps_ex2(fmaf(v[j], 1.925963033500011e-8f, v[j] * 1.4426950216293335f));
      ev[j] = e[j] * v[j];
    }
    s0f += (e[0] + e[1]) + (e[2] + e[3]);
    if (ent_out) s1f += (ev[0] + ev[1]) + (ev[2] + ev[3]);     // sum e^v v feeds the entropy only (uniform branch)
    if (KL) svf += (v[0] + v[1]) + (v[2] + v[3]);
  }
  double s0 = static_cast<double>(s0f), s1 = static_cast<double>(s1f), sv = static_cast<double>(svf);
  // KL: with a = gaussian + eps,  z = sum a,  A = sum a log a,  Bv = sum a v.  Outside a square of half-width `rad`
  // around the argmax a == eps exactly, so those pixels contribute eps * (count), eps log eps * (count) and eps * sum v;
  // the <= (2 rad + 3)^2 window pixels are re-read (L1-resident: this CTA just loaded them) and get the full expf / logf.
  double z = 0.0, A = 0.0, Bv = 0.0;
  if (KL) {
    const float far_thr = eps > 0.0f ? -logf(eps) + 17.6f : 1e30f;     // exp(-far_thr) < eps * 2^-25
    const int rad = static_cast<int>(fminf(ceilf(sqrtf(far_thr / inv2s2)), 65536.0f)) + 1;   // +1: the centre is (k + 0.5) up to rounding
    const int r_c = static_cast<int>(idx >> lgw), c_c = static_cast<int>(idx & (W - 1));
    const int rlo = max(0, r_c - rad), rhi = min(H - 1, r_c + rad), clo = max(0, c_c - rad), chi = min(W - 1, c_c + rad);
    const int bw = chi - clo + 1, bn = (rhi - rlo + 1) * bw;
    const float* pl = maps + static_cast<int64_t>(blockIdx.x) * n;
    const float inv_bw = 1.0f / static_cast<float>(bw);
    for (int i = threadIdx.x; i < bn; i += PS_THREADS) {
      // i / bw without the ~40-instruction integer division: exact for i < 2^20, bw < 2^10 (error of the product << 0.5 / bw)
      const int q = __float2int_rd((static_cast<float>(i) + 0.5f) * inv_bw);
      const int r = rlo + q, c = clo + (i - q * bw);
      const float v = __ldg(pl + (r << lgw) + c) - mx;
      const float dr = (static_cast<float>(r) + 0.5f) - cr, dc = (static_cast<float>(c) + 0.5f) - cc;
      const float a = expf(-(dc * dc + dr * dr) * inv2s2) + eps;       // g * (1 / num_subjects) with one subject
      z += static_cast<double>(a);
      A += static_cast<double>(a * logf(a));
      Bv += static_cast<double>(a * v) - static_cast<double>(eps) * static_cast<double>(v);
    }
    if (threadIdx.x == 0) {
      const double nfar = static_cast<double>(n - bn);
      z += nfar * static_cast<double>(eps);
      A += nfar * static_cast<double>(eps * logf(eps));
    }
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  s0 = warp_sum(s0);
  s1 = warp_sum(s1);
  if (KL) {
    sv = warp_sum(sv);
    z = warp_sum(z);
    A = warp_sum(A);
    Bv = warp_sum(Bv);
  }
  if (lane == 0) {
    dsum[0][warp] = s0; dsum[1][warp] = s1;
    if (KL) { dsum[2][warp] = z; dsum[3][warp] = A; dsum[4][warp] = Bv; dsum[5][warp] = sv; }
  }
  __syncthreads();
  if (warp == 0) {
    double r[6];
#pragma unroll
    for (int q = 0; q < (KL ? 6 : 2); ++q) r[q] = warp_sum(lane < PS_THREADS / 32 ? dsum[q][lane] : 0.0);
    if (lane == 0) {
      if (ent_out) ent_out[blockIdx.x] = static_cast<float>(log(r[0]) - r[1] / r[0]);
      if (KL) {
        const double bv = r[4] + static_cast<double>(eps) * r[5];       // sum a v over the whole plane
        kl_out[blockIdx.x] = static_cast<float>((r[3] - bv) / r[2] - log(r[2]) + log(r[0]));
      }
    }
  }
}

template <int VPT>
static void launch_plane_stats_reg(const float* maps, int N, int W, float sigma, float eps, float* kl, float* ent,
                                   int64_t* am, cudaStream_t st) {
  if (kl) plane_stats_reg_kernel<VPT, true><<<(::skp::note_launch(), N), PS_THREADS, 0, st>>>(maps, W, sigma, eps, kl, ent, am);
  else plane_stats_reg_kernel<VPT, false><<<(::skp::note_launch(), N), PS_THREADS, 0, st>>>(maps, W, sigma, eps, kl, ent, am);
}

__global__ void gaussian_circles_kernel(const float* __restrict__ pos, int S, int N, int size, float sigma, float* out) {
  const int n = blockIdx.y;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= size * size) return;
  const float rr = static_cast<float>(i / size) + 0.5f, cx = static_cast<float>(i % size) + 0.5f;
  const float den = 2.0f * sigma * sigma;
  float g = 0.0f;
  for (int s = 0; s < S; ++s) {
    float pr = __fmul_rn(pos[(static_cast<int64_t>(s) * N + n) * 2 + 0], static_cast<float>(size));
    float pc = __fmul_rn(pos[(static_cast<int64_t>(s) * N + n) * 2 + 1], static_cast<float>(size));
    float dc = cx - pc, dr = rr - pr;
    float d2 = __fadd_rn(__fmul_rn(dc, dc), __fmul_rn(dr, dr));
    g += expf(__fdiv_rn(-1.0f * d2, den));
  }
  out[static_cast<int64_t>(n) * size * size + i] = S == 1 ? g : g / static_cast<float>(S);
}

// =========================================================================================
// ascending rank (torch.argsort(...)[:top_k]) — single-CTA bitonic sort of (score, index) keys
// =========================================================================================
__global__ void rank_ascending_kernel(const float* __restrict__ score, int N, int top_k, int64_t* out) {
  extern __shared__ unsigned long long keys[];
  score += static_cast<int64_t>(blockIdx.x) * N;      // one CTA per score vector (batched launches)
  out += static_cast<int64_t>(blockIdx.x) * top_k;
  int P = 1;
  while (P < N) P <<= 1;
  for (int i = threadIdx.x; i < P; i += blockDim.x) {
    unsigned long long k = ~0ull;
    if (i < N) {
      float v = score[i];
      if (v == 0.0f) v = 0.0f;   // -0 == +0
      uint32_t b = __float_as_uint(v);
      uint32_t ord = (v != v) ? 0xffffffffu : ((b & 0x80000000u) ? ~b : (b | 0x80000000u));
      k = (static_cast<unsigned long long>(ord) << 32) | static_cast<uint32_t>(i);
    }
    keys[i] = k;
  }
  __syncthreads();
  for (int k = 2; k <= P; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = threadIdx.x; i < P; i += blockDim.x) {
        int ixj = i ^ j;
        if (ixj > i) {
          unsigned long long a = keys[i], b = keys[ixj];
          bool up = (i & k) == 0;
          if ((a > b) == up) { keys[i] = b; keys[ixj] = a; }
        }
      }
      __syncthreads();
    }
  }
  for (int i = threadIdx.x; i < top_k; i += blockDim.x) out[i] = static_cast<int64_t>(keys[i] & 0xffffffffull);
}

// =========================================================================================
// furthest point sampling (ptp_utils.py:115-159) — fp32 arithmetic identical to the tensor ops
// =========================================================================================
__device__ __forceinline__ float fps_dist(float ar, float ac, float br, float bc) {
  float dr = __fsub_rn(ar, br), dc = __fsub_rn(ac, bc);
  return __fsqrt_rn(__fadd_rn(__fmul_rn(dr, dr), __fmul_rn(dc, dc)));
}

__global__ void fps_kernel(const int64_t* __restrict__ argmax_flat, int N, int H, int W,
                           const int64_t* __restrict__ cand, int n_cand, int top_k, int64_t* out, int32_t* n_out) {
  extern __shared__ float sm[];
  float* lr = sm;                 // [n_cand] row location
  float* lc = sm + n_cand;        // [n_cand] col location
  int* tok = reinterpret_cast<int*>(sm + 2 * n_cand);      // [n_cand] token id
  int* chosen = tok + n_cand;                               // [top_k] candidate positions
  __shared__ unsigned long long scratch[32];
  __shared__ int n_sel;
  argmax_flat += static_cast<int64_t>(blockIdx.x) * N;   // one CTA per image (batched launches)
  cand += static_cast<int64_t>(blockIdx.x) * n_cand;
  out += static_cast<int64_t>(blockIdx.x) * top_k;
  n_out += blockIdx.x;
  for (int i = threadIdx.x; i < n_cand; i += blockDim.x) {
    int t = static_cast<int>(cand[i]);
    tok[i] = t;
    int64_t f = argmax_flat[t];
    lr[i] = __fdiv_rn(static_cast<float>(f / W) + 0.5f, static_cast<float>(H));
    lc[i] = __fdiv_rn(static_cast<float>(f % W) + 0.5f, static_cast<float>(H));
  }
  __syncthreads();
  // furthest pair: strict '>' over pairs in (i<j) lexicographic order == max distance, first pair on ties
  unsigned long long best = 0ull;
  const int npair = n_cand * n_cand;
  for (int q = threadIdx.x; q < npair; q += blockDim.x) {
    int i = q / n_cand, j = q % n_cand;
    if (j <= i) continue;
    float dd = fps_dist(lr[i], lc[i], lr[j], lc[j]);
    unsigned long long k = (static_cast<unsigned long long>(__float_as_uint(dd)) << 32) |
                           static_cast<unsigned long long>(0xffffffffu - static_cast<uint32_t>(q));
    best = best > k ? best : k;
  }
  best = block_max_u64(best, scratch);
  if (threadIdx.x == 0) {
    uint32_t q = 0xffffffffu - static_cast<uint32_t>(best & 0xffffffffull);
    chosen[0] = q / n_cand;
    chosen[1] = q % n_cand;
    n_sel = 2;
  }
  __syncthreads();
  for (int round = 0; round < top_k - 2; ++round) {
    const int ns = n_sel;
    unsigned long long bk = 0ull;
    for (int i = threadIdx.x; i < n_cand; i += blockDim.x) {
      bool skip = false;
      for (int s = 0; s < ns; ++s) skip |= (tok[chosen[s]] == tok[i]);
      if (skip) continue;
      float m = INFINITY;
      for (int s = 0; s < ns; ++s) m = fminf(m, fps_dist(lr[i], lc[i], lr[chosen[s]], lc[chosen[s]]));
      // distances are >= 0 so the raw bit pattern orders them; +1 keeps 0 above the "none" key
      unsigned long long k = ((static_cast<unsigned long long>(__float_as_uint(m)) + 1ull) << 32) |
                             static_cast<unsigned long long>(0xffffffffu - static_cast<uint32_t>(i));
      bk = bk > k ? bk : k;
    }
    bk = block_max_u64(bk, scratch);
    if (threadIdx.x == 0 && bk != 0ull) {
      chosen[n_sel] = static_cast<int>(0xffffffffu - static_cast<uint32_t>(bk & 0xffffffffull));
      n_sel = n_sel + 1;
    }
    __syncthreads();
  }
  for (int i = threadIdx.x; i < n_sel; i += blockDim.x) out[i] = tok[chosen[i]];
  if (threadIdx.x == 0) *n_out = n_sel;
}

// =========================================================================================
// sharpening + equivariance losses (optimize.py:157-206) on the K selected planes
// =========================================================================================
struct InvGrid {   // un-warp sampling position for output pixel (y,x): affine_grid + grid_sample
  float ix, iy;
};
__device__ __forceinline__ InvGrid inv_grid(const float* th, int y, int x, int H, int W) {
  float xs = (2.0f * x + 1.0f) / W - 1.0f;
  float ys = (2.0f * y + 1.0f) / H - 1.0f;
  float gx = th[0] * xs + th[1] * ys + th[2];
  float gy = th[3] * xs + th[4] * ys + th[5];
  InvGrid g;
  g.ix = ((gx + 1.0f) * W - 1.0f) * 0.5f;
  g.iy = ((gy + 1.0f) * H - 1.0f) * 0.5f;
  return g;
}
__device__ __forceinline__ float sample_bilinear_zeros(const float* __restrict__ p, int H, int W, float ix, float iy) {
  float x0f = floorf(ix), y0f = floorf(iy);
  int x0 = static_cast<int>(x0f), y0 = static_cast<int>(y0f);
  float fx = ix - x0f, fy = iy - y0f;
  float v = 0.0f;
  if (y0 >= 0 && y0 < H) {
    if (x0 >= 0 && x0 < W) v += p[y0 * W + x0] * (1.0f - fx) * (1.0f - fy);
    if (x0 + 1 >= 0 && x0 + 1 < W) v += p[y0 * W + x0 + 1] * fx * (1.0f - fy);
  }
  if (y0 + 1 >= 0 && y0 + 1 < H) {
    if (x0 >= 0 && x0 < W) v += p[(y0 + 1) * W + x0] * (1.0f - fx) * fy;
    if (x0 + 1 >= 0 && x0 + 1 < W) v += p[(y0 + 1) * W + x0 + 1] * fx * fy;
  }
  return v;
}

// grid (K, B): plane k of image b.  Batched launches (B > 1): maps_a / maps_b (B, planes, H, W), idx (B, K) or one shared
// row (idx_stride 0), theta_inv (B, 2, 3), outputs (B, K, ...).  g_dev (backward, optional): (B, 2) DEVICE upstream
// gradients (d total / d mean sharpening, d total / d mean equivariance) of each image, multiplied into ca / cb here so
// that the autograd glue never reads a scalar back to the host.
template <bool kBackward>
__global__ void keypoint_losses_kernel(const float* __restrict__ maps_a, const float* __restrict__ maps_b,
                                       const int64_t* __restrict__ idx, int64_t idx_stride, int planes, int H, int W,
                                       const float* __restrict__ theta_inv, float sigma, int num_subjects, float* loss_parts,
                                       float ca, float cb, const float* __restrict__ g_dev, float* d_a_sel, float* d_b_sel) {
  __shared__ unsigned long long scratch[32];
  __shared__ double dscratch[32];
  const int n = H * W;
  const int img = blockIdx.y, K = gridDim.x;
  const int64_t t = idx[img * idx_stride + blockIdx.x] + static_cast<int64_t>(img) * planes;
  const float* a = maps_a + t * n;
  const float* b = maps_b ? maps_b + t * n : nullptr;
  theta_inv = theta_inv ? theta_inv + 6 * img : nullptr;
  const int64_t oplane = static_cast<int64_t>(img) * K + blockIdx.x;
  if (kBackward && g_dev) {
    ca *= g_dev[2 * img];
    cb *= g_dev[2 * img + 1];
  }
  Discs d;
  d.n = 0;
  d.rad2 = static_cast<float>((0.05 * static_cast<double>(H)) * (0.05 * static_cast<double>(H)));
  float cr[8], cc[8];
  for (int s = 0; s < num_subjects; ++s) {
    unsigned long long key = block_masked_argmax(a, H, W, d, scratch);
    uint32_t i = argmax_key_index(key);
    float r = static_cast<float>(i / W) + 0.5f, c = static_cast<float>(i % W) + 0.5f;
    d.r[d.n] = r; d.c[d.n] = c; d.n++;
    // sharpening_loss divides by attn_map.shape[-1] (= W) and gaussian_circles multiplies by size = H
    cr[s] = __fmul_rn(__fdiv_rn(r, static_cast<float>(W)), static_cast<float>(H));
    cc[s] = __fmul_rn(__fdiv_rn(c, static_cast<float>(W)), static_cast<float>(H));
    __syncthreads();
  }
  float th[6];
  if (b) {
#pragma unroll
    for (int j = 0; j < 6; ++j) th[j] = theta_inv[j];
  }
  const float inv2s2 = 1.0f / (2.0f * sigma * sigma);
  const float invS = 1.0f / static_cast<float>(num_subjects);
  double ss = 0.0, se = 0.0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const int y = i / W, x = i % W;
    const float av = a[i];
    float g = 0.0f;
    for (int s = 0; s < num_subjects; ++s) {
      float dc = (x + 0.5f) - cc[s], dr = (y + 0.5f) - cr[s];
      g += expf(-(dc * dc + dr * dr) * inv2s2);
    }
    g *= invS;
    const float ds = av - g;
    float de = 0.0f;
    InvGrid ig;
    if (b) {
      ig = inv_grid(th, y, x, H, W);
      de = av - sample_bilinear_zeros(b, H, W, ig.ix, ig.iy);
    }
    if (!kBackward) {
      ss += static_cast<double>(ds) * ds;
      se += static_cast<double>(de) * de;
    } else {
      d_a_sel[oplane * n + i] = ca * ds + cb * de;
      if (b && d_b_sel) {
        float* db = d_b_sel + oplane * n;
        const float gco = -cb * de;
        float x0f = floorf(ig.ix), y0f = floorf(ig.iy);
        int x0 = static_cast<int>(x0f), y0 = static_cast<int>(y0f);
        float fx = ig.ix - x0f, fy = ig.iy - y0f;
        if (y0 >= 0 && y0 < H) {
          if (x0 >= 0 && x0 < W) atomicAdd(db + y0 * W + x0, gco * (1.0f - fx) * (1.0f - fy));
          if (x0 + 1 >= 0 && x0 + 1 < W) atomicAdd(db + y0 * W + x0 + 1, gco * fx * (1.0f - fy));
        }
        if (y0 + 1 >= 0 && y0 + 1 < H) {
          if (x0 >= 0 && x0 < W) atomicAdd(db + (y0 + 1) * W + x0, gco * (1.0f - fx) * fy);
          if (x0 + 1 >= 0 && x0 + 1 < W) atomicAdd(db + (y0 + 1) * W + x0 + 1, gco * fx * fy);
        }
      }
    }
  }
  if (!kBackward) {
    ss = block_sum(ss, dscratch);
    se = block_sum(se, dscratch);
    if (threadIdx.x == 0) {
      loss_parts[oplane * 2 + 0] = static_cast<float>(ss);
      loss_parts[oplane * 2 + 1] = static_cast<float>(se);
    }
  }
}

}  // namespace skp

// =========================================================================================
// C ABI
// =========================================================================================
using namespace skp;

extern "C" int skp_plane_argmax(const float* maps, int N, int H, int W, int64_t* flat_idx, float* maxval, void* stream) {
  SKP_REQUIRE(maps && flat_idx && N > 0 && H > 0 && W > 0, "skp_plane_argmax: bad arguments");
  SKP_REQUIRE(static_cast<int64_t>(H) * W < (1ll << 32), "skp_plane_argmax: plane too large");
  cudaStream_t st = as_stream(stream);
  const int64_t plane = static_cast<int64_t>(H) * W;
  SKP_CUDA(cudaMemsetAsync(flat_idx, 0, sizeof(int64_t) * N, st));
  // enough CTAs to fill the machine even for a handful of large planes; chunk starts stay 16B aligned
  int chunks = static_cast<int>((plane + 16383) / 16384);
  int want = (4 * num_sms() + N - 1) / N;
  chunks = chunks < want ? chunks : want;
  chunks = chunks < 1 ? 1 : chunks;
  while (chunks > 1 && ((plane + chunks - 1) / chunks) % 4 != 0) --chunks;
  dim3 grid(chunks, N);
  plane_argmax_kernel<<<(::skp::note_launch(), grid), 256, 0, st>>>(maps, plane, reinterpret_cast<unsigned long long*>(flat_idx));
  SKP_LAUNCH_CHECK();
  argmax_decode_kernel<<<(::skp::note_launch(), (N + 127) / 128), 128, 0, st>>>(reinterpret_cast<unsigned long long*>(flat_idx), N, maxval);
  SKP_LAUNCH_CHECK();
  return SKP_OK;
}

extern "C" int skp_find_k_max_pixels(const float* maps, int N, int H, int W, int num, float* out, void* stream) {
  SKP_REQUIRE(maps && out && N > 0 && H > 0 && W > 0 && num >= 1 && num <= 8, "skp_find_k_max_pixels: bad arguments");
  k_max_pixels_kernel<<<(::skp::note_launch(), N), 512, 0, as_stream(stream)>>>(maps, N, H, W, num, out);
  SKP_LAUNCH_CHECK();
  return SKP_OK;
}

extern "C" int skp_weighted_avg_readout(float* maps, int N, int H, int W, float distance, int mutate, float* out, void* stream) {
  SKP_REQUIRE(maps && out && N > 0 && H > 0 && W > 0, "skp_weighted_avg_readout: bad arguments");
  weighted_avg_kernel<<<(::skp::note_launch(), N), 512, 0, as_stream(stream)>>>(maps, H, W, distance, mutate, out);
  SKP_LAUNCH_CHECK();
  return SKP_OK;
}

extern "C" int skp_plane_stats(const float* maps, int N, int H, int W, float sigma, float epsilon, int num_subjects,
                               float* kl, float* entropy, int64_t* argmax_out, void* stream) {
  SKP_REQUIRE(maps && N > 0 && H > 0 && W > 0 && num_subjects >= 1 && num_subjects <= 8, "skp_plane_stats: bad arguments");
  SKP_REQUIRE(kl || entropy || argmax_out, "skp_plane_stats: no output requested");
  const int64_t n = static_cast<int64_t>(H) * W;
  if (num_subjects == 1 && (W & (W - 1)) == 0 && (reinterpret_cast<uintptr_t>(maps) & 15) == 0) {
    cudaStream_t st = as_stream(stream);
    bool done = true;
    switch (n) {
      case PS_THREADS * 4 * 1: launch_plane_stats_reg<1>(maps, N, W, sigma, epsilon, kl, entropy, argmax_out, st); break;
      case PS_THREADS * 4 * 2: launch_plane_stats_reg<2>(maps, N, W, sigma, epsilon, kl, entropy, argmax_out, st); break;
      case PS_THREADS * 4 * 4: launch_plane_stats_reg<4>(maps, N, W, sigma, epsilon, kl, entropy, argmax_out, st); break;
      case PS_THREADS * 4 * 8: launch_plane_stats_reg<8>(maps, N, W, sigma, epsilon, kl, entropy, argmax_out, st); break;
      default: done = false;
    }
    if (done) {
      SKP_LAUNCH_CHECK();
      return SKP_OK;
    }
  }
  plane_stats_kernel<<<(::skp::note_launch(), N), 512, 0, as_stream(stream)>>>(maps, H, W, sigma, epsilon, num_subjects, kl, entropy, argmax_out);
  SKP_LAUNCH_CHECK();
  return SKP_OK;
}

extern "C" int skp_gaussian_circles(const float* pos, int S, int N, int size, float sigma, float* out, void* stream) {
  SKP_REQUIRE(pos && out && S >= 1 && N > 0 && size > 0, "skp_gaussian_circles: bad arguments");
  dim3 grid((size * size + 255) / 256, N);
  gaussian_circles_kernel<<<(::skp::note_launch(), grid), 256, 0, as_stream(stream)>>>(pos, S, N, size, sigma, out);
  SKP_LAUNCH_CHECK();
  return SKP_OK;
}

extern "C" int skp_rank_ascending_batched(const float* score, int B, int N, int top_k, int64_t* out, void* stream) {
  SKP_REQUIRE(score && out && B > 0 && N > 0 && N <= 4096 && top_k >= 0 && top_k <= N,
              "skp_rank_ascending: need B > 0, 0 < N <= 4096, top_k <= N");
  int P = 1;
  while (P < N) P <<= 1;
  rank_ascending_kernel<<<(::skp::note_launch(), B), 512, P * sizeof(unsigned long long), as_stream(stream)>>>(score, N, top_k, out);
  SKP_LAUNCH_CHECK();
  return SKP_OK;
}
extern "C" int skp_rank_ascending(const float* score, int N, int top_k, int64_t* out, void* stream) {
  return skp_rank_ascending_batched(score, 1, N, top_k, out, stream);
}

extern "C" int skp_furthest_point_sampling_batched(const int64_t* argmax_flat, int B, int N, int H, int W,
                                                   const int64_t* candidates, int n_cand, int top_k, int64_t* out,
                                                   int32_t* n_out, void* stream) {
  SKP_REQUIRE(argmax_flat && candidates && out && n_out, "skp_furthest_point_sampling: null pointer");
  SKP_REQUIRE(n_cand >= 2 && n_cand <= 1024 && top_k >= 2 && top_k <= n_cand && N > 0 && B > 0,
              "skp_furthest_point_sampling: need 2 <= top_k <= n_cand <= 1024 (the reference raises NameError below 2 candidates)");
  size_t smem = sizeof(float) * 2 * n_cand + sizeof(int) * (n_cand + top_k);
  fps_kernel<<<(::skp::note_launch(), B), 256, smem, as_stream(stream)>>>(argmax_flat, N, H, W, candidates, n_cand, top_k, out, n_out);
  SKP_LAUNCH_CHECK();
  return SKP_OK;
}
extern "C" int skp_furthest_point_sampling(const int64_t* argmax_flat, int N, int H, int W, const int64_t* candidates,
                                           int n_cand, int top_k, int64_t* out, int32_t* n_out, void* stream) {
  return skp_furthest_point_sampling_batched(argmax_flat, 1, N, H, W, candidates, n_cand, top_k, out, n_out, stream);
}

extern "C" int skp_keypoint_losses_fwd(const float* maps_a, const float* maps_b, const int64_t* idx, int K, int H, int W,
                                       const float* theta_inv, float sigma, int num_subjects, float* loss_parts, void* stream) {
  SKP_REQUIRE(maps_a && idx && loss_parts && K > 0 && H > 0 && W > 0, "skp_keypoint_losses_fwd: bad arguments");
  SKP_REQUIRE(!maps_b || theta_inv, "skp_keypoint_losses_fwd: theta_inv required with maps_b");
  SKP_REQUIRE(num_subjects >= 1 && num_subjects <= 8, "skp_keypoint_losses_fwd: 1 <= num_subjects <= 8");
  keypoint_losses_kernel<false><<<(::skp::note_launch(), K), 512, 0, as_stream(stream)>>>(maps_a, maps_b, idx, 0, 0, H, W, theta_inv, sigma, num_subjects,
                                                                  loss_parts, 0.f, 0.f, nullptr, nullptr, nullptr);
  SKP_LAUNCH_CHECK();
  return SKP_OK;
}

extern "C" int skp_keypoint_losses_batched_fwd(const float* maps_a, const float* maps_b, const int64_t* idx, int64_t idx_stride,
                                               int B, int planes, int K, int H, int W, const float* theta_inv, float sigma,
                                               int num_subjects, float* loss_parts, void* stream) {
  SKP_REQUIRE(maps_a && idx && loss_parts && B > 0 && B <= 65535 && K > 0 && planes > 0 && H > 0 && W > 0,
              "skp_keypoint_losses_batched_fwd: bad arguments");
  SKP_REQUIRE(!maps_b || theta_inv, "skp_keypoint_losses_batched_fwd: theta_inv required with maps_b");
  SKP_REQUIRE(num_subjects >= 1 && num_subjects <= 8, "skp_keypoint_losses_batched_fwd: 1 <= num_subjects <= 8");
  keypoint_losses_kernel<false><<<(::skp::note_launch(), dim3(K, B)), 512, 0, as_stream(stream)>>>(
      maps_a, maps_b, idx, idx_stride, planes, H, W, theta_inv, sigma, num_subjects, loss_parts, 0.f, 0.f, nullptr, nullptr, nullptr);
  SKP_LAUNCH_CHECK();
  return SKP_OK;
}

extern "C" int skp_keypoint_losses_batched_bwd(const float* maps_a, const float* maps_b, const int64_t* idx, int64_t idx_stride,
                                               int B, int planes, int K, int H, int W, const float* theta_inv, float sigma,
                                               int num_subjects, const float* g, float* d_a_sel, float* d_b_sel, void* stream) {
  SKP_REQUIRE(maps_a && idx && d_a_sel && g && B > 0 && B <= 65535 && K > 0 && planes > 0 && H > 0 && W > 0,
              "skp_keypoint_losses_batched_bwd: bad arguments");
  SKP_REQUIRE(!maps_b || theta_inv, "skp_keypoint_losses_batched_bwd: theta_inv required with maps_b");
  SKP_REQUIRE(num_subjects >= 1 && num_subjects <= 8, "skp_keypoint_losses_batched_bwd: 1 <= num_subjects <= 8");
  const float nrm = 2.0f / (static_cast<float>(K) * H * W);
  keypoint_losses_kernel<true><<<(::skp::note_launch(), dim3(K, B)), 512, 0, as_stream(stream)>>>(
      maps_a, maps_b, idx, idx_stride, planes, H, W, theta_inv, sigma, num_subjects, nullptr, nrm, nrm, g, d_a_sel, d_b_sel);
  SKP_LAUNCH_CHECK();
  return SKP_OK;
}

extern "C" int skp_keypoint_losses_bwd(const float* maps_a, const float* maps_b, const int64_t* idx, int K, int H, int W,
                                       const float* theta_inv, float sigma, int num_subjects, float g_sharp, float g_equiv,
                                       float* d_a_sel, float* d_b_sel, void* stream) {
  SKP_REQUIRE(maps_a && idx && d_a_sel && K > 0 && H > 0 && W > 0, "skp_keypoint_losses_bwd: bad arguments");
  SKP_REQUIRE(!maps_b || theta_inv, "skp_keypoint_losses_bwd: theta_inv required with maps_b");
  SKP_REQUIRE(num_subjects >= 1 && num_subjects <= 8, "skp_keypoint_losses_bwd: 1 <= num_subjects <= 8");
  const float nrm = 2.0f / (static_cast<float>(K) * H * W);
  keypoint_losses_kernel<true><<<(::skp::note_launch(), K), 512, 0, as_stream(stream)>>>(maps_a, maps_b, idx, 0, 0, H, W, theta_inv, sigma, num_subjects,
                                                                 nullptr, g_sharp * nrm, g_equiv * nrm, nullptr, d_a_sel, d_b_sel);
  SKP_LAUNCH_CHECK();
  return SKP_OK;
}
