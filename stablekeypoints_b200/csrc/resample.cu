// Resampling kernels: bilinear / bicubic upsample, affine warps and the TTA ensemble.
// All bandwidth-bound: one thread per output element (or per 4 along x), coalesced stores,
// gathers served by L1/L2 (sources are 4x..64x smaller than the outputs).
#include "common.cuh"

namespace skp {

// F.interpolate(mode="bilinear", align_corners=False): (N,h,w) -> (N,H,W)
__global__ void bilinear_upsample_kernel(const float* __restrict__ in, int h, int w, float* __restrict__ out, int H, int W) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y;
  const int n = blockIdx.z;
  if (x >= W) return;
  int y0, y1, x0, x1;
  float ly, lx;
  bilinear_coeffs(y, static_cast<float>(h) / H, h, &y0, &y1, &ly);
  bilinear_coeffs(x, static_cast<float>(w) / W, w, &x0, &x1, &lx);
  const float* p = in + static_cast<int64_t>(n) * h * w;
  float v = (1.0f - ly) * ((1.0f - lx) * p[y0 * w + x0] + lx * p[y0 * w + x1]) +
            ly * ((1.0f - lx) * p[y1 * w + x0] + lx * p[y1 * w + x1]);
  out[(static_cast<int64_t>(n) * H + y) * W + x] = v;
}

// Same, W % 4 == 0: one CTA per (plane, band of rows), four outputs per thread written as one 16-byte store; the
// source plane stays in L1, so the kernel is bound by the output write (the per-pixel version launched one
// 128-thread CTA per half row and reached 19 % of the copy bandwidth).
constexpr int BILINEAR_ROWS = 32;
__global__ void __launch_bounds__(256) bilinear_upsample_vec4_kernel(const float* __restrict__ in, int h, int w,
                                                                       float* __restrict__ out, int H, int W) {
  const int n = blockIdx.y;
  const int ybase = blockIdx.x * BILINEAR_ROWS;
  const int rows = min(BILINEAR_ROWS, H - ybase);
  const int w4 = W >> 2;
  const float* p = in + static_cast<int64_t>(n) * h * w;
  float4* o = reinterpret_cast<float4*>(out + (static_cast<int64_t>(n) * H + ybase) * W);
  const float ry = static_cast<float>(h) / H, rx = static_cast<float>(w) / W;
  if (blockDim.x % w4 == 0) {
    // every thread keeps one column group for all its rows: the x taps / weights are computed once
    const int x4 = threadIdx.x % w4, ystep = blockDim.x / w4;
    int x0[4], x1[4];
    float lx[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) bilinear_coeffs(4 * x4 + j, rx, w, &x0[j], &x1[j], &lx[j]);
    // ... and walks CONSECUTIVE output rows, keeping the x-interpolated values of the two source rows it last touched:
    // at x2 / x4 upsampling a source row serves 2-4 output rows twice (as upper and as lower tap), so the 16 gathers per
    // 16-byte store drop to ~4 (the kernel was LSU-bound at 55 % of the copy bandwidth, not store-bound)
    const int per = (rows + ystep - 1) / ystep;
    const int yb = (threadIdx.x / w4) * per, ye = min(rows, yb + per);
    int cy0 = -1, cy1 = -1;
    float a0[4], a1[4];
    auto xrow = [&](int yy_src, float (&a)[4]) {
      const float* r = p + yy_src * w;
#pragma unroll
      for (int j = 0; j < 4; ++j) a[j] = (1.0f - lx[j]) * __ldg(r + x0[j]) + lx[j] * __ldg(r + x1[j]);
    };
    for (int yy = yb; yy < ye; ++yy) {
      int y0, y1;
      float ly;
      bilinear_coeffs(ybase + yy, ry, h, &y0, &y1, &ly);
      if (y0 != cy0) {
        if (y0 == cy1) {
#pragma unroll
          for (int j = 0; j < 4; ++j) a0[j] = a1[j];
        } else {
          xrow(y0, a0);
        }
        cy0 = y0;
      }
      if (y1 != cy1) {
        if (y1 == cy0) {
#pragma unroll
          for (int j = 0; j < 4; ++j) a1[j] = a0[j];
        } else {
          xrow(y1, a1);
        }
        cy1 = y1;
      }
      o[yy * w4 + x4] = make_float4((1.0f - ly) * a0[0] + ly * a1[0], (1.0f - ly) * a0[1] + ly * a1[1],
                                    (1.0f - ly) * a0[2] + ly * a1[2], (1.0f - ly) * a0[3] + ly * a1[3]);
    }
    return;
  }
  for (int i = threadIdx.x; i < rows * w4; i += blockDim.x) {
    const int yy = i / w4, x4 = i - yy * w4;
    int y0, y1;
    float ly;
    bilinear_coeffs(ybase + yy, ry, h, &y0, &y1, &ly);
    const float* r0 = p + y0 * w;
    const float* r1 = p + y1 * w;
    float v[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int x0, x1;
      float lx;
      bilinear_coeffs(4 * x4 + j, rx, w, &x0, &x1, &lx);
      v[j] = (1.0f - ly) * ((1.0f - lx) * __ldg(r0 + x0) + lx * __ldg(r0 + x1)) +
             ly * ((1.0f - lx) * __ldg(r1 + x0) + lx * __ldg(r1 + x1));
    }
    o[i] = make_float4(v[0], v[1], v[2], v[3]);
  }
}

// F.interpolate(mode="bicubic", align_corners=False) on channels-last data (B,s,s,C)->(B,R,R,C)
__global__ void bicubic_upsample_nhwc_kernel(const float* __restrict__ in, int s, int C, float* __restrict__ out, int R) {
  const int ox = blockIdx.x % R;
  const int c = (blockIdx.x / R) * blockDim.x + threadIdx.x;
  const int oy = blockIdx.y;
  const int pix = oy * R + ox;         // output pixel
  const int b = blockIdx.z;
  if (c >= C) return;
  int iy, ix;
  float wy[4], wx[4];
  const float ratio = static_cast<float>(s) / R;
  bicubic_coeffs(oy, ratio, &iy, wy);
  bicubic_coeffs(ox, ratio, &ix, wx);
  const float* p = in + static_cast<int64_t>(b) * s * s * C + c;
  float acc = 0.0f;
#pragma unroll
  for (int a = 0; a < 4; ++a) {
    int yy = min(max(iy - 1 + a, 0), s - 1);
    float row = 0.0f;
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      int xx = min(max(ix - 1 + e, 0), s - 1);
      row += wx[e] * p[static_cast<int64_t>(yy * s + xx) * C];
    }
    acc += wy[a] * row;
  }
  out[(static_cast<int64_t>(b) * R * R + pix) * C + c] = acc;
}

__device__ __forceinline__ void affine_src(const float* th, int y, int x, int H, int W, float* ix, float* iy) {
  float xs = (2.0f * x + 1.0f) / W - 1.0f;
  float ys = (2.0f * y + 1.0f) / H - 1.0f;
  float gx = th[0] * xs + th[1] * ys + th[2];
  float gy = th[3] * xs + th[4] * ys + th[5];
  *ix = ((gx + 1.0f) * W - 1.0f) * 0.5f;
  *iy = ((gy + 1.0f) * H - 1.0f) * 0.5f;
}

// grid_sample(img, affine_grid(theta)) bilinear / zeros / align_corners=False
template <bool kAdjoint>
__global__ void affine_warp_kernel(const float* __restrict__ src, const float* __restrict__ theta, int C, int H, int W,
                                   float* dst) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y;
  const int n = blockIdx.z;
  if (x >= W) return;
  float th[6];
#pragma unroll
  for (int j = 0; j < 6; ++j) th[j] = theta[n * 6 + j];
  float ix, iy;
  affine_src(th, y, x, H, W, &ix, &iy);
  const float x0f = floorf(ix), y0f = floorf(iy);
  const int x0 = static_cast<int>(x0f), y0 = static_cast<int>(y0f);
  const float fx = ix - x0f, fy = iy - y0f;
  const bool vx0 = x0 >= 0 && x0 < W, vx1 = x0 + 1 >= 0 && x0 + 1 < W;
  const bool vy0 = y0 >= 0 && y0 < H, vy1 = y0 + 1 >= 0 && y0 + 1 < H;
  const float w00 = (1.0f - fx) * (1.0f - fy), w01 = fx * (1.0f - fy), w10 = (1.0f - fx) * fy, w11 = fx * fy;
  const int64_t plane = static_cast<int64_t>(H) * W;
  for (int c = 0; c < C; ++c) {
    const int64_t base = (static_cast<int64_t>(n) * C + c) * plane;
    if (!kAdjoint) {
      const float* p = src + base;
      float v = 0.0f;
      if (vy0 && vx0) v += p[y0 * W + x0] * w00;
      if (vy0 && vx1) v += p[y0 * W + x0 + 1] * w01;
      if (vy1 && vx0) v += p[(y0 + 1) * W + x0] * w10;
      if (vy1 && vx1) v += p[(y0 + 1) * W + x0 + 1] * w11;
      dst[base + y * W + x] = v;
    } else {
      const float g = src[base + y * W + x];
      float* p = dst + base;
      if (vy0 && vx0) atomicAdd(p + y0 * W + x0, g * w00);
      if (vy0 && vx1) atomicAdd(p + y0 * W + x0 + 1, g * w01);
      if (vy1 && vx0) atomicAdd(p + (y0 + 1) * W + x0, g * w10);
      if (vy1 && vx1) atomicAdd(p + (y0 + 1) * W + x0 + 1, g * w11);
    }
  }
}

// TTA accumulation (eval.py:260-266): for each output pixel, over the A augmentations of this
// call: num += unwarp(ones), sum_k += unwarp(upsample(maps_k)).  The (K,H,W) upsampled maps are
// never materialised: each of the 4 un-warp taps is itself a 4-tap bilinear read of the
// feature-resolution plane (L2 resident).
__global__ void tta_accumulate_kernel(const float* __restrict__ maps, const float* __restrict__ theta_inv, int A, int K,
                                      int h, int w, float* sum_, float* num_, int H, int W) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y;
  if (x >= W) return;
  const int64_t plane = static_cast<int64_t>(H) * W;
  const int64_t o = static_cast<int64_t>(y) * W + x;
  const int64_t lplane = static_cast<int64_t>(h) * w;
  constexpr int KMAX = 32;
  float acc[KMAX];
#pragma unroll
  for (int k = 0; k < KMAX; ++k) acc[k] = 0.0f;
  float cover = 0.0f;
  const bool same = (h == H && w == W);
  const float ry = static_cast<float>(h) / H, rx = static_cast<float>(w) / W;
  for (int a = 0; a < A; ++a) {
    float th[6];
#pragma unroll
    for (int j = 0; j < 6; ++j) th[j] = theta_inv[a * 6 + j];
    float ix, iy;
    affine_src(th, y, x, H, W, &ix, &iy);
    const float x0f = floorf(ix), y0f = floorf(iy);
    const int x0 = static_cast<int>(x0f), y0 = static_cast<int>(y0f);
    const float fx = ix - x0f, fy = iy - y0f;
    const float wgt[4] = {(1.0f - fx) * (1.0f - fy), fx * (1.0f - fy), (1.0f - fx) * fy, fx * fy};
    const float* base = maps + static_cast<int64_t>(a) * K * lplane;
    for (int t = 0; t < 4; ++t) {
      const int X = x0 + (t & 1), Y = y0 + (t >> 1);
      if (X < 0 || X >= W || Y < 0 || Y >= H) continue;
      cover += wgt[t];
      int sy0 = Y, sy1 = Y, sx0 = X, sx1 = X;
      float ly = 0.0f, lx = 0.0f;
      if (!same) {
        bilinear_coeffs(Y, ry, h, &sy0, &sy1, &ly);
        bilinear_coeffs(X, rx, w, &sx0, &sx1, &lx);
      }
      const int o00 = sy0 * w + sx0, o01 = sy0 * w + sx1, o10 = sy1 * w + sx0, o11 = sy1 * w + sx1;
#pragma unroll
      for (int k = 0; k < KMAX; ++k) {
        if (k < K) {
          const float* p = base + k * lplane;
          float v = same ? p[o00]
                         : (1.0f - ly) * ((1.0f - lx) * p[o00] + lx * p[o01]) + ly * ((1.0f - lx) * p[o10] + lx * p[o11]);
          acc[k] += wgt[t] * v;
        }
      }
    }
  }
#pragma unroll
  for (int k = 0; k < KMAX; ++k) {
    if (k < K) {
      sum_[k * plane + o] += acc[k];
      num_[k * plane + o] += cover;
    }
  }
}

__global__ void tta_finalize_kernel(const float* __restrict__ sum_, const float* __restrict__ num_, int64_t n, float* out) {
  int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float v = sum_[i] / num_[i];
  out[i] = (v != v) ? 0.0f : v;
}

}  // namespace skp

using namespace skp;

extern "C" int skp_bilinear_upsample(const float* in, int N, int h, int w, float* out, int H, int W, void* stream) {
  SKP_REQUIRE(in && out && N > 0 && h > 0 && w > 0 && H > 0 && W > 0, "skp_bilinear_upsample: bad arguments");
  SKP_REQUIRE(H <= 65535 && N <= 65535, "skp_bilinear_upsample: grid too large");
  if ((W & 3) == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0) {
    dim3 grid4((H + BILINEAR_ROWS - 1) / BILINEAR_ROWS, N);
    bilinear_upsample_vec4_kernel<<<(::skp::note_launch(), grid4), 256, 0, as_stream(stream)>>>(in, h, w, out, H, W);
    SKP_LAUNCH_CHECK();
    return SKP_OK;
  }
  dim3 grid((W + 127) / 128, H, N);
  bilinear_upsample_kernel<<<(::skp::note_launch(), grid), 128, 0, as_stream(stream)>>>(in, h, w, out, H, W);
  SKP_LAUNCH_CHECK();
  return SKP_OK;
}

extern "C" int skp_bicubic_upsample_nhwc(const float* in, int B, int s, int C, float* out, int R, void* stream) {
  SKP_REQUIRE(in && out && B > 0 && s > 0 && C > 0 && R > 0, "skp_bicubic_upsample_nhwc: bad arguments");
  SKP_REQUIRE(B <= 65535 && R <= 65535, "skp_bicubic_upsample_nhwc: grid too large");
  dim3 grid(((C + 127) / 128) * R, R, B);
  bicubic_upsample_nhwc_kernel<<<(::skp::note_launch(), grid), 128, 0, as_stream(stream)>>>(in, s, C, out, R);
  SKP_LAUNCH_CHECK();
  return SKP_OK;
}

extern "C" int skp_affine_warp(const float* img, const float* theta, int N, int C, int H, int W, float* out, void* stream) {
  SKP_REQUIRE(img && theta && out && N > 0 && C > 0 && H > 0 && W > 0, "skp_affine_warp: bad arguments");
  SKP_REQUIRE(H <= 65535 && N <= 65535, "skp_affine_warp: grid too large");
  dim3 grid((W + 127) / 128, H, N);
  affine_warp_kernel<false><<<(::skp::note_launch(), grid), 128, 0, as_stream(stream)>>>(img, theta, C, H, W, out);
  SKP_LAUNCH_CHECK();
  return SKP_OK;
}

extern "C" int skp_affine_warp_bwd(const float* dout, const float* theta, int N, int C, int H, int W, float* dimg, void* stream) {
  SKP_REQUIRE(dout && theta && dimg && N > 0 && C > 0 && H > 0 && W > 0, "skp_affine_warp_bwd: bad arguments");
  SKP_REQUIRE(H <= 65535 && N <= 65535, "skp_affine_warp_bwd: grid too large");
  dim3 grid((W + 127) / 128, H, N);
  affine_warp_kernel<true><<<(::skp::note_launch(), grid), 128, 0, as_stream(stream)>>>(dout, theta, C, H, W, dimg);
  SKP_LAUNCH_CHECK();
  return SKP_OK;
}

extern "C" int skp_tta_accumulate(const float* maps, const float* theta_inv, int A, int K, int h, int w, float* sum_,
                                  float* num_, int H, int W, void* stream) {
  SKP_REQUIRE(maps && theta_inv && sum_ && num_ && A > 0 && K > 0 && K <= 32 && h > 0 && w > 0 && H > 0 && W > 0,
              "skp_tta_accumulate: bad arguments (K <= 32)");
  SKP_REQUIRE(H <= 65535, "skp_tta_accumulate: grid too large");
  dim3 grid((W + 127) / 128, H);
  tta_accumulate_kernel<<<(::skp::note_launch(), grid), 128, 0, as_stream(stream)>>>(maps, theta_inv, A, K, h, w, sum_, num_, H, W);
  SKP_LAUNCH_CHECK();
  return SKP_OK;
}

extern "C" int skp_tta_finalize(const float* sum_, const float* num_, int64_t n, float* out, void* stream) {
  SKP_REQUIRE(sum_ && num_ && out && n > 0, "skp_tta_finalize: bad arguments");
  tta_finalize_kernel<<<(::skp::note_launch(), static_cast<unsigned>((n + 255) / 256)), 256, 0, as_stream(stream)>>>(sum_, num_, n, out);
  SKP_LAUNCH_CHECK();
  return SKP_OK;
}
