// Shared helpers for the sm_100a kernels of libskp_b200.so.
#pragma once

#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdarg>
#include <cmath>

#include "../../include/skp_b200.h"

namespace skp {

// ---- error plumbing -------------------------------------------------------------------
void set_error(const char* fmt, ...);

#define SKP_REQUIRE(cond, ...)                    \
  do {                                            \
    if (!(cond)) {                                \
      ::skp::set_error(__VA_ARGS__);              \
      return SKP_ERR_INVALID;                     \
    }                                             \
  } while (0)

#define SKP_CUDA(expr)                                                              \
  do {                                                                              \
    cudaError_t _e = (expr);                                                        \
    if (_e != cudaSuccess) {                                                        \
      ::skp::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e),     \
                       __FILE__, __LINE__);                                         \
      return SKP_ERR_CUDA;                                                          \
    }                                                                               \
  } while (0)

#define SKP_LAUNCH_CHECK() SKP_CUDA(cudaPeekAtLastError())

// process-wide count of kernels this library has launched (skp_kernel_launches(); bench.py's gpu_launches)
int note_launch();

inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

// SM count of the CURRENT device (queried once per device ordinal; 148 on B200): sizes persistent / machine-filling grids
int num_sms();

// ---- warp / block reductions ------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ unsigned long long warp_max_u64(unsigned long long v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    unsigned long long w = __shfl_xor_sync(0xffffffffu, v, o);
    v = w > v ? w : v;
  }
  return v;
}

// block-wide reductions; `scratch` must hold >= 32 elements; result valid in all threads.
template <typename T>
__device__ __forceinline__ T block_sum(T v, T* scratch) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = (blockDim.x + 31) >> 5;
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) scratch[warp] = v;
  __syncthreads();
  T r = (lane < nwarp) ? scratch[lane] : T(0);
  r = warp_sum(r);
  return r;
}
__device__ __forceinline__ float block_max(float v, float* scratch) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = (blockDim.x + 31) >> 5;
  v = warp_max(v);
  __syncthreads();
  if (lane == 0) scratch[warp] = v;
  __syncthreads();
  float r = (lane < nwarp) ? scratch[lane] : -INFINITY;
  return warp_max(r);
}
__device__ __forceinline__ unsigned long long block_max_u64(unsigned long long v, unsigned long long* scratch) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = (blockDim.x + 31) >> 5;
  v = warp_max_u64(v);
  __syncthreads();
  if (lane == 0) scratch[warp] = v;
  __syncthreads();
  unsigned long long r = (lane < nwarp) ? scratch[lane] : 0ull;
  return warp_max_u64(r);
}

// ---- argmax key: max value, FIRST index on ties (torch.argmax semantics; NaN counts as max) ----
__device__ __forceinline__ unsigned long long argmax_key(float v, uint32_t idx) {
  uint32_t b = __float_as_uint(v);
  uint32_t ord = (v != v) ? 0xffffffffu : ((b & 0x80000000u) ? ~b : (b | 0x80000000u));
  return (static_cast<unsigned long long>(ord) << 32) | static_cast<unsigned long long>(0xffffffffu - idx);
}
__device__ __forceinline__ uint32_t argmax_key_index(unsigned long long key) {
  return 0xffffffffu - static_cast<uint32_t>(key & 0xffffffffull);
}
__device__ __forceinline__ float argmax_key_value(unsigned long long key) {
  uint32_t ord = static_cast<uint32_t>(key >> 32);
  if (ord == 0xffffffffu) return __uint_as_float(0x7fc00000u);
  uint32_t b = (ord & 0x80000000u) ? (ord & 0x7fffffffu) : ~ord;
  return __uint_as_float(b);
}

// ---- interpolation coefficient helpers (SURVEY.md Appendix A == ATen upsample kernels) ----
// bicubic (Keys, A = -0.75), align_corners=False: source index and 4 weights for output o.
__device__ __host__ __forceinline__ void bicubic_coeffs(int o, float ratio /* in/out */, int* i0, float w[4]) {
  const float A = -0.75f;
  float src = (o + 0.5f) * ratio - 0.5f;
  float fl = floorf(src);
  float t = src - fl;
  *i0 = static_cast<int>(fl);
  float u;
  u = t + 1.0f;  w[0] = ((A * u - 5.0f * A) * u + 8.0f * A) * u - 4.0f * A;
  u = t;         w[1] = ((A + 2.0f) * u - (A + 3.0f)) * u * u + 1.0f;
  u = 1.0f - t;  w[2] = ((A + 2.0f) * u - (A + 3.0f)) * u * u + 1.0f;
  u = 2.0f - t;  w[3] = ((A * u - 5.0f * A) * u + 8.0f * A) * u - 4.0f * A;
}
// bilinear, align_corners=False
__device__ __host__ __forceinline__ void bilinear_coeffs(int o, float ratio, int n_in, int* i0, int* i1, float* lam) {
  float src = (o + 0.5f) * ratio - 0.5f;
  src = src < 0.0f ? 0.0f : src;
  int a = static_cast<int>(src);
  a = a < n_in - 1 ? a : n_in - 1;
  *i0 = a;
  *i1 = a + 1 < n_in - 1 ? a + 1 : n_in - 1;
  *lam = src - a;
}

}  // namespace skp
