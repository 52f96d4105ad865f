// Probe (GPU box only): cycles per 32-element chunk of the backward's dense pass for a softmax warp, 2 warps per scheduler
// (8 warps per CTA, one CTA per SM), adding the instructions one class at a time:
//   0: ex2 only   1: + (s - m) and * nd   2: + bf16x2 pack   3: + 16-byte shared-memory stores (swizzled like the dS tile)
// MUFU-bound would be 32 ex2 x 2 warps x 8 cycles = 512 cycles per chunk.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I stablekeypoints_b200/csrc scripts/dense_probe.cu -o build/dense_probe
#include <cstdio>
#include <cstdint>
#include <cuda_bf16.h>
#include "ptx_sm100.cuh"
using namespace skp;

template <int MODE>
__global__ void __launch_bounds__(256) dense(long long* cyc, float* out, const float* in, float m, float nd, int iters) {
  __shared__ __align__(1024) uint8_t tile[32768];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = (warp & 3) * 32 + lane, ch = warp >> 2;
  float v[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] = in[(threadIdx.x * 32 + j) & 1023];
  float acc = 0.f;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    uint32_t pk[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      float e0, e1;
      if (MODE == 0) {
        e0 = ptx::ex2(v[2 * j]);
        e1 = ptx::ex2(v[2 * j + 1]);
      } else {
        e0 = ptx::ex2(v[2 * j] - m) * nd;
        e1 = ptx::ex2(v[2 * j + 1] - m) * nd;
      }
      if (MODE >= 2) {
        const __nv_bfloat162 h2 = __floats2bfloat162_rn(e0, e1);
        pk[j] = *reinterpret_cast<const uint32_t*>(&h2);
      } else {
        acc += e0 + e1;
        pk[j] = 0;
      }
      v[2 * j] = e0 * 0.5f + v[2 * j] * 0.25f;        // keep the chain alive across iterations (2 extra FFMA per pair in every mode)
      v[2 * j + 1] = e1 * 0.5f + v[2 * j + 1] * 0.25f;
    }
    if (MODE >= 3) {
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int cc = ch * 128 + 32 * (it & 3) + 8 * q;
        const uint32_t off = ((cc >> 6) * 16384 + row * 128 + ((((cc & 63) >> 3) ^ (row & 7)) << 4)) % (32768 - 16);
        *reinterpret_cast<uint4*>(tile + (off & ~15u)) = make_uint4(pk[4 * q], pk[4 * q + 1], pk[4 * q + 2], pk[4 * q + 3]);
      }
    } else if (MODE == 2) {
#pragma unroll
      for (int j = 0; j < 16; ++j) acc += __uint_as_float(pk[j]);
    }
  }
  const long long t1 = clock64();
  out[blockIdx.x * 256 + threadIdx.x] = acc + v[0] + tile[threadIdx.x];
  if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}

template <int MODE>
void run(long long* cyc, float* out, const float* in, const char* what) {
  const int iters = 2048;
  dense<MODE><<<148, 256>>>(cyc, out, in, 0.3f, -0.7f, iters);
  cudaDeviceSynchronize();
  dense<MODE><<<148, 256>>>(cyc, out, in, 0.3f, -0.7f, iters);
  cudaError_t err = cudaDeviceSynchronize();
  long long h = 0;
  cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  printf("mode %d (%s): %.1f cycles per 32-element chunk (2 warps per scheduler)  [%s]\n", MODE, what, double(h) / iters, cudaGetErrorString(err));
}

int main() {
  long long* cyc;
  float *out, *in;
  cudaMalloc(&cyc, 8);
  cudaMalloc(&out, 148 * 256 * 4);
  cudaMalloc(&in, 4096);
  cudaMemset(in, 0, 4096);
  run<0>(cyc, out, in, "ex2 + 2 FFMA");
  run<1>(cyc, out, in, "+ sub, mul");
  run<2>(cyc, out, in, "+ bf16x2 pack");
  run<3>(cyc, out, in, "+ STS.128");
  return 0;
}
