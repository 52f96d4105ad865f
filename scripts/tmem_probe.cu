// Probe (GPU box only): tcgen05.ld / tcgen05.st throughput per SM as a function of warp count and vector width,
// and whether tensor-core MMAs running concurrently slow the loads down.  The forward kernel's softmax reads each
// logit tile three times from TMEM; this decides whether that is the bound.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -std=c++17 -I stablekeypoints_b200/csrc scripts/tmem_probe.cu -o build/tmem_probe
#include <cstdio>
#include <cstdlib>
#include <cuda_bf16.h>
#include "ptx_sm100.cuh"
using namespace skp;

template <int W>
__device__ __forceinline__ void ld_n(uint32_t taddr, uint32_t& sink) {
  if constexpr (W == 16) {
    uint32_t v[16];
    ptx::tmem_ld16(taddr, v);
    ptx::tmem_ld_wait();
#pragma unroll
    for (int j = 0; j < 16; ++j) sink ^= v[j];
  } else {
    uint32_t v[32];
    ptx::tmem_ld32(taddr, v);
    ptx::tmem_ld_wait();
#pragma unroll
    for (int j = 0; j < 32; ++j) sink ^= v[j];
  }
}

// mode 0: x16 loads, one in flight; 1: x32 loads, one in flight; 2: x16 loads, two in flight; 3: x32 stores
__global__ void __launch_bounds__(512, 1) probe(int mode, int iters, int nwarps, long long* cycles, uint32_t* out) {
  __shared__ uint32_t tptr;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) { ptx::tmem_alloc(ptx::smem_u32(&tptr), 512); ptx::tmem_relinquish(); }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tm = tptr;
  const uint32_t t_lane = tm + (static_cast<uint32_t>((warp & 3) * 32) << 16);
  const int grp = warp >> 2;                      // warps sharing a lane quadrant take different column ranges
  const int ngrp = (nwarps + 3) / 4;
  const int span = 512 / ngrp;                    // columns per warp
  uint32_t sink = 0;
  __syncthreads();
  const long long t0 = clock64();
  if (warp < nwarps) {
    const uint32_t c0 = grp * span;
    if (mode == 0) {
      for (int it = 0; it < iters; ++it)
        for (int c = 0; c < span; c += 16) ld_n<16>(t_lane + c0 + c, sink);
    } else if (mode == 1) {
      for (int it = 0; it < iters; ++it)
        for (int c = 0; c < span; c += 32) ld_n<32>(t_lane + c0 + c, sink);
    } else if (mode == 2) {
      for (int it = 0; it < iters; ++it)
        for (int c = 0; c < span; c += 32) {
          uint32_t a[16], b[16];
          ptx::tmem_ld16(t_lane + c0 + c, a);
          ptx::tmem_ld16(t_lane + c0 + c + 16, b);
          ptx::tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j) sink ^= a[j] ^ b[j];
        }
    } else {
      uint32_t v[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] = threadIdx.x + j;
      for (int it = 0; it < iters; ++it)
        for (int c = 0; c < span; c += 32) { ptx::tmem_st32(t_lane + c0 + c, v); }
      ptx::tmem_st_wait();
    }
  }
  __syncthreads();
  const long long t1 = clock64();
  if (threadIdx.x == 0 && blockIdx.x == 0) cycles[0] = t1 - t0;
  out[blockIdx.x * blockDim.x + threadIdx.x] = sink;
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 0) { ptx::tc_fence_after(); ptx::tmem_dealloc(tm, 512); }
}

int main() {
  long long* cyc;
  uint32_t* out;
  cudaMalloc(&cyc, 8);
  cudaMalloc(&out, 148 * 512 * 4);
  const char* names[4] = {"ld x16 (1 in flight)", "ld x32 (1 in flight)", "ld x16 (2 in flight)", "st x32"};
  for (int mode = 0; mode < 4; ++mode)
    for (int nw : {4, 8, 16}) {
      const int iters = 200;
      probe<<<148, 512>>>(mode, iters, nw, cyc, out);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
      long long h;
      cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
      // every warp quadrant group covers its span; all groups together cover 128 lanes x 512 cols x 4 B = 256 KB per iteration
      // when nwarps >= 4 (with 4 warps: one group spanning all 512 columns)
      const double bytes = 128.0 * 512 * 4 * iters;
      printf("%-22s warps %2d : %9lld cycles  %7.1f B/clk/SM\n", names[mode], nw, h, bytes / h);
    }
  return 0;
}
