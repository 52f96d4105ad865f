// Probe (GPU box only): issue-to-completion rate of SS-mode tcgen05.mma kind::f16 M=128, K=16 for N = 64 / 128 / 256
// (128B-swizzled K-major operands, the four 32-byte K slices of a 64-wide block in turn), with one and with two CTAs
// per SM, and the tcgen05.ld 32x32b.x32 rate with 4 / 8 / 16 warps per CTA.  Sizes the forward kernel's tile choice.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -std=c++17 -I stablekeypoints_b200/csrc scripts/mma_rate_probe.cu -o gpurun_out/mma_rate_probe
#include <cstdio>
#include <cstdint>
#include <cuda_bf16.h>
#include "ptx_sm100.cuh"
using namespace skp;

constexpr int ITERS = 4096;

template <int N>
__global__ void __launch_bounds__(128) mma_rate(long long* cyc, int tmem_cols) {
  extern __shared__ uint8_t raw[];
  const uint32_t base = (ptx::smem_u32(raw) + 1023u) & ~1023u;
  uint8_t* bp = raw + (base - ptx::smem_u32(raw));
  const uint32_t a = base, b = base + 16384;             // A 128 x 64 bf16, B 256 x 64 bf16
  uint32_t* tptr = reinterpret_cast<uint32_t*>(bp + 16384 + 32768);
  const uint32_t bar = base + 16384 + 32768 + 16;
  for (int i = threadIdx.x; i < (16384 + 32768) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(bp)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) { ptx::mbar_init(bar, 1); ptx::fence_barrier_init(); }
  if (threadIdx.x < 32) { ptx::tmem_alloc(ptx::smem_u32(tptr), tmem_cols); ptx::tmem_relinquish(); }
  ptx::fence_proxy_async();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(tptr);
  if (threadIdx.x == 0) {
    constexpr uint32_t idesc = ptx::umma_idesc_bf16(128, N);
    const long long t0 = clock64();
    for (int it = 0; it < ITERS; ++it) {
      const int kk = it & 3;
      ptx::umma_bf16(tmem, ptx::umma_desc_sw128(a + kk * 32), ptx::umma_desc_sw128(b + kk * 32), idesc, 1u);
    }
    ptx::umma_commit(bar);
    ptx::mbar_wait(bar, 0u);
    const long long t1 = clock64();
    if (blockIdx.x == 0) cyc[0] = t1 - t0;
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) { ptx::tc_fence_after(); ptx::tmem_dealloc(tmem, tmem_cols); }
}

__global__ void __launch_bounds__(512) ld_rate(long long* cyc, float* out) {
  __shared__ uint32_t tptr;
  if (threadIdx.x < 32) { ptx::tmem_alloc(ptx::smem_u32(&tptr), 512); ptx::tmem_relinquish(); }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = tptr;
  const int warp = threadIdx.x >> 5;
  const uint32_t t_lane = tmem + (static_cast<uint32_t>((warp & 3) * 32) << 16) + (warp >> 2) * 64;
  float s = 0.f;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < 1024; ++it) {
    uint32_t v[32];
    ptx::tmem_ld32(t_lane + (it & 1) * 32, v);
    ptx::tmem_ld_wait();
    s += __uint_as_float(v[0]) + __uint_as_float(v[17]) + __uint_as_float(v[31]);
  }
  const long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
  ptx::tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) { ptx::tc_fence_after(); ptx::tmem_dealloc(tmem, 512); }
}

template <int N>
void run(int ctas_per_sm, long long* cyc) {
  const int smem = 16384 + 32768 + 64 + 1024;
  cudaFuncSetAttribute(mma_rate<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  mma_rate<N><<<148 * ctas_per_sm, 128, smem>>>(cyc, 512 / ctas_per_sm);
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  mma_rate<N><<<148 * ctas_per_sm, 128, smem>>>(cyc, 512 / ctas_per_sm);
  cudaEventRecord(e1);
  cudaError_t err = cudaDeviceSynchronize();
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  long long h = 0;
  cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  printf("N=%3d  %d CTA/SM: %7.1f cycles per MMA in CTA 0 (clock64), kernel %.3f ms = %.1f ns per MMA per SM  [%s]\n", N, ctas_per_sm,
         double(h) / ITERS, ms, ms * 1e6 / (ITERS * ctas_per_sm), cudaGetErrorString(err));
}

int main() {
  long long* cyc;
  float* out;
  cudaMalloc(&cyc, 8);
  cudaMalloc(&out, 148 * 512 * 4);
  for (int c = 1; c <= 2; ++c) {
    run<64>(c, cyc);
    run<128>(c, cyc);
    run<256>(c, cyc);
  }
  for (int threads : {128, 256, 512}) {
    ld_rate<<<148, threads>>>(cyc, out);
    cudaError_t err = cudaDeviceSynchronize();
    long long h = 0;
    cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    printf("tcgen05.ld 32x32b.x32 + wait, %2d warps/CTA: %.1f cycles per load per warp, %.1f per load per SMSP  [%s]\n", threads / 32,
           double(h) / 1024, double(h) / 1024 / (threads / 128), cudaGetErrorString(err));
  }
  return 0;
}
