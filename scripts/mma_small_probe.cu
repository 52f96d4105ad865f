// Probe (GPU box only): how long the tensor pipe takes per SS-mode tcgen05.mma kind::f16 (M=128, K=16) for the SMALL
// products of the backward kernel: N = 16 (selected tokens' logits) and N = 64 (logit-gradient window), K-major and
// MN-major operands, accumulating into ONE TMEM accumulator or rotating over 2 / 4 of them, and M = 64 x N = 256.
// Answers whether the ~100 cycles per instruction seen in the kernel's issue trace are a same-accumulator dependency.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -std=c++17 -I stablekeypoints_b200/csrc scripts/mma_small_probe.cu -o build/mma_small_probe
#include <cstdio>
#include <cstdint>
#include <cuda_bf16.h>
#include "ptx_sm100.cuh"
using namespace skp;

constexpr int ITERS = 4096;

__device__ __forceinline__ uint64_t mn_desc(uint32_t addr, uint32_t lbo, uint32_t sbo) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((addr & 0x3FFFFu) >> 4);
  d |= static_cast<uint64_t>(lbo >> 4) << 16;
  d |= static_cast<uint64_t>(sbo >> 4) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(2) << 61;
  return d;
}
__device__ __host__ constexpr uint32_t idesc_of(int M, int N, int a_mn, int b_mn) {
  return (1u << 4) | (1u << 7) | (1u << 10) | (static_cast<uint32_t>(a_mn) << 15) | (static_cast<uint32_t>(b_mn) << 16) |
         (static_cast<uint32_t>(N >> 3) << 17) | (static_cast<uint32_t>(M >> 4) << 24);
}

// MODE 0: K-major A and B (S / S_sel products); 1: MN-major A and B (dz product: A = 64 KB dS tile, B = 16 KB W tile)
template <int M, int N, int NACC, int MODE>
__global__ void __launch_bounds__(128) mma_rate(long long* cyc) {
  extern __shared__ uint8_t raw[];
  const uint32_t base = (ptx::smem_u32(raw) + 1023u) & ~1023u;
  uint8_t* bp = raw + (base - ptx::smem_u32(raw));
  const uint32_t a = base, b = base + 65536;             // A up to 64 KB, B up to 32 KB
  uint32_t* tptr = reinterpret_cast<uint32_t*>(bp + 65536 + 32768);
  const uint32_t bar = base + 65536 + 32768 + 16;
  for (int i = threadIdx.x; i < (65536 + 32768) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(bp)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) { ptx::mbar_init(bar, 1); ptx::fence_barrier_init(); }
  if (threadIdx.x < 32) { ptx::tmem_alloc(ptx::smem_u32(tptr), 512); ptx::tmem_relinquish(); }
  ptx::fence_proxy_async();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(tptr);
  if (threadIdx.x == 0) {
    constexpr uint32_t idesc = idesc_of(M, N, MODE, MODE);
    const long long t0 = clock64();
#pragma unroll 8
    for (int it = 0; it < ITERS; ++it) {
      const int kk = it & 7;
      const uint32_t acc = tmem + (it % NACC) * (N < 32 ? 32 : N);
      if (MODE == 0)
        ptx::umma_bf16(acc, ptx::umma_desc_sw128(a + (kk & 3) * 32), ptx::umma_desc_sw128(b + (kk & 3) * 32), idesc, 1u);
      else
        ptx::umma_bf16(acc, mn_desc(a + kk * 2048, 16384, 1024), mn_desc(b + kk * 2048, 16384, 1024), idesc, 1u);
    }
    const long long t_issue = clock64();
    ptx::umma_commit(bar);
    ptx::mbar_wait(bar, 0u);
    const long long t1 = clock64();
    if (blockIdx.x == 0) { cyc[0] = t1 - t0; cyc[1] = t_issue - t0; }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) { ptx::tc_fence_after(); ptx::tmem_dealloc(tmem, 512); }
}

template <int M, int N, int NACC, int MODE>
void run(long long* cyc) {
  const int smem = 65536 + 32768 + 64 + 1024;
  cudaFuncSetAttribute(mma_rate<M, N, NACC, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  mma_rate<M, N, NACC, MODE><<<148, 128, smem>>>(cyc);
  cudaDeviceSynchronize();
  mma_rate<M, N, NACC, MODE><<<148, 128, smem>>>(cyc);
  cudaError_t err = cudaDeviceSynchronize();
  long long h[2] = {0, 0};
  cudaMemcpy(h, cyc, 16, cudaMemcpyDeviceToHost);
  printf("M=%3d N=%3d  %s  %d accumulator(s): %6.1f cycles per MMA to completion, %6.1f to issue  [%s]\n", M, N,
         MODE ? "MN-major" : "K-major ", NACC, double(h[0]) / ITERS, double(h[1]) / ITERS, cudaGetErrorString(err));
}


// The backward kernel's per-unit MMA mix issued back to back by one thread: 12 x (N=16) + 12 x (N=256) K-major, 16 x (N=64) MN-major
// (2.87 k cycles if every instruction ran at its isolated rate), optionally with 8 more warps streaming 16-byte shared-memory
// stores (the dense pass writes 64 KB of dS per unit) or with the MN-major A operand read from the region they write.
template <int STORE_WARPS>
__global__ void __launch_bounds__(128 + 32 * STORE_WARPS) unit_mix(long long* cyc, int units, int sts_per_unit) {
  extern __shared__ uint8_t raw[];
  const uint32_t base = (ptx::smem_u32(raw) + 1023u) & ~1023u;
  uint8_t* bp = raw + (base - ptx::smem_u32(raw));
  const uint32_t a = base, b = base + 32768, ds = base + 65536;   // A 32 KB (W hi+lo), B 32 KB (window hi+lo), dS 64 KB
  uint32_t* tptr = reinterpret_cast<uint32_t*>(bp + 131072);
  const uint32_t bar = base + 131072 + 16;
  for (int i = threadIdx.x; i < 131072 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(bp)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) { ptx::mbar_init(bar, 1); ptx::fence_barrier_init(); }
  if (threadIdx.x < 32) { ptx::tmem_alloc(ptx::smem_u32(tptr), 512); ptx::tmem_relinquish(); }
  ptx::fence_proxy_async();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(tptr);
  if (threadIdx.x == 0) {
    constexpr uint32_t id16 = idesc_of(128, 16, 0, 0), id256 = idesc_of(128, 256, 0, 0), id64 = idesc_of(128, 64, 1, 1);
    const long long t0 = clock64();
    for (int u = 0; u < units; ++u) {
#pragma unroll
      for (int i = 0; i < 12; ++i)
        ptx::umma_bf16(tmem + 448, ptx::umma_desc_sw128(a + (i & 3) * 32 + (i & 4) * 4096), ptx::umma_desc_sw128(b + (i & 3) * 32), id16, 1u);
#pragma unroll
      for (int i = 0; i < 12; ++i)
        ptx::umma_bf16(tmem, ptx::umma_desc_sw128(a + (i & 3) * 32 + (i & 4) * 4096), ptx::umma_desc_sw128(b + (i & 3) * 32), id256, 1u);
#pragma unroll
      for (int i = 0; i < 16; ++i)
        ptx::umma_bf16(tmem + 256 + (i >> 3) * 64, mn_desc(ds + (i >> 3) * 32768 + (i & 7) * 2048, 16384, 1024), mn_desc(a + (i & 7) * 2048, 16384, 1024), id64, 1u);
    }
    ptx::umma_commit(bar);
    ptx::mbar_wait(bar, 0u);
    const long long t1 = clock64();
    if (blockIdx.x == 0) cyc[0] = t1 - t0;
  } else if (threadIdx.x >= 128) {
    // store warps: sts_per_unit 16-byte stores per thread per ~unit, paced by clock
    const int t = threadIdx.x - 128;
    uint4 v = make_uint4(t, t, t, t);
    for (int u = 0; u < units; ++u) {
      for (int i = 0; i < sts_per_unit; ++i) {
        const uint32_t off = ((i * 256 + t) * 16) & 65535;
        *reinterpret_cast<uint4*>(bp + 65536 + off) = v;
      }
      const long long t0 = clock64();
      while (clock64() - t0 < 2000) {}
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) { ptx::tc_fence_after(); ptx::tmem_dealloc(tmem, 512); }
}

template <int STORE_WARPS>
void run_mix(long long* cyc, int sts) {
  const int smem = 131072 + 64 + 1024, units = 256;
  cudaFuncSetAttribute(unit_mix<STORE_WARPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  unit_mix<STORE_WARPS><<<148, 128 + 32 * STORE_WARPS, smem>>>(cyc, units, sts);
  cudaDeviceSynchronize();
  unit_mix<STORE_WARPS><<<148, 128 + 32 * STORE_WARPS, smem>>>(cyc, units, sts);
  cudaError_t err = cudaDeviceSynchronize();
  long long h = 0;
  cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  printf("unit mix (12 x N16 + 12 x N256 + 16 x N64 MN-major), %d store warps x %d STS.128 per unit: %7.1f cycles per unit  [%s]\n", STORE_WARPS, sts,
         double(h) / units, cudaGetErrorString(err));
}

int main() {
  long long* cyc;
  cudaMalloc(&cyc, 16);
  run<128, 16, 1, 0>(cyc);
  run<128, 16, 2, 0>(cyc);
  run<128, 16, 4, 0>(cyc);
  run<128, 32, 1, 0>(cyc);
  run<128, 64, 1, 0>(cyc);
  run<128, 64, 2, 0>(cyc);
  run<128, 128, 1, 0>(cyc);
  run<128, 256, 1, 0>(cyc);
  run<128, 64, 1, 1>(cyc);
  run<128, 64, 2, 1>(cyc);
  run<128, 64, 4, 1>(cyc);
  run<128, 128, 1, 1>(cyc);
  run<128, 128, 2, 1>(cyc);
  run<64, 256, 1, 0>(cyc);
  run<64, 128, 1, 0>(cyc);
  run_mix<0>(cyc, 0);
  run_mix<8>(cyc, 16);
  run_mix<8>(cyc, 64);
  return 0;
}
