// Probe (GPU box only): validates the MN-major shared-memory descriptors the backward kernel relies on.
//   case 1 (dK-like):  D[t][d] = sum_p dS[p][t] * Q[p][d]   A = dS stored [p][t] (MN-major A), B = Q stored [p][d] (MN-major B)
//   case 2 (dQ-like):  D[p][d] = sum_t dS[p][t] * K[t][d]   A = dS stored [p][t] (K-major A),  B = K stored [t][d] (MN-major B)
//   case 3 (dQ, TS):   same product with A = dS read from TENSOR MEMORY (bf16 pairs packed per 32-bit column, lane = pixel),
//                      B as in case 2; also times 16 SS-mode vs 16 TS-mode M=128 N=64 K=16 MMAs.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -std=c++17 -I stablekeypoints_b200/csrc scripts/umma_probe.cu -o gpurun_out/umma_probe
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cmath>
#include <cuda_bf16.h>
#include "ptx_sm100.cuh"
using namespace skp;

__device__ __forceinline__ uint64_t desc_sw128(uint32_t addr, uint32_t lbo, uint32_t sbo) {
  uint64_t d = 0;
  d |= (uint64_t)((addr & 0x3FFFFu) >> 4);
  d |= (uint64_t)(lbo >> 4) << 16;
  d |= (uint64_t)(sbo >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
__device__ __host__ constexpr uint32_t idesc(int M, int N, int a_mn, int b_mn) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
// rows x cols bf16 row-major (cols multiple of 64) -> slabs of [rows][64] with 128B swizzle
__device__ void fill_sw128(uint8_t* dst, const __nv_bfloat16* src, int rows, int cols) {
  for (int e = threadIdx.x; e < rows * cols / 8; e += blockDim.x) {
    int r = e / (cols / 8), c8 = e % (cols / 8);
    int slab = c8 / 8, c = c8 % 8;
    uint4 v = *reinterpret_cast<const uint4*>(src + (size_t)r * cols + c8 * 8);
    *reinterpret_cast<uint4*>(dst + (size_t)slab * rows * 128 + r * 128 + ((c ^ (r & 7)) * 16)) = v;
  }
}
// S: [128 p][256 t], Q: [128 p][64 d], Kt: [256 t][64 d]; out1: [256 t][64 d], out2: [128 p][64 d]
__device__ __forceinline__ void umma_bf16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__global__ void __launch_bounds__(128, 1) probe(const __nv_bfloat16* S, const __nv_bfloat16* Q, const __nv_bfloat16* Kt, float* out1, float* out2,
                                                 float* out3, long long* cyc) {
  extern __shared__ uint8_t raw[];
  uint32_t base = (ptx::smem_u32(raw) + 1023u) & ~1023u;
  uint8_t* bp = raw + (base - ptx::smem_u32(raw));
  uint8_t* sS = bp;                 // 4 slabs x 16 KB
  uint8_t* sQ = bp + 65536;         // 16 KB
  uint8_t* sK = bp + 65536 + 16384; // 256 rows x 128 B = 32 KB
  uint32_t* tptr = reinterpret_cast<uint32_t*>(bp + 65536 + 16384 + 32768);
  uint32_t bar = base + 65536 + 16384 + 32768 + 16;
  fill_sw128(sS, S, 128, 256);
  fill_sw128(sQ, Q, 128, 64);
  fill_sw128(sK, Kt, 256, 64);
  if (threadIdx.x == 0) { ptx::mbar_init(bar, 1); ptx::fence_barrier_init(); }
  if (threadIdx.x < 32) { ptx::tmem_alloc(ptx::smem_u32(tptr), 512); ptx::tmem_relinquish(); }
  ptx::fence_proxy_async();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  uint32_t tm = *tptr;
  const uint32_t aS = base, aQ = base + 65536, aK = base + 65536 + 16384;
  {   // case 3 operand: row p of dS as 128 packed columns at TMEM columns 256..383 (tokens 2c, 2c+1 in column c)
    const int row = threadIdx.x;
    const uint32_t tl = tm + ((uint32_t)((threadIdx.x >> 5) * 32) << 16) + 256;
    for (int c = 0; c < 4; ++c) {
      uint32_t pk[32];
      for (int j = 0; j < 32; ++j) {
        const __nv_bfloat162 h2 = __halves2bfloat162(S[(size_t)row * 256 + 64 * c + 2 * j], S[(size_t)row * 256 + 64 * c + 2 * j + 1]);
        pk[j] = *reinterpret_cast<const uint32_t*>(&h2);
      }
      ptx::tmem_st32(tl + 32 * c, pk);
    }
    ptx::tmem_st_wait();
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
  }
  if (threadIdx.x == 0) {
    // case 1: two M halves (t 0..127 -> cols 0..63, t 128..255 -> cols 64..127), K = 128 p in 8 steps of 16
    for (int mh = 0; mh < 2; ++mh)
      for (int k = 0; k < 8; ++k) {
        uint64_t a = desc_sw128(aS + mh * 2 * 16384 + k * 2048, 16384, 1024);
        uint64_t b = desc_sw128(aQ + k * 2048, 16384, 1024);
        ptx::umma_bf16(tm + mh * 64, a, b, idesc(128, 64, 1, 1), k != 0);
      }
    // case 2: M = 128 p, N = 64 d, K = 256 t in 16 steps
    for (int k = 0; k < 16; ++k) {
      uint64_t a = desc_sw128(aS + (k / 4) * 16384 + (k % 4) * 32, 16, 1024);
      uint64_t b = desc_sw128(aK + k * 2048, 16384, 1024);
      ptx::umma_bf16(tm + 128, a, b, idesc(128, 64, 0, 1), k != 0);
    }
    ptx::umma_commit(bar);
    ptx::mbar_wait(bar, 0);
    ptx::tc_fence_after();
    // timing: 16 SS-mode MMAs (case 2 again) vs 16 TS-mode MMAs (case 3), issue + completion
    long long t0 = clock64();
    for (int k = 0; k < 16; ++k) {
      uint64_t a = desc_sw128(aS + (k / 4) * 16384 + (k % 4) * 32, 16, 1024);
      uint64_t b = desc_sw128(aK + k * 2048, 16384, 1024);
      ptx::umma_bf16(tm + 128, a, b, idesc(128, 64, 0, 1), k != 0);
    }
    ptx::umma_commit(bar);
    ptx::mbar_wait(bar, 1);
    ptx::tc_fence_after();
    long long t1 = clock64();
    for (int k = 0; k < 16; ++k) {
      uint64_t b = desc_sw128(aK + k * 2048, 16384, 1024);
      umma_bf16_ts(tm + 192, tm + 256 + 8 * k, b, idesc(128, 64, 0, 1), k != 0);
    }
    ptx::umma_commit(bar);
    ptx::mbar_wait(bar, 0);
    ptx::tc_fence_after();
    long long t2 = clock64();
    cyc[0] = t1 - t0;
    cyc[1] = t2 - t1;
  }
  __syncthreads();
  ptx::tc_fence_after();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, row = warp * 32 + lane;
  uint32_t v[32];
  const uint32_t tl = tm + ((uint32_t)(warp * 32) << 16);
  for (int c = 0; c < 8; ++c) {
    ptx::tmem_ld32(tl + 32 * c, v);
    ptx::tmem_ld_wait();
    for (int j = 0; j < 32; ++j) {
      int col = 32 * c + j;
      if (col < 64) out1[(size_t)row * 64 + col] = __uint_as_float(v[j]);
      else if (col < 128) out1[(size_t)(128 + row) * 64 + col - 64] = __uint_as_float(v[j]);
      else if (col < 192) out2[(size_t)row * 64 + col - 128] = __uint_as_float(v[j]);
      else out3[(size_t)row * 64 + col - 192] = __uint_as_float(v[j]);
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) ptx::tmem_dealloc(tm, 512);
}

int main() {
  const int P = 128, T = 256, D = 64;
  std::vector<__nv_bfloat16> S(P * T), Q(P * D), K(T * D);
  std::vector<float> Sf(P * T), Qf(P * D), Kf(T * D);
  srand(1);
  auto rnd = [] { return (rand() % 2001 - 1000) / 1000.0f; };
  for (int i = 0; i < P * T; ++i) { S[i] = __float2bfloat16(rnd()); Sf[i] = __bfloat162float(S[i]); }
  for (int i = 0; i < P * D; ++i) { Q[i] = __float2bfloat16(rnd()); Qf[i] = __bfloat162float(Q[i]); }
  for (int i = 0; i < T * D; ++i) { K[i] = __float2bfloat16(rnd()); Kf[i] = __bfloat162float(K[i]); }
  __nv_bfloat16 *dS, *dQ, *dK; float *o1, *o2, *o3; long long* dcyc;
  cudaMalloc(&o3, P * D * 4); cudaMalloc(&dcyc, 16);
  cudaMalloc(&dS, S.size() * 2); cudaMalloc(&dQ, Q.size() * 2); cudaMalloc(&dK, K.size() * 2);
  cudaMalloc(&o1, T * D * 4); cudaMalloc(&o2, P * D * 4);
  cudaMemcpy(dS, S.data(), S.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dQ, Q.data(), Q.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dK, K.data(), K.size() * 2, cudaMemcpyHostToDevice);
  const int smem = 65536 + 16384 + 32768 + 64 + 1024;
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  probe<<<1, 128, smem>>>(dS, dQ, dK, o1, o2, o3, dcyc);
  cudaError_t e = cudaDeviceSynchronize();
  printf("kernel: %s\n", cudaGetErrorString(e));
  std::vector<float> h1(T * D), h2(P * D), h3(P * D);
  long long hc[2];
  cudaMemcpy(h3.data(), o3, h3.size() * 4, cudaMemcpyDeviceToHost);
  cudaMemcpy(hc, dcyc, 16, cudaMemcpyDeviceToHost);
  cudaMemcpy(h1.data(), o1, h1.size() * 4, cudaMemcpyDeviceToHost);
  cudaMemcpy(h2.data(), o2, h2.size() * 4, cudaMemcpyDeviceToHost);
  double e1 = 0, e2 = 0, e3 = 0;
  for (int t = 0; t < T; ++t) for (int d = 0; d < D; ++d) {
    double r = 0; for (int p = 0; p < P; ++p) r += (double)Sf[p * T + t] * Qf[p * D + d];
    e1 = fmax(e1, fabs(r - h1[t * D + d]));
  }
  for (int p = 0; p < P; ++p) for (int d = 0; d < D; ++d) {
    double r = 0; for (int t = 0; t < T; ++t) r += (double)Sf[p * T + t] * Kf[t * D + d];
    e2 = fmax(e2, fabs(r - h2[p * D + d]));
    e3 = fmax(e3, fabs(r - h3[p * D + d]));
  }
  printf("case3 (A from tensor memory, MN-major B) max err %.3e\n", e3);
  printf("16 MMAs M=128 N=64 K=16: SS mode %lld cycles, TS mode %lld cycles (issue to completion)\n", hc[0], hc[1]);
  printf("case1 (MN-major A, MN-major B) max err %.3e\ncase2 (K-major A, MN-major B) max err %.3e\n", e1, e2);
  printf("%s\n", (e1 < 1e-3 && e2 < 1e-3 && e3 < 1e-3) ? "PROBE OK" : "PROBE FAIL");
  return 0;
}
