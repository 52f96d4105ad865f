// Probe (GPU box only): issue rate per SM sub-partition of the instructions the softmax warps are made of
// (3-register FFMA, packed FFMA2 / FADD2, FADD, FMNMX3, MUFU.EX2), 4 warps per scheduler, 8 independent chains.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -std=c++17 -I stablekeypoints_b200/csrc scripts/pipe_probe.cu -o build/pipe_probe
#include <cstdio>
#include <cstdint>
#include "ptx_sm100.cuh"
using namespace skp;

constexpr int ITERS = 2048;

template <int MODE>
__global__ void __launch_bounds__(512, 1) probe(float a, float b, long long* cycles, float* out) {
  float x[8];
  uint64_t y[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    x[j] = threadIdx.x * 0.001f + j;
    y[j] = ptx::pack_f32x2(x[j], x[j] + 0.5f);
  }
  const uint64_t a2 = ptx::pack_f32x2(a, a), b2 = ptx::pack_f32x2(b, b);
  __syncthreads();
  const long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (MODE == 0) x[j] = fmaf(x[j], a, b);                       // FFMA, 3 register sources
      if (MODE == 1) y[j] = ptx::fma_f32x2(y[j], a2, b2);           // FFMA2
      if (MODE == 2) x[j] = x[j] + a;                               // FADD
      if (MODE == 3) y[j] = ptx::add_f32x2(y[j], a2);               // FADD2
      if (MODE == 4) x[j] = ptx::max3(x[j], a, b);                  // FMNMX3
      if (MODE == 5) x[j] = ptx::ex2(x[j]);                         // MUFU.EX2
      if (MODE == 6) x[j] = fmaxf(x[j], a);                         // FMNMX
      if (MODE == 7) x[j] = fmaf(x[j], 1.0001f, b);                 // FFMA with an immediate
    }
  }
  const long long t1 = clock64();
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    float lo, hi;
    ptx::unpack_f32x2(y[j], lo, hi);
    s += x[j] + lo + hi;
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) cycles[0] = t1 - t0;
}

template <int MODE>
void run(const char* name, int elems, long long* cyc, float* out) {
  probe<MODE><<<148, 512>>>(0.999f, 0.001f, cyc, out);
  cudaDeviceSynchronize();
  long long h;
  cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  // 4 warps per scheduler x 8 instructions x ITERS
  const double per_instr = double(h) / (4.0 * 8 * ITERS);
  printf("%-28s %8lld cycles  %.2f cycles / warp-instruction / scheduler  (%.2f per fp32 element)\n", name, h, per_instr, per_instr / elems);
}

int main() {
  long long* cyc;
  float* out;
  cudaMalloc(&cyc, 8);
  cudaMalloc(&out, 148 * 512 * 4);
  run<0>("FFMA (3 registers)", 1, cyc, out);
  run<7>("FFMA (immediate)", 1, cyc, out);
  run<1>("FFMA2", 2, cyc, out);
  run<2>("FADD", 1, cyc, out);
  run<3>("FADD2", 2, cyc, out);
  run<4>("FMNMX3", 2, cyc, out);
  run<6>("FMNMX", 1, cyc, out);
  run<5>("MUFU.EX2", 1, cyc, out);
  return 0;
}
