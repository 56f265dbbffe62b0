// micro-benchmark: issue rate of FFMA vs FFMA2 (packed fp32x2) on sm_100a, alone and mixed with ALU ops
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c){ uint64_t d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
template <int MODE>
__global__ void __launch_bounds__(512) k(float* out, int iters, float a, float b) {
  float x[8]; uint64_t y[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) { x[i] = threadIdx.x * 0.001f + i; y[i] = (uint64_t)__float_as_uint(x[i]) | ((uint64_t)__float_as_uint(x[i] + 1.f) << 32); }
  const uint64_t a2 = (uint64_t)__float_as_uint(a) | ((uint64_t)__float_as_uint(a) << 32);
  const uint64_t b2 = (uint64_t)__float_as_uint(b) | ((uint64_t)__float_as_uint(b) << 32);
  uint32_t z = threadIdx.x;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (MODE == 0) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(a), "f"(b));
      if (MODE == 1) y[i] = fma2(y[i], a2, b2);
      if (MODE == 2) { asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(a), "f"(b)); asm volatile("lop3.b32 %0, %0, %1, 0x55, 0x96;" : "+r"(z) : "r"(it)); }
      if (MODE == 3) { y[i] = fma2(y[i], a2, b2); asm volatile("lop3.b32 %0, %0, %1, 0x55, 0x96;" : "+r"(z) : "r"(it)); }
    }
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += x[i] + __uint_as_float((uint32_t)y[i]) + __uint_as_float((uint32_t)(y[i] >> 32));
  out[blockIdx.x * blockDim.x + threadIdx.x] = s + z;
}
template <int MODE> void run(const char* name, float* out) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int iters = 20000, grid = 148 * 4;
  k<MODE><<<grid, 512>>>(out, 100, 1.0001f, 0.5f);
  cudaEventRecord(e0);
  k<MODE><<<grid, 512>>>(out, iters, 1.0001f, 0.5f);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  const double winstr = (double)grid * 16 * iters * 8;  // warp-level FMA-type instructions
  printf("%-22s %8.3f ms  %7.2f warp-instr/clk/SM (fma-type, at 1.9 GHz)\n", name, ms, winstr / (ms * 1e-3) / 148 / 1.9e9);
}
int main() {
  float* out; cudaMalloc(&out, 148 * 4 * 512 * 4);
  run<0>("FFMA", out); run<1>("FFMA2", out); run<2>("FFMA + LOP3", out); run<3>("FFMA2 + LOP3", out);
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
