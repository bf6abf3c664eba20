#include <cstdio>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <cstdint>
template <int MODE>
__global__ void k(float* out, int iters, float seed) {
  float acc[8];
  uint32_t h[8];
  for (int i = 0; i < 8; ++i) { acc[i] = seed + i * 0.01f + threadIdx.x * 1e-3f; h[i] = 0x3c003800u + i + threadIdx.x; }
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (MODE == 0) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(acc[i]));
      if (MODE == 1) asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(h[i]));
      if (MODE == 2) asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(h[i]));
      if (MODE == 3) { uint64_t v; asm volatile("mov.b64 %0, {%1, %2};" : "=l"(v) : "f"(acc[i]), "f"(acc[(i+1)&7]));
                       asm volatile("fma.rn.f32x2 %0, %0, %0, %0;" : "+l"(v)); float lo, hi; asm volatile("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); acc[i] = lo + hi; }
      if (MODE == 4) asm volatile("fma.rn.f32 %0, %0, %0, %0;" : "+f"(acc[i]));
    }
  }
  float s = 0; for (int i = 0; i < 8; ++i) s += acc[i] + __uint_as_float(h[i]);
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int MODE> void run(const char* name, float* d) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  int iters = 4096;
  k<MODE><<<148 * 4, 512>>>(d, 16, 0.5f);
  cudaEventRecord(a);
  k<MODE><<<148 * 4, 512>>>(d, iters, 0.5f);
  cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  double ops = 148.0 * 4 * 512 * (double)iters * 8;
  printf("%s: %.3f ms  %.2f Gop-lanes/s  per SM per clk @1.9GHz: %.2f\n", name, ms, ops / ms / 1e6, ops / (ms * 1e-3) / 148 / 1.9e9);
}
int main() {
  float* d; cudaMalloc(&d, 148 * 4 * 512 * 4);
  run<0>("ex2.f32", d); run<1>("ex2.f16x2 (2 results/op)", d); run<2>("ex2.bf16x2 (2 results/op)", d);
  run<3>("fma.f32x2 (+mov)", d); run<4>("fma.f32", d);
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
}
