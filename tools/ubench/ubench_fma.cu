// Microbenchmark: FFMA vs FFMA2 (fma.rn.f32x2) issue throughput per SM on sm_100a.
#include <cstdio>
#include <cuda_runtime.h>
template <int MODE>
__global__ void k(float* out, int iters, float s) {
  float a[16];
  unsigned long long p[8];
#pragma unroll
  for (int i = 0; i < 16; ++i) a[i] = threadIdx.x * 0.001f + i;
#pragma unroll
  for (int i = 0; i < 8; ++i) asm("mov.b64 %0, {%1,%2};" : "=l"(p[i]) : "f"(a[2 * i]), "f"(a[2 * i + 1]));
  unsigned long long ss;
  asm("mov.b64 %0, {%1,%2};" : "=l"(ss) : "f"(s), "f"(s));
  for (int it = 0; it < iters; ++it) {
    if (MODE == 0) {
#pragma unroll
      for (int i = 0; i < 16; ++i) a[i] = fmaf(a[i], s, 1.0f);
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) asm volatile("fma.rn.f32x2 %0, %0, %1, %1;" : "+l"(p[i]) : "l"(ss));
    }
  }
  float r = 0;
  if (MODE == 0) { for (int i = 0; i < 16; ++i) r += a[i]; }
  else { for (int i = 0; i < 8; ++i) { float lo, hi; asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(p[i])); r += lo + hi; } }
  out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}
int main() {
  float* out; cudaMalloc(&out, 148 * 8 * 1024 * 4);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int mode = 0; mode < 2; ++mode) for (int warps : {4, 8, 16, 32}) {
    const int iters = 20000; float ms;
    for (int rep = 0; rep < 2; ++rep) {
      cudaEventRecord(e0);
      if (mode == 0) k<0><<<148 * 2, warps * 16>>>(out, iters, 0.999f); else k<1><<<148 * 2, warps * 16>>>(out, iters, 0.999f);
      cudaEventRecord(e1); cudaEventSynchronize(e1); cudaEventElapsedTime(&ms, e0, e1);
    }
    double fma = 2.0 * 148 * warps * 16 * 16.0 * iters;  // scalar FMAs
    printf("%s warps/SM=%d  %.3f ms  %.1f TFLOP/s (fp32 FMA)\n", mode ? "FFMA2" : "FFMA ", warps, ms, 2 * fma / ms / 1e9);
  }
  return 0;
}
