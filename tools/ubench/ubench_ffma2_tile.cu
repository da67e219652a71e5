// Microbenchmark: register-only 8x8 (and 8x4) outer-product tiles with FFMA2 / FFMA on sm_100a:
// how close to the 74 TFLOP/s fp32 pipe can the issue + register file get with N warps per SM?
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ void fma2(u64& d, u64 a, u64 b) { asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(d) : "l"(a), "l"(b)); }

template <int I, int J, int MODE>   // MODE 0: FFMA2 (I x J pairs), 1: FFMA (I x J scalars)
__global__ void __launch_bounds__(128) k(float* out, int iters, const u64* src) {
  u64 a[I], b[J], acc[I][J];
#pragma unroll
  for (int i = 0; i < I; ++i) a[i] = src[threadIdx.x + i * 128];
#pragma unroll
  for (int j = 0; j < J; ++j) b[j] = src[threadIdx.x + (I + j) * 128];
#pragma unroll
  for (int i = 0; i < I; ++i)
#pragma unroll
    for (int j = 0; j < J; ++j) acc[i][j] = 0;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int j = 0; j < J; ++j)
#pragma unroll
      for (int i = 0; i < I; ++i) {
        if (MODE == 0) fma2(acc[i][j], a[i], b[j]);
        else {
          float x = __uint_as_float((unsigned)acc[i][j]);
          x = fmaf(__uint_as_float((unsigned)a[i]), __uint_as_float((unsigned)b[j]), x);
          acc[i][j] = __float_as_uint(x);
        }
      }
    // rotate operands so nothing is loop-invariant for the compiler
    u64 t = a[0];
#pragma unroll
    for (int i = 0; i + 1 < I; ++i) a[i] = a[i + 1];
    a[I - 1] = t;
  }
  u64 r = 0;
#pragma unroll
  for (int i = 0; i < I; ++i)
#pragma unroll
    for (int j = 0; j < J; ++j) r ^= acc[i][j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = (float)r;
}

template <int I, int J, int MODE>
void run(const char* name, float* out, const u64* src) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int ctas : {1, 2, 3, 4}) {   // CTAs of 128 threads per SM
    const int iters = 4000; float ms = 0;
    for (int rep = 0; rep < 2; ++rep) {
      cudaEventRecord(e0);
      k<I, J, MODE><<<148 * ctas, 128>>>(out, iters, src);
      cudaEventRecord(e1); cudaEventSynchronize(e1); cudaEventElapsedTime(&ms, e0, e1);
    }
    const double fma = (double)148 * ctas * 128 * iters * I * J * (MODE == 0 ? 2 : 1);
    printf("%s %dx%d  warps/SM=%2d  %.3f ms  %.1f TFLOP/s\n", name, I, J, ctas * 4, ms, 2 * fma / ms / 1e9);
  }
}
int main() {
  float* out; u64* src;
  cudaMalloc(&out, 148 * 4 * 128 * 4); cudaMalloc(&src, 128 * 16 * 8); cudaMemset(src, 0, 128 * 16 * 8);
  run<8, 8, 0>("FFMA2", out, src);
  run<8, 4, 0>("FFMA2", out, src);
  run<6, 8, 0>("FFMA2", out, src);
  run<8, 8, 1>("FFMA ", out, src);
  run<8, 4, 1>("FFMA ", out, src);
  cudaError_t e = cudaDeviceSynchronize();
  printf("%s\n", cudaGetErrorString(e));
  return 0;
}
