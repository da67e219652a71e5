// Microbenchmark: latency of mbarrier.try_wait / test_wait on a phase that completed long ago (the common case of
// the TMA rings here: the data of the next slab arrived while the previous one was being consumed).
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

template <int MODE>
__global__ void k(int iters, long long* cycles, unsigned* sink) {
  __shared__ unsigned long long bar;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&bar)), "r"(1) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  unsigned acc = 0;
  const long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
    unsigned ok;
    if (MODE == 0)
      asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                   : "=r"(ok) : "r"(smem_u32(&bar)), "r"(1) : "memory");
    else
      asm volatile("{\n.reg .pred p;\nmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                   : "=r"(ok) : "r"(smem_u32(&bar)), "r"(1) : "memory");
    acc += ok;
  }
  const long long t1 = clock64();
  if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
  sink[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

int main() {
  long long* cyc; unsigned* sink;
  cudaMallocManaged(&cyc, 8); cudaMalloc(&sink, 148 * 1024 * 4);
  const int iters = 10000;
  for (int warps : {1, 4, 8}) {
    k<0><<<148, warps * 32>>>(iters, cyc, sink); cudaDeviceSynchronize();
    k<0><<<148, warps * 32>>>(iters, cyc, sink); cudaDeviceSynchronize();
    printf("warps %d  try_wait  (completed phase): %.1f clk per dependent call\n", warps, (double)*cyc / iters);
    k<1><<<148, warps * 32>>>(iters, cyc, sink); cudaDeviceSynchronize();
    printf("warps %d  test_wait (completed phase): %.1f clk per dependent call\n", warps, (double)*cyc / iters);
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
