// Microbenchmark: what a predicated LDS.128 costs in the shared-memory data pipe when only some (or none) of the
// warp's quarter-warps are active.  Each iteration issues 8 x [one full LDS.128, every quarter reading its own
// 128-byte row = 4 wavefronts] + 8 x [one LDS.128 predicated on (active_mask >> quarter) & 1].
// Prints SM-cycles per (full + predicated) pair; "none" = the loop without the second load.
#include <cstdio>
#include <cuda_runtime.h>

template <bool SECOND>
__global__ void __launch_bounds__(1024) k(unsigned active_mask, int iters, float* out, long long* cycles) {
  extern __shared__ __align__(128) unsigned char sm[];
  for (int i = threadIdx.x; i < 65536 / 4; i += blockDim.x) reinterpret_cast<float*>(sm)[i] = 1.0f;
  __syncthreads();
  const int lane = threadIdx.x & 31, q = lane >> 3, s = lane & 7;
  const unsigned base = (unsigned)__cvta_generic_to_shared(sm) + q * 128 + s * 16;
  const unsigned base2 = base + 32768 + 512;
  const int on = (active_mask >> q) & 1;
  float x = 0, y = 0, z = 0, w = 0, x2 = 0, y2 = 0, z2 = 0, w2 = 0;
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    const unsigned a = base + ((it & 3) << 12);
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      asm volatile("ld.volatile.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(x), "=f"(y), "=f"(z), "=f"(w) : "r"(a + u * 1024));
      if (SECOND)
        asm volatile("{\n.reg .pred p;\nsetp.ne.s32 p, %5, 0;\n@p ld.volatile.shared.v4.f32 {%0,%1,%2,%3}, [%4];\n}"
                     : "+f"(x2), "+f"(y2), "+f"(z2), "+f"(w2) : "r"(base2 + u * 1024), "r"(on));
    }
  }
  const long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = x + y + z + w + x2 + y2 + z2 + w2;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}

int main() {
  float* out; long long* cyc;
  cudaMalloc(&out, 148 * 1024 * 4); cudaMallocManaged(&cyc, 8);
  cudaFuncSetAttribute(k<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  cudaFuncSetAttribute(k<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  const int iters = 4096;
  for (int warps : {8, 16, 32}) {
    k<false><<<148, warps * 32, 65536>>>(0u, iters, out, cyc);
    cudaDeviceSynchronize();
    k<false><<<148, warps * 32, 65536>>>(0u, iters, out, cyc);
    cudaDeviceSynchronize();
    printf("warps %2d  none: %.2f clk per full LDS.128 per SM\n", warps, (double)*cyc / ((double)iters * 8 * warps));
    for (unsigned m : {0u, 1u, 3u, 5u, 15u}) {
      k<true><<<148, warps * 32, 65536>>>(m, iters, out, cyc);
      cudaDeviceSynchronize();
      printf("warps %2d  predicated LDS.128, active quarters mask 0x%x: %.2f clk per pair per SM\n", warps, m,
             (double)*cyc / ((double)iters * 8 * warps));
    }
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
