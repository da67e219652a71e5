// Microbenchmark: shared-memory cost of LDS.128 / LDS.64 / LDS.32 access patterns on sm_100a.
// One CTA per SM; every warp hammers one pattern; prints SM-cycles per warp-level load instruction.
#include <cstdio>
#include <cuda_runtime.h>

__host__ __device__ inline unsigned pattern_addr(int p, int lane) {
  const int q = lane >> 3, s = lane & 7, h = lane >> 4;   // rows are 128 B
  switch (p) {
    case 0: return 0;                                            // all lanes the same 16 B
    case 1: return h * 128 + h * 16;                             // one address per half warp
    case 2: return q * 128 + q * 16;                             // one address per quarter warp
    case 3: return s * 128 + ((3 ^ s) << 4);                     // 8 distinct rows per quarter (swizzled), same in all quarters
    case 4: return (lane & 15) * 128 + ((3 ^ (lane & 7)) << 4);  // 16 distinct rows, same in both halves
    case 5: return lane * 128 + ((3 ^ (lane & 7)) << 4);         // 32 distinct rows
    case 6: return (q >> 1) * 1024 + s * 128 + ((3 ^ s) << 4);   // quarters pairwise equal
    case 7: return q * 128 + s * 16;                             // each quarter reads one full 128 B row
    case 8: return (lane >> 2) * 128 + (((lane >> 2) & 7) << 4); // 8 distinct addresses, 4 lanes each
    case 9: return lane * 16;                                    // fully contiguous 512 B
    case 10: return (lane & 3) * 128 + ((lane & 3) << 4);          // 4 distinct, strided lanes (lane & 3)
    case 11: return ((lane >> 1) & 7) * 128 + (((lane >> 1) & 7) << 4);  // 8 distinct, adjacent pairs, halves equal
    case 12: return (lane & 1) * 128 + ((lane & 1) << 4);          // 2 distinct, alternating lanes
    case 13: return (lane >> 1) * 128 + (((lane >> 1) & 7) << 4);  // 16 distinct, adjacent pairs
    case 14: return ((lane & 15) >> 2) * 128 + (((lane & 15) >> 2) << 4);  // 4 distinct (groups of 4), halves equal
    case 15: return (lane & 7) * 16;                               // 8 distinct chunks of ONE 128 B row, quarters equal
    case 16: return ((lane >> 2) & 1) * 128 + (((lane >> 2) & 1) << 4);    // 2 distinct, groups of 4 alternate
    case 17: return (lane & 1) * 16 + (lane >> 4) * 128;           // 2 adjacent chunks alternating, per half
    default: return 0;
  }
}

template <int BYTES>
__global__ void __launch_bounds__(1024) k(int p, int iters, float* out, long long* cycles) {
  extern __shared__ __align__(128) unsigned char sm[];
  for (int i = threadIdx.x; i < 32768 / 4; i += blockDim.x) reinterpret_cast<float*>(sm)[i] = 1.0f;
  __syncthreads();
  const unsigned base = (unsigned)__cvta_generic_to_shared(sm) + pattern_addr(p, threadIdx.x & 31);
  float x = 0, y = 0, z = 0, w = 0;
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    const unsigned a = base + ((it & 3) << 12);
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      // asm volatile: the load is issued whether or not its result is consumed
      if (BYTES == 16) asm volatile("ld.volatile.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(x), "=f"(y), "=f"(z), "=f"(w) : "r"(a + u * 2048));
      else if (BYTES == 8) asm volatile("ld.volatile.shared.v2.f32 {%0,%1}, [%2];" : "=f"(x), "=f"(y) : "r"(a + u * 2048));
      else asm volatile("ld.volatile.shared.f32 %0, [%1];" : "=f"(x) : "r"(a + u * 2048));
    }
  }
  const long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = x + y + z + w;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}

int main() {
  float* out; long long* cyc;
  cudaMalloc(&out, 148 * 1024 * 4); cudaMallocManaged(&cyc, 8);
  cudaFuncSetAttribute(k<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  cudaFuncSetAttribute(k<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  cudaFuncSetAttribute(k<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int iters = 4000;
  for (int warps : {32})
    for (int bytes : {16, 8, 4})
      for (int p = 0; p < 18; ++p) {
        float ms = 0;
        for (int rep = 0; rep < 2; ++rep) {
          *cyc = 0;
          cudaEventRecord(e0);
          if (bytes == 16) k<16><<<148, warps * 32, 65536>>>(p, iters, out, cyc);
          else if (bytes == 8) k<8><<<148, warps * 32, 65536>>>(p, iters, out, cyc);
          else k<4><<<148, warps * 32, 65536>>>(p, iters, out, cyc);
          cudaEventRecord(e1);
          cudaError_t e = cudaDeviceSynchronize();
          if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
          cudaEventElapsedTime(&ms, e0, e1);
        }
        const double n = (double)iters * 8 * warps;
        printf("warps %2d LDS.%-3d pattern %d: %.2f cyc/instr (clock64)  %.2f cyc/instr (events @1.965GHz)\n", warps,
               bytes * 8, p, (double)*cyc / n, ms * 1e-3 * 1.965e9 / n);
      }
  return 0;
}
