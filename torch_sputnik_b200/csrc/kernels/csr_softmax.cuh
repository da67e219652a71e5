// csr_softmax.cuh -- row-wise softmax over CSR values (forward and backward), HBM-bound.
//
// Semantics of sputnik::SparseSoftmaxKernel (third_party/sputnik/sputnik/softmax/
// sparse_softmax.cu.cc:24-69): per row, y = exp(x - max) * (1 / sum(exp(x - max))) with expf
// and one reciprocal.  The reference makes three passes over the row from global memory and
// launches once per replica (src/softmax_cuda.cu:35-43); here one warp keeps a row of up to
// 32*REG values in registers so every value is read once and written once, replicas ride on
// grid.y, and the caller-side 1/sqrt(d) (modules/sparse_attention.py:72) can be folded in as
// `scale`.  Rows longer than 32*REG fall back to the three-pass loop (L1/L2 catches re-reads).
#pragma once
#include "common.cuh"

namespace tsp {

// T: storage type of the values (float, __half, __nv_bfloat16); max / exp / sum are fp32.
template <typename T>
struct SoftmaxArgsT {
  int m, nnz;
  float scale;
  const int* row_indices;  // may be null
  const int* row_offsets;
  const T* x;   // forward: values; backward: y
  const T* dy;  // backward only
  T* out;
};
using SoftmaxArgs = SoftmaxArgsT<float>;

template <typename T>
__device__ __forceinline__ float ld_once(const T* p) {
  if constexpr (sizeof(T) == 4) return ld_stream_f(reinterpret_cast<const float*>(p));
  else return Elem<T>::to_f(*p);
}
template <typename T>
__device__ __forceinline__ void st_once(T* p, float v) {
  if constexpr (sizeof(T) == 4) st_stream_f(reinterpret_cast<float*>(p), v);
  else *p = Elem<T>::from_f(v);
}

template <int REG, int WARPS, typename T = float>
__global__ void __launch_bounds__(WARPS * 32) csr_softmax_kernel(const SoftmaxArgsT<T> a) {
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int item = blockIdx.x * WARPS + warp;
  if (item >= a.m) return;
  const int row = a.row_indices ? __ldg(a.row_indices + item) : item;
  const int beg = __ldg(a.row_offsets + row);
  const int len = __ldg(a.row_offsets + row + 1) - beg;
  if (len <= 0) return;
  const size_t base = (size_t)blockIdx.y * a.nnz + beg;
  const T* __restrict__ in = a.x + base;
  T* __restrict__ out = a.out + base;
  const float scale = a.scale;

  if (len <= 32 * REG) {
    float v[REG];
    float mx = -INFINITY;
#pragma unroll
    for (int i = 0; i < REG; ++i) {
      const int idx = lane + 32 * i;
      v[i] = idx < len ? ld_once(in + idx) * scale : -INFINITY;
      mx = fmaxf(mx, v[i]);
    }
    mx = warp_max(mx);
    float sum = 0.f;
#pragma unroll
    for (int i = 0; i < REG; ++i) {
      if (32 * i < len) {        // warp-uniform: register slots past the row do no work
        v[i] = expf(v[i] - mx);  // expf(-inf) == 0 for the padding lanes
        sum += v[i];
      }
    }
    sum = warp_sum(sum);
    const float inv = 1.0f / sum;
#pragma unroll
    for (int i = 0; i < REG; ++i) {
      const int idx = lane + 32 * i;
      if (idx < len) st_once(out + idx, v[i] * inv);
    }
  } else {
    float mx = -INFINITY;
    for (int idx = lane; idx < len; idx += 32) mx = fmaxf(mx, Elem<T>::to_f(in[idx]) * scale);
    mx = warp_max(mx);
    float sum = 0.f;
    for (int idx = lane; idx < len; idx += 32) sum += expf(Elem<T>::to_f(in[idx]) * scale - mx);
    sum = warp_sum(sum);
    const float inv = 1.0f / sum;
    for (int idx = lane; idx < len; idx += 32)
      out[idx] = Elem<T>::from_f(expf(Elem<T>::to_f(in[idx]) * scale - mx) * inv);
  }
}

// dx = scale * y * (dy - sum(dy * y)) per row.
template <int REG, int WARPS>
__global__ void __launch_bounds__(WARPS * 32) csr_softmax_bwd_kernel(const SoftmaxArgs a) {
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int item = blockIdx.x * WARPS + warp;
  if (item >= a.m) return;
  const int row = a.row_indices ? __ldg(a.row_indices + item) : item;
  const int beg = __ldg(a.row_offsets + row);
  const int len = __ldg(a.row_offsets + row + 1) - beg;
  if (len <= 0) return;
  const size_t base = (size_t)blockIdx.y * a.nnz + beg;
  const float* __restrict__ y = a.x + base;
  const float* __restrict__ dy = a.dy + base;
  float* __restrict__ dx = a.out + base;

  if (len <= 32 * REG) {
    float yv[REG], gv[REG];
    float dot = 0.f;
#pragma unroll
    for (int i = 0; i < REG; ++i) {
      const int idx = lane + 32 * i;
      yv[i] = idx < len ? ld_stream_f(y + idx) : 0.f;
      gv[i] = idx < len ? ld_stream_f(dy + idx) : 0.f;
      dot = fmaf(yv[i], gv[i], dot);
    }
    dot = warp_sum(dot);
#pragma unroll
    for (int i = 0; i < REG; ++i) {
      const int idx = lane + 32 * i;
      if (idx < len) st_stream_f(dx + idx, a.scale * yv[i] * (gv[i] - dot));
    }
  } else {
    float dot = 0.f;
    for (int idx = lane; idx < len; idx += 32) dot = fmaf(__ldg(y + idx), __ldg(dy + idx), dot);
    dot = warp_sum(dot);
    for (int idx = lane; idx < len; idx += 32)
      dx[idx] = a.scale * __ldg(y + idx) * (__ldg(dy + idx) - dot);
  }
}

}  // namespace tsp
