// csr_softmax.cuh -- row-wise softmax over CSR values (forward and backward), HBM-bound.
//
// Semantics of sputnik::SparseSoftmaxKernel (third_party/sputnik/sputnik/softmax/
// sparse_softmax.cu.cc:24-69): per row, y = exp(x - max) * (1 / sum(exp(x - max))) with one reciprocal
// (exp through the hardware EX2, see the kernel).  The reference makes three passes over the row from global memory and
// launches once per replica (src/softmax_cuda.cu:35-43); here one warp keeps a row of up to
// 32*REG values in registers so every value is read once and written once -- as aligned 16-byte
// vectors (4 fp32 / 8 fp16 or bf16 values per lane-load) -- replicas ride on grid.y, and the
// caller-side 1/sqrt(d) (modules/sparse_attention.py:72) can be folded in as `scale`.  Rows longer
// than 32*REG fall back to the three-pass loop (L1/L2 catches re-reads).
// Summation order: lane l holds values 4l .. 4l+3 (+128 per vector) instead of l, l+32, ...; the
// lane sums then go through the same 1,2,4,8,16 butterfly as the reference's (sparse_softmax.cu.cc:46-59).
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
  MaskTable masks;   // many masks in one launch (forward kernel)
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

constexpr int kSoftmaxRowsPerWarp = 4;

template <int REG, int WARPS, typename T = float>
__global__ void __launch_bounds__(WARPS * 32) csr_softmax_kernel(const SoftmaxArgsT<T> a) {
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  // Rows are taken in storage order, not through the row swizzle: every row costs the same 8 bytes per
  // value whatever its length, consecutive rows are consecutive in memory (the warps of a CTA stream one
  // contiguous region), and one dependent load (row_indices -> row_offsets -> values) less heads each row.
  // A warp takes kSoftmaxRowsPerWarp consecutive rows, one after another: with one row per warp a CTA lived
  // for ~7000 cycles and the kernel was bound by CTA turnover (65536 CTAs at the attention shape), not by HBM.
  for (int rr = 0; rr < kSoftmaxRowsPerWarp; ++rr) {
  const int row = (blockIdx.x * WARPS + warp) * kSoftmaxRowsPerWarp + rr;
  if (row >= a.m) return;
  const MaskView mv = mask_view(a.masks, blockIdx.y, a.nnz);
  const int* __restrict__ row_offsets = a.row_offsets + mv.topo_rows * (a.m + 1);
  const int beg = __ldg(row_offsets + row);
  const int len = __ldg(row_offsets + row + 1) - beg;
  if (len <= 0) continue;
  const size_t base = (size_t)mv.values + beg;
  const T* __restrict__ in = a.x + base;
  T* __restrict__ out = a.out + base;
  // exp(x) is evaluated as 2^(x log2 e) with the hardware EX2 (the reference uses expf, sparse_softmax.cu.cc:52;
  // the difference is below 2 ulp of the probability): `scale` carries the log2 e, the maximum is in log2 units.
  const float scale = a.scale * 1.4426950408889634f;

  // Vector path: the row is read and written as ALIGNED 16-byte vectors (4 fp32 / 8 16-bit values per lane
  // and load), starting at the 16-byte boundary at or before its first value; the up to V - 1 foreign
  // values in the first and last vector are masked on the way in and never written (their vectors are
  // stored element-wise).  A 16-byte-aligned vector that starts inside an allocation ends inside it.
  constexpr int V = 16 / (int)sizeof(T), NV = REG / V;
  const int mis = (int)((reinterpret_cast<uintptr_t>(in) & 15u) / sizeof(T));
  if (NV > 0 && len + mis <= 32 * REG) {
    const int vlen = len + mis;
    const T* const vin = in - mis;
    T* const vout = out - mis;
    float v[NV > 0 ? NV : 1][V];
    float mx = -INFINITY;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const int e0 = (lane + 32 * j) * V;
#pragma unroll
      for (int t = 0; t < V; ++t) v[j][t] = -INFINITY;
      if (e0 < vlen) {
        unpack_raw<T, V>(v[j], ld_stream_u4(vin + e0));
#pragma unroll
        for (int t = 0; t < V; ++t) {
          v[j][t] = (e0 + t >= mis && e0 + t < vlen) ? v[j][t] * scale : -INFINITY;
          mx = fmaxf(mx, v[j][t]);
        }
      }
    }
    mx = warp_max(mx);
    float sum = 0.f;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      if (32 * j * V < vlen) {   // warp-uniform
#pragma unroll
        for (int t = 0; t < V; ++t) {
          v[j][t] = exp2f(v[j][t] - mx);   // exp2f(-inf) == 0 for masked slots
          sum += v[j][t];
        }
      }
    }
    sum = warp_sum(sum);
    const float inv = 1.0f / sum;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const int e0 = (lane + 32 * j) * V;
      if (e0 < vlen) {
#pragma unroll
        for (int t = 0; t < V; ++t) v[j][t] *= inv;
        if (e0 >= mis && e0 + V <= vlen) {
          if constexpr (sizeof(T) == 4) {
            st_stream_u4(vout + e0, make_uint4(__float_as_uint(v[j][0]), __float_as_uint(v[j][1]),
                                               __float_as_uint(v[j][2]), __float_as_uint(v[j][3])));
          } else {
            st_stream_u4(vout + e0, make_uint4(pack_pair<T>(v[j][0], v[j][1]), pack_pair<T>(v[j][2], v[j][3]),
                                               pack_pair<T>(v[j][4], v[j][5]), pack_pair<T>(v[j][6], v[j][7])));
          }
        } else {
#pragma unroll
          for (int t = 0; t < V; ++t)
            if (e0 + t >= mis && e0 + t < vlen) st_once(vout + e0 + t, v[j][t]);
        }
      }
    }
  } else if (len <= 32 * REG) {
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
        v[i] = exp2f(v[i] - mx);  // exp2f(-inf) == 0 for the padding lanes
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
    for (int idx = lane; idx < len; idx += 32) sum += exp2f(Elem<T>::to_f(in[idx]) * scale - mx);
    sum = warp_sum(sum);
    const float inv = 1.0f / sum;
    for (int idx = lane; idx < len; idx += 32)
      out[idx] = Elem<T>::from_f(exp2f(Elem<T>::to_f(in[idx]) * scale - mx) * inv);
  }
  }   // rows of this warp
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
