// sddmm_gather.cuh -- general CSR-sampled dense-dense matmul (variant 1, "gather").
//
// out[j] = <lhs[row_of(j), :], rhs[column_indices[j], :]>  for every nonzero j of the mask.
//
// Work decomposition (new, not the reference's): the FLAT nonzero array is cut into chunks of
// 32 consecutive nonzeros, one chunk per warp, so the load is balanced by construction with no
// empty blocks and no dependence on the row swizzle (the reference launches (m/4) x (n/32)
// blocks and lets the strips past a row's length exit, third_party/sputnik/sputnik/sddmm/
// cuda_sddmm.cu.cc:69,207-209).  Inside a warp LPN lanes (8, 16 or 32, by k) share one
// nonzero, each lane owning a VEC-wide slice of the inner dimension; the per-lane partial
// sums of the chunk are combined with a transposing butterfly (NV values x LPN lanes in
// NV-1 shuffles) instead of one 5-step all-reduce per nonzero (all_reduce.h:61-80).
// Arithmetic matches sputnik::CudaSddmmKernel (cuda_sddmm.cu.cc:34-165): fp32 FMA along k per
// lane, then fp32 tree sum across lanes.
#pragma once
#include "spmm_gather.cuh"  // StagedGuard

namespace tsp {

// T: storage type of lhs / rhs / out (float, __half, __nv_bfloat16); arithmetic is fp32.
template <typename T>
struct SddmmArgsT {
  int m, k, n, nnz;
  const int* row_offsets;
  const int* column_indices;
  const T* lhs;  // [replicas, m, k]
  const T* rhs;  // [replicas, n, k]
  T* out;        // [replicas, nnz]
  // guard_flag != null: this launch is the unsorted-columns fallback of a staged launch and runs
  // only if *guard_flag != 0 (see StagedGuard, spmm_gather.cuh).
  const unsigned* guard_flag = nullptr;
  // reduce_reps > 0: out is ONE [nnz] vector holding the sum over that many replicas (the
  // weight gradient of a sparse linear layer); the staged kernels fold the replica loop into
  // their k loop, grid.y is 1.
  int reduce_reps = 0;
  MaskTable masks;   // many masks in one launch (rows / gather kernels only); a.nnz is then the LARGEST mask's count
};
using SddmmArgs = SddmmArgsT<float>;

// Row containing flat nonzero j: largest i with row_offsets[i] <= j (skips empty rows).
__device__ __forceinline__ int row_of_nonzero(const int* __restrict__ row_offsets, int m, int j) {
  int lo = 0, hi = m;  // invariant: row_offsets[lo] <= j < row_offsets[hi]
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (__ldg(row_offsets + mid) <= j) lo = mid; else hi = mid;
  }
  return lo;
}

// Reduce NV per-lane partials across the LPN lanes of a group; on return lane l (within its
// group) holds in vals[0] the total of value index l.  NV == LPN.
template <int LPN>
__device__ __forceinline__ void transpose_reduce(float (&vals)[LPN], int lane) {
#pragma unroll
  for (int s = LPN / 2; s >= 1; s >>= 1) {
    const bool upper = (lane & s) != 0;
#pragma unroll
    for (int i = 0; i < s; ++i) {
      const float keep = upper ? vals[i + s] : vals[i];
      const float send = upper ? vals[i] : vals[i + s];
      vals[i] = keep + __shfl_xor_sync(kFullMask, send, s);
    }
  }
}

template <int VEC, int LPN, int WARPS, typename T = float>
__global__ void __launch_bounds__(WARPS * 32) sddmm_gather_kernel(const SddmmArgsT<T> a) {
  constexpr int G = 32 / LPN;  // nonzeros processed side by side in a warp
  constexpr int NPG = LPN;     // nonzeros per group per chunk (G * NPG == 32)
  constexpr int KSTEP = LPN * VEC;

  if (fallback_not_needed(a.guard_flag)) return;
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int chunk = blockIdx.x * WARPS + warp;
  const int j0 = chunk * 32;
  const int rep = blockIdx.y;
  const MaskView mv = mask_view(a.masks, rep, a.nnz);
  const int nnz = mv.nnz;
  if (j0 >= nnz) return;
  const int* __restrict__ row_offsets = a.row_offsets + mv.topo_rows * (a.m + 1);
  const int* __restrict__ column_indices = a.column_indices + mv.cols;
  const int k = a.k;
  // reduce_reps > 0: one output vector, the replica loop runs here (grid.y == 1)
  const int nrep = a.reduce_reps > 0 ? a.reduce_reps : 1;

  const int grp = lane / LPN;
  const int sub = lane % LPN;

  // Lane t of the warp resolves (row, column) of nonzero j0 + t once.
  int my_row = 0, my_col = 0;
  {
    const int j = j0 + lane;
    if (j < nnz) {
      my_row = row_of_nonzero(row_offsets, a.m, j);
      my_col = __ldg(column_indices + j);
    }
  }

  float acc[NPG];
#pragma unroll
  for (int i = 0; i < NPG; ++i) acc[i] = 0.f;

  // Group g owns nonzeros j0 + g*NPG + i, i in [0, NPG).
  for (int rr = 0; rr < nrep; ++rr) {
  const T* __restrict__ lhs = a.lhs + (size_t)(rep + rr) * a.m * k;
  const T* __restrict__ rhs = a.rhs + (size_t)(rep + rr) * a.n * k;
  for (int kk = 0; kk < k; kk += KSTEP) {
    const int kofs = kk + sub * VEC;
    const bool kin = kofs < k;  // VEC == 4 implies k % 4 == 0, so a vector is all-in or all-out
    int prev_row = -1;
    float av[VEC];
#pragma unroll
    for (int i = 0; i < NPG; ++i) {
      const int src = grp * NPG + i;
      const int r = __shfl_sync(kFullMask, my_row, src);
      const int c = __shfl_sync(kFullMask, my_col, src);
      if (kin && (j0 + src) < nnz) {
        if (r != prev_row) {
          load_elems<T, VEC>(av, lhs + (size_t)r * k + kofs);
          prev_row = r;
        }
        float b[VEC];
        load_elems<T, VEC>(b, rhs + (size_t)c * k + kofs);
#pragma unroll
        for (int v = 0; v < VEC; ++v) acc[i] = fmaf(av[v], b[v], acc[i]);
      }
    }
  }
  }

  transpose_reduce<LPN>(acc, lane);
  const int j = j0 + grp * NPG + sub;
  if (j < nnz) a.out[mv.values + j] = Elem<T>::from_f(acc[0]);
}

// out[j] = sum_r scratch[r, j] (ascending r).
__global__ void __launch_bounds__(256)
reduce_replicas_kernel(const float* __restrict__ scratch, int nnz, int replicas, float* __restrict__ out) {
  const int j = blockIdx.x * 256 + threadIdx.x;
  if (j >= nnz) return;
  float s = 0.f;
  for (int r = 0; r < replicas; ++r) s += ld_stream_f(scratch + (size_t)r * nnz + j);
  out[j] = s;
}

template <int VEC, int LPN, typename T = float>
inline void launch_sddmm_gather(const SddmmArgsT<T>& a, int replicas, cudaStream_t stream) {
  constexpr int WARPS = 4;
  dim3 grid(ceil_div(ceil_div(a.nnz, 32), WARPS), a.reduce_reps > 0 ? 1 : replicas, 1);
  sddmm_gather_kernel<VEC, LPN, WARPS, T><<<grid, WARPS * 32, 0, stream>>>(a);
}

}  // namespace tsp
