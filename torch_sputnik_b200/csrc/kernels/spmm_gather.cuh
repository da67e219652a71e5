// spmm_gather.cuh -- general CSR SpMM (variant 1, "gather").
//
// One warp owns one sparse row x one tile of 32*VEC*U output columns and walks the row's
// nonzeros in order, gathering the matching row segment of the dense operand straight from
// global memory (L1/L2).  It makes no assumption about the order of column indices inside a
// row, the alignment of n, or duplicates, so it is the path for every input the staged kernel
// (spmm_staged.cuh) does not accept, and for small problems where staging cannot pay.
//
// Computes what sputnik::SpmmKernel computes (third_party/sputnik/sputnik/spmm/cuda_spmm.cu.cc:
// 106-283): sequential fp32 FMA accumulation over the row's nonzeros per output element, then
// the optional bias[row] + ReLU epilogue (:267-277).  The decomposition is new: replicas ride
// on grid.z (the reference loops on the host, src/spmm_cuda.cu:48-57), (value, column) pairs
// are staged per warp in shared memory as 16-byte pairs-of-pairs, and address math is 64-bit
// (the reference's 32-bit byte offsets overflow at k*n*4 >= 2^31, sparse_tile.h:93,122).
#pragma once
#include "common.cuh"

namespace tsp {

// T: storage type of values / dense / out (float, __half, __nv_bfloat16); arithmetic is fp32.
template <typename T>
struct SpmmArgsT {
  int m, k, n, nnz;
  long long values_stride;  // nnz, or 0 when one values vector is shared by all replicas
  const int* row_indices;   // may be null (identity)
  const T* values;
  const int* row_offsets;
  const int* column_indices;
  const T* dense;
  const float* bias;  // may be null; always fp32 (as sputnik::CudaSpmmBiasRelu, cuda_spmm.h:64-72,115-127)
  int relu;
  T* out;
  // value_perm != null: nonzero p takes values[value_perm[p]] -- the CSR arrays describe the
  // TRANSPOSE of the matrix the values are stored for (perm from tsp_csr_transpose_f32), so a
  // backward pass multiplies by A^T without materialising the permuted [replicas, nnz] values.
  const int* value_perm = nullptr;
  // guard_flag != null: this launch is the unsorted-columns fallback of a staged launch and runs
  // only if the sorted-columns check raised *guard_flag (a device word owned by the call's plan).
  const unsigned* guard_flag = nullptr;
  MaskTable masks;   // many masks in one launch (gather kernel only)
};
using SpmmArgs = SpmmArgsT<float>;

// "Columns are not sorted" verdict of ONE topology: a device word owned by the topology's plan
// (plan.cuh) -- persistent for a planned topology, stream-ordered scratch for an unplanned call.
// flag == null: the host already knows the columns are sorted; no device check, no fallback.
struct StagedGuard {
  const unsigned* flag;
};
__device__ __forceinline__ bool staged_declines(const StagedGuard& g) {
  return g.flag != nullptr && *reinterpret_cast<const volatile unsigned*>(g.flag) != 0u;
}
__device__ __forceinline__ bool fallback_not_needed(const unsigned* guard_flag) {
  return guard_flag != nullptr && *reinterpret_cast<const volatile unsigned*>(guard_flag) == 0u;
}

template <int VEC>
struct VecT;
template <>
struct VecT<1> {
  using type = float;
};
template <>
struct VecT<2> {
  using type = float2;
};
template <>
struct VecT<4> {
  using type = float4;
};

template <int VEC>
__device__ __forceinline__ void load_vec(float (&dst)[VEC], const float* p) {
  if constexpr (VEC == 4) {
    float4 v = __ldg(reinterpret_cast<const float4*>(p));
    dst[0] = v.x; dst[1] = v.y; dst[2] = v.z; dst[3] = v.w;
  } else if constexpr (VEC == 2) {
    float2 v = __ldg(reinterpret_cast<const float2*>(p));
    dst[0] = v.x; dst[1] = v.y;
  } else {
    dst[0] = __ldg(p);
  }
}

template <int VEC>
__device__ __forceinline__ void store_vec(float* p, const float (&src)[VEC]) {
  if constexpr (VEC == 4) {
    *reinterpret_cast<float4*>(p) = make_float4(src[0], src[1], src[2], src[3]);
  } else if constexpr (VEC == 2) {
    *reinterpret_cast<float2*>(p) = make_float2(src[0], src[1]);
  } else {
    *p = src[0];
  }
}

// OFF32: the staged pair carries the operand row's offset in VECTOR units (col * n / VEC; n % VEC
// == 0 on the vector paths, fits 32 bits when k * n * 4 < 2^31, checked on the host) instead of
// the column, so a gather address is ONE wide multiply-add onto the lane's own 64-bit base
// (IMAD.WIDE.U32) instead of a 64-bit multiply + adds per load -- the gather loop was
// issue-bound on that arithmetic (profiles/r01_spmm_gather_c4_ncu_summary.txt).
template <int VEC, int U, int WARPS, bool OFF32>
__global__ void __launch_bounds__(WARPS * 32) spmm_gather_kernel(const SpmmArgs a) {
  __shared__ __align__(16) int2 s_pairs[WARPS][32];
  if (fallback_not_needed(a.guard_flag)) return;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int item = blockIdx.x * WARPS + warp;
  if (item >= a.m) return;
  const int rep = blockIdx.z;
  const MaskView mv = mask_view(a.masks, rep, a.nnz);
  const int row = a.row_indices ? __ldg(a.row_indices + mv.topo_rows * a.m + item) : item;

  const int n = a.n;
  const int n0 = blockIdx.y * (32 * VEC * U) + lane * VEC;
  const float* __restrict__ dense = a.dense + (size_t)rep * a.k * n;
  using V = typename VecT<VEC>::type;
  const V* lane_base[U];   // this lane's slice of operand row 0, per column tile
#pragma unroll
  for (int u = 0; u < U; ++u) lane_base[u] = reinterpret_cast<const V*>(dense + n0 + u * 32 * VEC);
  const unsigned row_vecs = (unsigned)n / VEC;
  const float* __restrict__ vals = a.masks.prefix ? a.values + mv.values : a.values + (size_t)rep * a.values_stride;
  const int* __restrict__ cols = a.column_indices + mv.cols;

  bool ok[U];
#pragma unroll
  for (int u = 0; u < U; ++u) ok[u] = (n0 + u * 32 * VEC) < n;

  float acc[U][VEC];
#pragma unroll
  for (int u = 0; u < U; ++u)
#pragma unroll
    for (int v = 0; v < VEC; ++v) acc[u][v] = 0.f;

  const int* __restrict__ row_offsets = a.row_offsets + mv.topo_rows * (a.m + 1);
  const int beg = __ldg(row_offsets + row);
  const int end = __ldg(row_offsets + row + 1);

  int2* sp = s_pairs[warp];
  for (int base = beg; base < end; base += 32) {
    const int cnt = min(32, end - base);
    int2 p = make_int2(0, 0);
    if (lane < cnt) {
      p.x = __float_as_int(__ldg(vals + (a.value_perm ? __ldg(a.value_perm + base + lane) : base + lane)));
      p.y = __ldg(cols + base + lane);
      if (OFF32) p.y = (int)((unsigned)p.y * row_vecs);
    }
    __syncwarp();
    sp[lane] = p;
    __syncwarp();

    int j = 0;
    for (; j + 4 <= cnt; j += 4) {
      const int4 q0 = *reinterpret_cast<const int4*>(sp + j);
      const int4 q1 = *reinterpret_cast<const int4*>(sp + j + 2);
      const float av[4] = {__int_as_float(q0.x), __int_as_float(q0.z), __int_as_float(q1.x),
                           __int_as_float(q1.z)};
      const int ac[4] = {q0.y, q0.w, q1.y, q1.w};
      float b[4][U][VEC];
#pragma unroll
      for (int t = 0; t < 4; ++t)
#pragma unroll
        for (int u = 0; u < U; ++u)
          if (ok[u]) {
            if (OFF32)
              load_vec<VEC>(b[t][u], reinterpret_cast<const float*>(lane_base[u] + (unsigned)ac[t]));
            else
              load_vec<VEC>(b[t][u], dense + (size_t)ac[t] * n + n0 + u * 32 * VEC);
          }
#pragma unroll
      for (int t = 0; t < 4; ++t)
#pragma unroll
        for (int u = 0; u < U; ++u)
          if (ok[u])
#pragma unroll
            for (int v = 0; v < VEC; ++v) acc[u][v] = fmaf(av[t], b[t][u][v], acc[u][v]);
    }
    for (; j < cnt; ++j) {
      const int2 q = sp[j];
      const float av = __int_as_float(q.x);
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (ok[u]) {
          float b[VEC];
          if (OFF32)
            load_vec<VEC>(b, reinterpret_cast<const float*>(lane_base[u] + (unsigned)q.y));
          else
            load_vec<VEC>(b, dense + (size_t)q.y * n + n0 + u * 32 * VEC);
#pragma unroll
          for (int v = 0; v < VEC; ++v) acc[u][v] = fmaf(av, b[v], acc[u][v]);
        }
    }
  }

  const float bias = a.bias ? __ldg(a.bias + row) : 0.f;
  float* __restrict__ out = a.out + ((size_t)rep * a.m + row) * n;
#pragma unroll
  for (int u = 0; u < U; ++u)
    if (ok[u]) {
#pragma unroll
      for (int v = 0; v < VEC; ++v) {
        float x = acc[u][v] + bias;
        acc[u][v] = a.relu ? fmaxf(x, 0.f) : x;
      }
      store_vec<VEC>(out + n0 + u * 32 * VEC, acc[u]);
    }
}

// Rows per CTA (benchmark knob "spmm_warps"; 0 = default).  With a narrow dense operand the rows
// of a CTA walk the column space in near lockstep, so a wide CTA keeps the operand rows they
// share in L1 ("implicit banding").
extern std::atomic<int> g_spmm_warps;

template <int VEC, int U, int WARPS>
inline void launch_spmm_gather_w(const SpmmArgs& a, int replicas, cudaStream_t stream) {
  dim3 grid(ceil_div(a.m, WARPS), ceil_div(a.n, 32 * VEC * U), replicas);
  if ((unsigned long long)a.k * (unsigned long long)a.n * 4ull < (1ull << 31))
    spmm_gather_kernel<VEC, U, WARPS, true><<<grid, WARPS * 32, 0, stream>>>(a);
  else
    spmm_gather_kernel<VEC, U, WARPS, false><<<grid, WARPS * 32, 0, stream>>>(a);
}

template <int VEC, int U>
inline void launch_spmm_gather(const SpmmArgs& a, int replicas, cudaStream_t stream) {
  int w = g_spmm_warps.load();
  if (w == 0) w = (a.n <= 32 * VEC * U && a.n <= 128 && a.k >= 1024) ? 16 : 4;
  switch (w) {
    case 8: launch_spmm_gather_w<VEC, U, 8>(a, replicas, stream); break;
    case 16: launch_spmm_gather_w<VEC, U, 16>(a, replicas, stream); break;
    case 32: launch_spmm_gather_w<VEC, U, 32>(a, replicas, stream); break;
    default: launch_spmm_gather_w<VEC, U, 4>(a, replicas, stream); break;
  }
}

}  // namespace tsp
