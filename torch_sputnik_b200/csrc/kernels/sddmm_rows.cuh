// sddmm_rows.cuh -- SDDMM, variant 5 ("rows"): the path for a SHORT inner dimension
// (k = 32, 64, 96 or 128: attention scores, out = sample(Q @ K^T) with head_dim 64).
//
// out[j] = <lhs[row_of(j), :], rhs[col(j), :]>.  With k = 64 an rhs row is 256 bytes; the general
// gather kernel (variant 1) then spends ~25 instructions per nonzero on bookkeeping (two shuffles
// and a 64-bit address per 4 FMAs, a 16-lane butterfly per 16 nonzeros) and is issue- and
// latency-bound (profiles/r01_sddmm_gather_c4_ncu_summary.txt).  Here:
//
//  * one warp owns one mask row and keeps its lhs k-slices in registers for the whole row: lane
//    `sub` of every quarter-warp holds the 16-byte chunks sub, sub+8, ... of lhs[row, :];
//  * the row's nonzeros are taken 32 at a time (one coalesced load of column indices); in step g
//    quarter q works on nonzero 4g+q: each lane reads the same chunks of rhs[col, :], so a
//    quarter-warp LDG.128 is one contiguous 128-byte request, and the address is ONE wide
//    multiply-add (column x row-vectors onto the lane's base pointer, IMAD.WIDE.U32);
//  * a transposing butterfly over the 8 lanes of a quarter (7 shuffles per 32 nonzeros instead
//    of 3 per nonzero) leaves nonzero 4*sub+q in lane (q, sub); the 32 results go out in one
//    128-byte segment;
//  * a CTA is 16 consecutive rows: with ascending columns they sweep the rhs rows in near
//    lockstep, so the rows they share are L1 hits ("implicit banding").
//
// Arithmetic: per lane two fp32 FMA chains (even / odd k) in ascending k inside its chunks, their sum, then a fp32 tree over the 8
// lanes -- the shape of the reference's (all_reduce.h:61-80), a different association.
// Works for any column order (no sortedness assumption).  Needs k % 32 == 0, k <= 128,
// 16-byte aligned operands and n * k * 4 < 2^31.
// 16-bit storage (T = __half / __nv_bfloat16, fp32 products and sums, result rounded once): the
// same kernel with 8 elements per 16-byte chunk -- k % 64 == 0, k <= 256; a k = 64 row is ONE
// 128-byte request, half the L1 wavefronts per FMA of the fp32 layout.
#pragma once
#include "sddmm_gather.cuh"

namespace tsp {

// T = storage type; a quarter-warp reads 128 bytes of a row per step = 8 lanes x E elements
// (E = 4 fp32 / 8 fp16 or bf16), so k = 8 * E * KC: 16-bit operands need half the steps per nonzero.
template <int KC, int WARPS, typename T = float>
__global__ void __launch_bounds__(WARPS * 32) sddmm_rows_kernel(const SddmmArgsT<T> a) {
  constexpr int E = Elem<T>::kPer16;
  constexpr int K = 8 * E * KC;
  if (fallback_not_needed(a.guard_flag)) return;   // queued behind a band launch whose guard held
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * WARPS + warp;
  if (row >= a.m) return;
  const int rep = blockIdx.y;
  const int q = lane >> 3, sub = lane & 7;

  const MaskView mv = mask_view(a.masks, rep, a.nnz);
  const int* __restrict__ row_offsets = a.row_offsets + mv.topo_rows * (a.m + 1);
  const int beg = __ldg(row_offsets + row), end = __ldg(row_offsets + row + 1);
  if (beg >= end) return;
  const int* __restrict__ cols = a.column_indices + mv.cols;
  T* __restrict__ out = a.out + mv.values;
  // this lane's chunk `sub` of rhs row 0 (16-byte units); chunk cc of row c is at + c * (8 * KC) + cc * 8
  const T* const rbase = a.rhs + (size_t)rep * a.n * K + sub * E;
  const T* const lrow = a.lhs + ((size_t)rep * a.m + row) * K + sub * E;
  // fp32: lhs slices widened once; 16-bit: kept packed (4 registers per chunk) for the mixed-precision FMA
  float lv[KC][E];
  uint4 lraw[KC];
#pragma unroll
  for (int cc = 0; cc < KC; ++cc) {
    if constexpr (sizeof(T) == 4) load_elems<T, E>(lv[cc], lrow + cc * 8 * E);
    else lraw[cc] = __ldg(reinterpret_cast<const uint4*>(lrow + cc * 8 * E));
  }
  // Even and odd elements of the lane's chunks are summed separately as one packed fma.rn.f32x2
  // chain (half the issue slots), then added -- the SAME association as sddmm_band (band.cuh), so the
  // two kernels the heuristic chooses between for a short inner dimension agree bit for bit.
  auto dot_row = [&](const unsigned o) {   // o: rhs row offset in 16-byte units
    float s0 = 0.f, s1 = 0.f;
#pragma unroll
    for (int cc = 0; cc < KC; ++cc) {
      if constexpr (sizeof(T) == 4) {
        float x[E];
        load_elems<T, E>(x, rbase + ((size_t)o + cc * 8) * E);
#pragma unroll
        for (int e = 0; e < E; e += 2) ffma2(s0, s1, lv[cc][e], lv[cc][e + 1], x[e], x[e + 1]);
      } else {   // 16-bit storage: one chain of mixed-precision FMAs over the packed pairs, ascending k
        const uint4 x = __ldg(reinterpret_cast<const uint4*>(rbase + ((size_t)o + cc * 8) * E));
        s0 = fma_pair_16<T>(lraw[cc].x, x.x, s0);
        s0 = fma_pair_16<T>(lraw[cc].y, x.y, s0);
        s0 = fma_pair_16<T>(lraw[cc].z, x.z, s0);
        s0 = fma_pair_16<T>(lraw[cc].w, x.w, s0);
      }
    }
    return s0 + s1;
  };

  // The column batch of the NEXT 32 nonzeros is requested one batch ahead: its latency (an L2 hit for a
  // streamed index array) hides under this batch's gathers instead of heading every batch.
  unsigned nextoff = beg + lane < end ? (unsigned)__ldg(cols + beg + lane) * (unsigned)(8 * KC) : 0u;
  for (int p = beg; p < end; p += 32) {
    const int cnt = min(32, end - p);
    // rhs row offset of nonzero p + lane in 16-byte units (row 0 for the lanes past the row's end:
    // a valid address, the result is never stored)
    const unsigned myoff = nextoff;
    nextoff = p + 32 + lane < end ? (unsigned)__ldg(cols + p + 32 + lane) * (unsigned)(8 * KC) : 0u;
    float acc[8];
    if (cnt == 32) {
#pragma unroll
      for (int g = 0; g < 8; ++g) {
        acc[g] = dot_row(__shfl_sync(kFullMask, myoff, 4 * g + q));
      }
    } else {
#pragma unroll
      for (int g = 0; g < 8; ++g) {
        acc[g] = 0.f;
        if (4 * g < cnt) {   // warp-uniform
          acc[g] = dot_row(__shfl_sync(kFullMask, myoff, 4 * g + q));
        }
      }
    }
    transpose_reduce<8>(acc, sub);   // lane (q, sub) now holds nonzero 4 * sub + q of the batch
    const int j = 4 * sub + q;
    if (j < cnt) out[p + j] = Elem<T>::from_f(acc[0]);
  }
}

template <typename T>
inline bool sddmm_rows_applicable(const SddmmArgsT<T>& a) {
  constexpr int chunk = 8 * Elem<T>::kPer16;   // elements per quarter-warp step: 32 fp32 / 64 16-bit
  if (a.k % chunk != 0 || a.k > 4 * chunk || a.k < chunk) return false;
  if (!aligned16(a.lhs) || !aligned16(a.rhs)) return false;
  return (unsigned long long)a.n * (unsigned long long)a.k * sizeof(T) < (1ull << 31);
}

extern std::atomic<int> g_sddmm_warps;   // benchmark knob "sddmm_warps": rows per CTA (0 = default 16)

template <int KC, typename T>
inline void launch_sddmm_rows_kc(const SddmmArgsT<T>& a, int replicas, cudaStream_t stream) {
  const int w = g_sddmm_warps.load();
  if (w == 4) sddmm_rows_kernel<KC, 4, T><<<dim3(ceil_div(a.m, 4), replicas), 128, 0, stream>>>(a);
  else if (w == 8) sddmm_rows_kernel<KC, 8, T><<<dim3(ceil_div(a.m, 8), replicas), 256, 0, stream>>>(a);
  else if (w == 32) sddmm_rows_kernel<KC, 32, T><<<dim3(ceil_div(a.m, 32), replicas), 1024, 0, stream>>>(a);
  else sddmm_rows_kernel<KC, 16, T><<<dim3(ceil_div(a.m, 16), replicas), 512, 0, stream>>>(a);
}

template <typename T>
inline void launch_sddmm_rows(const SddmmArgsT<T>& a, int replicas, cudaStream_t stream) {
  switch (a.k / (8 * Elem<T>::kPer16)) {
    case 1: launch_sddmm_rows_kc<1, T>(a, replicas, stream); break;
    case 2: launch_sddmm_rows_kc<2, T>(a, replicas, stream); break;
    case 3: launch_sddmm_rows_kc<3, T>(a, replicas, stream); break;
    default: launch_sddmm_rows_kc<4, T>(a, replicas, stream); break;
  }
}

}  // namespace tsp
