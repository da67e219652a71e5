// attention_rows.cuh -- fused sparse attention forward:
//     out[r, i, :] = sum_j softmax_j(scale * <q[r,i,:], k[r,col_j,:]>) * v[r, col_j, :]
// over the nonzeros j of mask row i -- the chain sddmm -> sparse_softmax -> spmm of
// modules/sparse_attention.py:66-82 in ONE pass, so the two [R, nnz] intermediates (scores and
// probabilities: 16 of the 18.5 bytes per nonzero that the three-kernel chain moves through HBM)
// never exist.
//
// Shape of the work (head_dim d = 32, 64, 96 or 128 in fp32; 64 .. 256 in fp16 / bf16 storage):
// the same as sddmm_rows.cuh, twice.
//  * one warp owns one mask row; lane `sub` of each quarter-warp keeps the 16-byte chunks
//    sub, sub+8, .. of q[row, :] (pre-multiplied by scale * log2 e) and of the output row;
//  * nonzeros are taken 32 at a time.  Score phase: in step g quarter `qq` dots its q chunks with
//    the same chunks of k[col(4g+qq), :] (one 128-byte request per quarter, one IMAD.WIDE per
//    address); a transposing butterfly leaves the score of nonzero 4*sub+qq in lane (qq, sub);
//  * online softmax across batches (running maximum, denominators and output rescaled by
//    2^(old max - new max) once per 32 nonzeros; one EX2 per nonzero);
//  * value phase: step g quarter qq fetches its nonzero's probability with one shuffle (from lane
//    (qq, g)) and accumulates p * v[col, chunks] -- the same addresses as the score phase in the
//    other operand;
//  * the four quarters' partial output rows are summed by shuffles at the end of the row and
//    scaled by 1 / denominator.
// Arithmetic: fp32 throughout; exp(x) is evaluated as 2^(x log2 e) with the hardware EX2 (the
// reference chain uses expf, sparse_softmax.cu.cc:52; the difference is below 2 ulp of the
// probability).  Any column order.  Empty rows produce zeros (as the chain does).
#pragma once
#include "sddmm_gather.cuh"

namespace tsp {

// T: storage type of q / k / v / out (float, __half, __nv_bfloat16); scores, probabilities and sums fp32.
template <typename T>
struct AttnArgsT {
  int m, n, d, nnz;
  const int* row_offsets;
  const int* column_indices;
  const T* q;    // [replicas, m, d]
  const T* k;    // [replicas, n, d]
  const T* v;    // [replicas, n, d]
  T* out;        // [replicas, m, d]
  float* lse;    // optional [replicas, m]: log(sum exp(scale * score)) per row (natural log)
  float scale;
};
using AttnArgs = AttnArgsT<float>;

// d = 8 * E * KC with E elements per 16-byte chunk (4 fp32 / 8 16-bit): a quarter-warp covers 128
// bytes of a k / v row per step, so 16-bit storage halves the L1 wavefronts per nonzero.
template <int KC, int WARPS, typename T = float>
__global__ void __launch_bounds__(WARPS * 32) attention_rows_kernel(const AttnArgsT<T> a) {
  constexpr int E = Elem<T>::kPer16;
  constexpr int D = 8 * E * KC;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * WARPS + warp;
  if (row >= a.m) return;
  const int rep = blockIdx.y;
  const int qq = lane >> 3, sub = lane & 7;

  const int beg = __ldg(a.row_offsets + row), end = __ldg(a.row_offsets + row + 1);
  T* const orow = a.out + ((size_t)rep * a.m + row) * D + sub * E;
  if (beg >= end) {
    if (qq == 0) {
      float z[E];
#pragma unroll
      for (int e = 0; e < E; ++e) z[e] = 0.f;
#pragma unroll
      for (int cc = 0; cc < KC; ++cc) store_elems<T, E>(orow + cc * 8 * E, z);
    }
    if (a.lse != nullptr && lane == 0) a.lse[(size_t)rep * a.m + row] = -INFINITY;
    return;
  }
  const int* __restrict__ cols = a.column_indices;
  const T* const kbase = a.k + (size_t)rep * a.n * D + sub * E;
  const T* const vbase = a.v + (size_t)rep * a.n * D + sub * E;
  const T* const qrow = a.q + ((size_t)rep * a.m + row) * D + sub * E;
  const float qs = a.scale * 1.4426950408889634f;
  float qv[KC][E], acc[KC][E];
#pragma unroll
  for (int cc = 0; cc < KC; ++cc) {
    load_elems<T, E>(qv[cc], qrow + cc * 8 * E);
#pragma unroll
    for (int e = 0; e < E; ++e) {
      qv[cc][e] *= qs;
      acc[cc][e] = 0.f;
    }
  }
  float run_max = -INFINITY;   // warp-uniform, in log2 units
  float denom = 0.f;           // this lane's share of the denominator, relative to run_max

  for (int p = beg; p < end; p += 32) {
    const int cnt = min(32, end - p);
    // rows of k / v in 16-byte units; lanes past the end of the row repeat its last nonzero (a valid
    // address whose probability is forced to zero below)
    const unsigned myoff = (unsigned)__ldg(cols + p + min(lane, cnt - 1)) * (unsigned)(8 * KC);
    float s[8];
#pragma unroll
    for (int g = 0; g < 8; ++g) {
      s[g] = 0.f;
      if (4 * g < cnt) {   // warp-uniform
        const unsigned off = __shfl_sync(kFullMask, myoff, 4 * g + qq);
        float t = 0.f;
#pragma unroll
        for (int cc = 0; cc < KC; ++cc) {
          float x[E];
          load_elems<T, E>(x, kbase + ((size_t)off + cc * 8) * E);
#pragma unroll
          for (int e = 0; e < E; ++e) t = fmaf(qv[cc][e], x[e], t);
        }
        s[g] = t;
      }
    }
    transpose_reduce<8>(s, sub);   // lane (qq, sub): score (log2 units) of nonzero 4 * sub + qq
    const float score = (4 * sub + qq < cnt) ? s[0] : -INFINITY;
    const float new_max = fmaxf(run_max, warp_max(score));
    const float corr = exp2f(run_max - new_max);   // 0 on the first batch (run_max = -inf)
    const float prob = exp2f(score - new_max);     // 0 for the padding lanes
    denom = fmaf(denom, corr, prob);
    run_max = new_max;
#pragma unroll
    for (int cc = 0; cc < KC; ++cc)
#pragma unroll
      for (int e = 0; e < E; ++e) acc[cc][e] *= corr;
#pragma unroll
    for (int g = 0; g < 8; ++g) {
      if (4 * g < cnt) {
        const unsigned off = __shfl_sync(kFullMask, myoff, 4 * g + qq);   // (cheaper than 8 live registers)
        const float w = __shfl_sync(kFullMask, prob, 8 * qq + g);
#pragma unroll
        for (int cc = 0; cc < KC; ++cc) {
          float x[E];
          load_elems<T, E>(x, vbase + ((size_t)off + cc * 8) * E);
#pragma unroll
          for (int e = 0; e < E; ++e) acc[cc][e] = fmaf(w, x[e], acc[cc][e]);
        }
      }
    }
  }

  const float total = warp_sum(denom);
  const float inv = 1.0f / total;
#pragma unroll
  for (int cc = 0; cc < KC; ++cc) {
    float t[E];
#pragma unroll
    for (int e = 0; e < E; ++e) {
      float x = acc[cc][e];
      x += __shfl_xor_sync(kFullMask, x, 8);
      x += __shfl_xor_sync(kFullMask, x, 16);
      t[e] = x * inv;
    }
    if (qq == (cc & 3)) store_elems<T, E>(orow + cc * 8 * E, t);
  }
  if (a.lse != nullptr && lane == 0)
    a.lse[(size_t)rep * a.m + row] = (run_max + log2f(total)) * 0.6931471805599453f;
}

template <typename T>
inline bool attention_rows_applicable(const AttnArgsT<T>& a) {
  constexpr int chunk = 8 * Elem<T>::kPer16;
  if (a.d % chunk != 0 || a.d > 4 * chunk || a.d < chunk) return false;
  if (!aligned16(a.q) || !aligned16(a.k) || !aligned16(a.v) || !aligned16(a.out)) return false;
  return (unsigned long long)a.n * (unsigned long long)a.d * sizeof(T) < (1ull << 31);
}

extern std::atomic<int> g_attn_warps;   // benchmark knob "attn_warps": rows per CTA (0 = default)

template <int KC, int WARPS, typename T>
inline void launch_attention_rows_w(const AttnArgsT<T>& a, int replicas, cudaStream_t stream) {
  attention_rows_kernel<KC, WARPS, T><<<dim3(ceil_div(a.m, WARPS), replicas), WARPS * 32, 0, stream>>>(a);
}

template <int KC, typename T>
inline void launch_attention_rows_kc(const AttnArgsT<T>& a, int replicas, cudaStream_t stream) {
  const int w = g_attn_warps.load();
  if (w == 4) launch_attention_rows_w<KC, 4, T>(a, replicas, stream);
  else if (w == 8) launch_attention_rows_w<KC, 8, T>(a, replicas, stream);
  else if (w == 16) launch_attention_rows_w<KC, 16, T>(a, replicas, stream);
  else launch_attention_rows_w<KC, 32, T>(a, replicas, stream);
}

template <typename T>
inline void launch_attention_rows(const AttnArgsT<T>& a, int replicas, cudaStream_t stream) {
  switch (a.d / (8 * Elem<T>::kPer16)) {
    case 1: launch_attention_rows_kc<1, T>(a, replicas, stream); break;
    case 2: launch_attention_rows_kc<2, T>(a, replicas, stream); break;
    case 3: launch_attention_rows_kc<3, T>(a, replicas, stream); break;
    default: launch_attention_rows_kc<4, T>(a, replicas, stream); break;
  }
}

}  // namespace tsp
