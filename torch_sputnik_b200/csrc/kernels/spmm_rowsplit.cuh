// spmm_rowsplit.cuh -- CSR SpMM for NARROW dense operands and for 16-bit storage ("row split").
//
// One warp owns one sparse row.  The gather kernel (spmm_gather.cuh) gives every lane its own
// slice of the output row and walks the row's nonzeros one at a time, so with a narrow operand
// (attention: n = 64 -> a 256-byte fp32 / 128-byte fp16 row of V) half or three quarters of the
// lanes have no columns to own.  Here the warp is cut into G = 32 / LPN groups of LPN lanes; a
// group covers the whole operand row with 16-byte vectors (LPN * 16 bytes) and the G groups work
// on G DIFFERENT nonzeros at once, so every load instruction moves 512 bytes whatever n is, and
// the per-nonzero bookkeeping (two shuffles) is shared by G nonzeros.  The G partial rows are
// summed with a butterfly at the end.
//
// T is the storage type (float, __half, __nv_bfloat16): 16-bit operands halve the bytes per FMA
// through the L1 data pipe, which is what bounds every gather-type kernel;
// products and sums are fp32, the result is rounded once on the way out -- the contract of the
// reference's half SpMM (cuda_spmm.cu.cc:511-570, cuda_spmm.h:75-113) without its even-nnz /
// even-n / short2-index restrictions.  EPL (elements per lane-load) falls back to 4/2/1 for
// operands whose rows are not 16-byte multiples, and operands wider than 32 * EPL columns are
// tiled over grid.y with LPN = 32.
//
// Arithmetic: per output element fp32 FMAs over the row's nonzeros in G interleaved chains
// (nonzero j goes to chain j mod G) then a tree over the chains; the reference has one chain
// (cuda_spmm.cu.cc:224-238).  Tolerance parity, not bit parity, with spmm_gather.
// Non-finite operands: a nonzero past the row's end is never multiplied (no 0 * Inf).
#pragma once
#include "spmm_gather.cuh"

namespace tsp {

// (Measured twice and dropped: sending a 16-bit value and its column between lanes as ONE packed word -- one
// shuffle per step instead of two -- is slower both with widening FMAs (2.51 vs 2.31 ms at the C4 shape) and
// with the mixed-precision FMA, where the value half needs no unpacking at all (2.19 vs 2.03 ms): the second
// shuffle is not what the kernel waits for.  profiles/r02_band_ab.txt)
template <typename T, int EPL, int LPN, int WARPS>
__global__ void __launch_bounds__(WARPS * 32) spmm_rowsplit_kernel(const SpmmArgsT<T> a) {
  constexpr int G = 32 / LPN;
  constexpr int UNR = LPN >= 4 ? 4 : LPN;   // gathered loads in flight per lane
  if (fallback_not_needed(a.guard_flag)) return;   // queued behind a band launch whose guard held
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int item = blockIdx.x * WARPS + warp;
  if (item >= a.m) return;
  const int row = a.row_indices ? __ldg(a.row_indices + item) : item;
  const int rep = blockIdx.z;
  const int grp = lane / LPN, sub = lane % LPN;

  const int n = a.n;
  const int n0 = blockIdx.y * (LPN * EPL) + sub * EPL;
  const bool col_ok = n0 < n;                      // n % EPL == 0: a lane's vector is all in or all out
  const unsigned row_units = (unsigned)(n / EPL);  // operand row stride in units of one lane-load
  const T* __restrict__ lane_base = a.dense + (size_t)rep * a.k * n + (col_ok ? n0 : 0);
  const T* __restrict__ vals = a.values + (size_t)rep * a.values_stride;
  const int* __restrict__ cols = a.column_indices;

  float acc[EPL];
#pragma unroll
  for (int e = 0; e < EPL; ++e) acc[e] = 0.f;

  const int beg = __ldg(a.row_offsets + row), end = __ldg(a.row_offsets + row + 1);
  // 16-bit storage with full vectors: the value travels between lanes as its raw 16 bits and meets the
  // packed operand pairs in mixed-precision FMAs (no widening instructions)
  constexpr bool MIXED = sizeof(T) == 2 && EPL == 8;
  // (value, offset) of nonzero idx for this lane; fetched one batch AHEAD so that the index / value latency
  // hides under the current batch's gathers
  float nextval = 0.f;
  unsigned nextoff = 0u;
  auto fetch = [&](int idx) {
    nextval = 0.f;
    nextoff = 0u;
    if (idx < end) {
      const T raw = vals[a.value_perm ? __ldg(a.value_perm + idx) : idx];
      if constexpr (MIXED) nextval = __uint_as_float((unsigned)*reinterpret_cast<const unsigned short*>(&raw));
      else nextval = Elem<T>::to_f(raw);
      nextoff = (unsigned)__ldg(cols + idx) * row_units;
    }
  };
  fetch(beg + lane);
  for (int base = beg; base < end; base += 32) {
    const int cnt = min(32, end - base);
    const float myval = nextval;
    const unsigned myoff = nextoff;
    fetch(base + 32 + lane);
    // step g: group `grp` takes nonzero g * G + grp of the batch
#pragma unroll
    for (int g0 = 0; g0 < LPN; g0 += UNR) {
      if (g0 * G >= cnt) break;   // warp-uniform
      float v[UNR];
      typename RawVec<EPL * sizeof(T)>::type raw[UNR];
      bool on[UNR];
#pragma unroll
      for (int u = 0; u < UNR; ++u) {
        const int src = (g0 + u) * G + grp;
        v[u] = __shfl_sync(kFullMask, myval, src);
        const unsigned o = __shfl_sync(kFullMask, myoff, src);
        on[u] = src < cnt && col_ok;
        if (on[u]) raw[u] = ldg_raw<T, EPL>(lane_base + (size_t)o * EPL);
      }
#pragma unroll
      for (int u = 0; u < UNR; ++u)
        if (on[u]) {
          if constexpr (MIXED) {
            const unsigned w = __float_as_uint(v[u]);
            fma_scalar_pair_16<T>(w, raw[u].x, acc[0], acc[1]);
            fma_scalar_pair_16<T>(w, raw[u].y, acc[2], acc[3]);
            fma_scalar_pair_16<T>(w, raw[u].z, acc[4], acc[5]);
            fma_scalar_pair_16<T>(w, raw[u].w, acc[6], acc[7]);
          } else {
            float b[EPL];
            unpack_raw<T, EPL>(b, raw[u]);
#pragma unroll
            for (int e = 0; e < EPL; ++e) acc[e] = fmaf(v[u], b[e], acc[e]);
          }
        }
    }
  }

#pragma unroll
  for (int s = LPN; s < 32; s <<= 1)
#pragma unroll
    for (int e = 0; e < EPL; ++e) acc[e] += __shfl_xor_sync(kFullMask, acc[e], s);

  if (grp == 0 && col_ok) {
    const float bias = a.bias ? __ldg(a.bias + row) : 0.f;
#pragma unroll
    for (int e = 0; e < EPL; ++e) {
      const float x = acc[e] + bias;
      acc[e] = a.relu ? fmaxf(x, 0.f) : x;
    }
    store_elems<T, EPL>(a.out + ((size_t)rep * a.m + row) * n + n0, acc);
  }
}

// Largest EPL (elements per lane-load, <= one 16-byte vector) the operand layout allows.
template <typename T>
inline int spmm_rowsplit_epl(const SpmmArgsT<T>& a) {
  const uintptr_t bits = reinterpret_cast<uintptr_t>(a.dense) | reinterpret_cast<uintptr_t>(a.out);
  for (int epl = Elem<T>::kPer16; epl > 1; epl >>= 1)
    if (a.n % epl == 0 && (bits % (epl * sizeof(T))) == 0) return epl;
  return 1;
}

// The row offset of an operand row is carried as a 32-bit count of lane-loads.
template <typename T>
inline bool spmm_rowsplit_applicable(const SpmmArgsT<T>& a) {
  const int epl = spmm_rowsplit_epl(a);
  return (unsigned long long)a.k * (unsigned long long)(a.n / epl) < (1ull << 32);
}

extern std::atomic<int> g_spmm_warps;

template <typename T, int EPL, int LPN>
inline void launch_spmm_rowsplit_l(const SpmmArgsT<T>& a, int replicas, cudaStream_t stream) {
  // 16 rows per CTA when the operand is narrow and tall: the rows of a CTA sweep the operand's rows
  // in near lockstep and share them in L1 (the "implicit banding" of spmm_gather.cuh)
  int w = g_spmm_warps.load();
  if (w == 0) w = (LPN < 32 && a.k >= 1024) ? 16 : 4;
  dim3 grid(1, ceil_div(a.n, LPN * EPL), replicas);
  if (w >= 16) {
    grid.x = ceil_div(a.m, 16);
    spmm_rowsplit_kernel<T, EPL, LPN, 16><<<grid, 16 * 32, 0, stream>>>(a);
  } else if (w == 8) {
    grid.x = ceil_div(a.m, 8);
    spmm_rowsplit_kernel<T, EPL, LPN, 8><<<grid, 8 * 32, 0, stream>>>(a);
  } else {
    grid.x = ceil_div(a.m, 4);
    spmm_rowsplit_kernel<T, EPL, LPN, 4><<<grid, 4 * 32, 0, stream>>>(a);
  }
}

template <typename T, int EPL>
inline void launch_spmm_rowsplit_e(const SpmmArgsT<T>& a, int replicas, cudaStream_t stream) {
  const int units = ceil_div(a.n, EPL);   // lane-loads per operand row
  if (units <= 4) launch_spmm_rowsplit_l<T, EPL, 4>(a, replicas, stream);
  else if (units <= 8) launch_spmm_rowsplit_l<T, EPL, 8>(a, replicas, stream);
  else if (units <= 16) launch_spmm_rowsplit_l<T, EPL, 16>(a, replicas, stream);
  else launch_spmm_rowsplit_l<T, EPL, 32>(a, replicas, stream);
}

template <typename T>
inline void launch_spmm_rowsplit(const SpmmArgsT<T>& a, int replicas, cudaStream_t stream) {
  const int epl = spmm_rowsplit_epl(a);
  if constexpr (sizeof(T) == 2) {
    if (epl == 8) return launch_spmm_rowsplit_e<T, 8>(a, replicas, stream);
  }
  if (epl >= 4) return launch_spmm_rowsplit_e<T, 4>(a, replicas, stream);
  if (epl == 2) return launch_spmm_rowsplit_e<T, 2>(a, replicas, stream);
  launch_spmm_rowsplit_e<T, 1>(a, replicas, stream);
}

}  // namespace tsp
