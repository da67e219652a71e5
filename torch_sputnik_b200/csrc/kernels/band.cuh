// band.cuh -- SDDMM and SpMM for a SHORT dense dimension (attention: head_dim 32..128 fp32,
// 64..256 in 16-bit storage) with the gathered operand STAGED IN SHARED MEMORY ("band" kernels).
//
// Why.  The row kernels (sddmm_rows.cuh, spmm_gather.cuh) gather K / V rows with LDG.  A warp-wide
// LDG.128 that touches four different 128-byte lines is replayed once per line at ~2 cycles per
// line (measured: 2.99 ms for the C4 SDDMM = 4.1 cycles per nonzero per SM, exactly 2 lines x 2.07
// cycles; profiles/r02_c4_band.txt), so the L1 path delivers ~62 bytes/clk/SM however the lanes are
// arranged.  Shared memory delivers 128 bytes/clk/SM for the same quarter-warp-per-row pattern.
// So: a CTA owns BM = WARPS * RPW consecutive mask rows of one replica and walks the COLUMN space in
// blocks of BN operand rows; each block (BN x D elements, contiguous in memory) is brought in by one
// bulk async copy (cp.async.bulk -> mbarrier complete_tx, 2-stage ring, the copy of block b+1 flies
// under the arithmetic of block b); a warp visits each of its RPW rows, takes the row's nonzeros
// whose column falls in the block (columns are ascending, so that is a prefix of what is left of
// the row) up to 32 at a time, a quarter-warp per nonzero: one conflict-free LDS.128 per lane per
// 128 bytes of operand row.  Blocks before the first / after the last column any row of the CTA
// holds are skipped (causal and banded masks touch about half the blocks).
//
//   sddmm_band: out[j] = <lhs[row, :], rhs[col_j, :]>; the lhs rows of a warp live in registers for
//               the whole kernel; a transposing butterfly per 32 nonzeros (as sddmm_rows.cuh).
//   spmm_band:  out[row, :] = sum_j values[j] * dense[col_j, :]; RPW x D/8 accumulators per lane,
//               the four quarters' partial rows summed by shuffles at the end; bias / ReLU epilogue;
//               optional value permutation (A^T @ B on a transposed topology).
//
// T = float, __half or __nv_bfloat16 (fp32 products and sums, one rounding of the result).
// Arithmetic: SDDMM as sddmm_rows.cuh; SpMM sums a row's nonzeros in four interleaved chains
// (nonzero j of a visit goes to chain j mod 4) then a tree -- tolerance parity with the gather
// kernel, not bit parity.  A nonzero is only ever multiplied with its own operand row (no 0 * Inf).
// Precondition: ascending columns inside each row (the plan's verdict; StagedGuard otherwise),
// D * sizeof(T) in {128, 256, 384, 512}, 16-byte aligned operands.
#pragma once
#include "sddmm_rows.cuh"
#include "spmm_gather.cuh"

namespace tsp {

constexpr int kBandStages = 2;
constexpr int kBandStageBytes = 64 * 1024;

template <typename T, int KC>
struct BandGeom {
  static constexpr int E = Elem<T>::kPer16;       // elements per 16-byte chunk
  static constexpr int D = 8 * E * KC;            // dense dimension handled
  static constexpr int ROWB = D * (int)sizeof(T); // bytes per operand row = 128 * KC
  static constexpr int ROW16 = 8 * KC;            // 16-byte chunks per operand row
  static int block_rows(int n) {                  // operand rows per staged block
    int bn = kBandStageBytes / ROWB;
    if (n < bn) bn = (n + 7) & ~7;
    return bn;
  }
  static size_t smem_bytes(int bn) { return (size_t)kBandStages * bn * ROWB + 2 * kBandStages * 8 + 16 + 128; }
};

// Ring bookkeeping shared by the two kernels.
struct BandRing {
  unsigned char* ring;
  unsigned long long* full;
  unsigned long long* empty;
  int* range;   // [2]: first / one-past-last column block any row of the CTA touches
};

template <int WARPS>
__device__ __forceinline__ BandRing band_setup(unsigned char* smem_dyn, int stage_bytes) {
  BandRing r;
  r.ring = smem_dyn + ((128u - (smem_u32(smem_dyn) & 127u)) & 127u);
  r.full = reinterpret_cast<unsigned long long*>(r.ring + (size_t)kBandStages * stage_bytes);
  r.empty = r.full + kBandStages;
  r.range = reinterpret_cast<int*>(r.empty + kBandStages);
  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < kBandStages; ++s) {
      mbar_init(r.full + s, 1);
      mbar_init(r.empty + s, WARPS);
    }
    mbar_fence_init();
    r.range[0] = 0x7fffffff;
    r.range[1] = 0;
  }
  return r;
}

// Thread 0: bring operand rows [cb * bn, min(n, (cb + 1) * bn)) of `base` into stage `it % STAGES`.
template <int ROWB>
__device__ __forceinline__ void band_issue(const BandRing& r, const unsigned char* base, int n, int bn, int cb,
                                           int it, int stage_bytes) {
  const int st = it % kBandStages, use = it / kBandStages;
  if (use > 0) mbar_wait(r.empty + st, (use - 1) & 1);
  const int rows = min(bn, n - cb * bn);
  const unsigned bytes = (unsigned)rows * ROWB;
  mbar_arrive_expect_tx(r.full + st, bytes);
  bulk_g2s(r.ring + (size_t)st * stage_bytes, base + (size_t)cb * bn * ROWB, bytes, r.full + st);
}

// ------------------------------------------------------------------------------------- SDDMM
template <typename T, int KC, int WARPS, int RPW>
__global__ void __launch_bounds__(WARPS * 32, 1)
sddmm_band_kernel(const SddmmArgsT<T> a, const StagedGuard guard, const int bn) {
  using G = BandGeom<T, KC>;
  constexpr int E = G::E, ROWB = G::ROWB, ROW16 = G::ROW16;
  if (staged_declines(guard)) return;
  extern __shared__ unsigned char smem_dyn[];
  const int stage_bytes = bn * ROWB;
  const BandRing ring = band_setup<WARPS>(smem_dyn, stage_bytes);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q = lane >> 3, sub = lane & 7;
  const int rep = blockIdx.y;
  const int row0 = blockIdx.x * (WARPS * RPW) + warp * RPW;
  const int* __restrict__ cols = a.column_indices;
  T* __restrict__ out = a.out + (size_t)rep * a.nnz;

  // cpre[i]: the columns of nonzeros pos[i] + lane of row i (INT_MAX past the row's end), fetched
  // one visit AHEAD -- right after a visit consumed its batch -- so the load's latency hides under
  // the other rows' visits and the next block's barrier instead of heading every visit.
  int pos[RPW], endp[RPW], cpre[RPW];
  float lv[RPW][KC][E];
  int lo = 0x7fffffff, hi = -1;
#pragma unroll
  for (int i = 0; i < RPW; ++i) {
    const int row = row0 + i;
    pos[i] = endp[i] = 0;
    cpre[i] = 0x7fffffff;
    if (row < a.m) {
      pos[i] = __ldg(a.row_offsets + row);
      endp[i] = __ldg(a.row_offsets + row + 1);
      if (pos[i] + lane < endp[i]) cpre[i] = __ldg(cols + pos[i] + lane);
      const T* lrow = a.lhs + ((size_t)rep * a.m + row) * G::D + sub * E;
#pragma unroll
      for (int cc = 0; cc < KC; ++cc) load_elems<T, E>(lv[i][cc], lrow + cc * 8 * E);
      if (pos[i] < endp[i]) {
        lo = min(lo, __ldg(cols + pos[i]));
        hi = max(hi, __ldg(cols + endp[i] - 1));
      }
    } else {
#pragma unroll
      for (int cc = 0; cc < KC; ++cc)
#pragma unroll
        for (int e = 0; e < E; ++e) lv[i][cc][e] = 0.f;
    }
  }
  __syncthreads();   // barriers and range initialised
  if (lane == 0 && hi >= 0) {
    atomicMin(ring.range, lo / bn);
    atomicMax(ring.range + 1, hi / bn + 1);
  }
  __syncthreads();
  const int cb_lo = ring.range[0], cb_hi = ring.range[1];
  if (cb_lo >= cb_hi) return;   // no nonzeros in this row block

  const unsigned char* const rhs = reinterpret_cast<const unsigned char*>(a.rhs + (size_t)rep * a.n * G::D);
  if (threadIdx.x == 0) band_issue<ROWB>(ring, rhs, a.n, bn, cb_lo, 0, stage_bytes);

  for (int cb = cb_lo, it = 0; cb < cb_hi; ++cb, ++it) {
    if (threadIdx.x == 0 && cb + 1 < cb_hi) band_issue<ROWB>(ring, rhs, a.n, bn, cb + 1, it + 1, stage_bytes);
    __syncwarp();
    const int st = it % kBandStages;
    mbar_wait(ring.full + st, (it / kBandStages) & 1);
    const unsigned char* const tile = ring.ring + (size_t)st * stage_bytes + sub * 16;
    const int blk_beg = cb * bn, blk_end = blk_beg + bn;
#pragma unroll
    for (int i = 0; i < RPW; ++i) {
      while (pos[i] < endp[i]) {   // warp-uniform
        const int c = cpre[i];
        const bool in = c < blk_end;
        const int cnt = __popc(__ballot_sync(kFullMask, in));
        if (cnt == 0) break;
        const unsigned myoff = in ? (unsigned)(c - blk_beg) * (unsigned)ROWB : 0u;   // byte offset of the row in the tile
        const int base = pos[i];
        pos[i] += cnt;
        cpre[i] = pos[i] + lane < endp[i] ? __ldg(cols + pos[i] + lane) : 0x7fffffff;   // next visit's batch
        // (the shuffles stay outside the count test: a collective under a branch the compiler cannot
        // prove uniform costs a WARPSYNC + BSSY/BSYNC pair per step)
        unsigned og[8];
#pragma unroll
        for (int g = 0; g < 8; ++g) og[g] = __shfl_sync(kFullMask, myoff, 4 * g + q);
        float acc[8];
#pragma unroll
        for (int g = 0; g < 8; ++g) {
          acc[g] = 0.f;
          if (4 * g < cnt) {   // warp-uniform
            const unsigned o = og[g];
            float s0 = 0.f, s1 = 0.f;   // even / odd elements: packed fma.rn.f32x2, half the issue slots
#pragma unroll
            for (int cc = 0; cc < KC; ++cc) {
              float x[E];
              unpack_raw<T, E>(x, *reinterpret_cast<const uint4*>(tile + o + cc * 128));
#pragma unroll
              for (int e = 0; e < E; e += 2) ffma2(s0, s1, lv[i][cc][e], lv[i][cc][e + 1], x[e], x[e + 1]);
            }
            acc[g] = s0 + s1;
          }
        }
        transpose_reduce<8>(acc, sub);   // lane (q, sub) now holds nonzero 4 * sub + q of the batch
        const int j = 4 * sub + q;
        if (j < cnt) out[base + j] = Elem<T>::from_f(acc[0]);
        if (cnt < 32) break;
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(ring.empty + st);
  }
}

// -------------------------------------------------------------------------------------- SpMM
template <typename T, int KC, int WARPS, int RPW>
__global__ void __launch_bounds__(WARPS * 32, 1)
spmm_band_kernel(const SpmmArgsT<T> a, const StagedGuard guard, const int bn) {
  using G = BandGeom<T, KC>;
  constexpr int E = G::E, ROWB = G::ROWB;
  if (staged_declines(guard)) return;
  extern __shared__ unsigned char smem_dyn[];
  const int stage_bytes = bn * ROWB;
  const BandRing ring = band_setup<WARPS>(smem_dyn, stage_bytes);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q = lane >> 3, sub = lane & 7;
  const int rep = blockIdx.y;
  const int item0 = blockIdx.x * (WARPS * RPW) + warp * RPW;
  const int* __restrict__ cols = a.column_indices;
  const T* __restrict__ vals = a.values + (size_t)rep * a.values_stride;

  // cpre / vpre: columns and values of nonzeros pos[i] + lane, fetched one visit ahead (see sddmm_band)
  int rows[RPW], pos[RPW], endp[RPW], cpre[RPW];
  float vpre[RPW];
  float acc[RPW][KC][E];
  int lo = 0x7fffffff, hi = -1;
  auto fetch = [&](int i) {
    const int idx = pos[i] + lane;
    cpre[i] = 0x7fffffff;
    vpre[i] = 0.f;
    if (idx < endp[i]) {
      cpre[i] = __ldg(cols + idx);
      vpre[i] = Elem<T>::to_f(vals[a.value_perm ? __ldg(a.value_perm + idx) : idx]);
    }
  };
#pragma unroll
  for (int i = 0; i < RPW; ++i) {
    rows[i] = -1;
    pos[i] = endp[i] = 0;
    cpre[i] = 0x7fffffff;
    vpre[i] = 0.f;
    if (item0 + i < a.m) {
      rows[i] = a.row_indices ? __ldg(a.row_indices + item0 + i) : item0 + i;
      pos[i] = __ldg(a.row_offsets + rows[i]);
      endp[i] = __ldg(a.row_offsets + rows[i] + 1);
      fetch(i);
      if (pos[i] < endp[i]) {
        lo = min(lo, __ldg(cols + pos[i]));
        hi = max(hi, __ldg(cols + endp[i] - 1));
      }
    }
#pragma unroll
    for (int cc = 0; cc < KC; ++cc)
#pragma unroll
      for (int e = 0; e < E; ++e) acc[i][cc][e] = 0.f;
  }
  __syncthreads();
  if (lane == 0 && hi >= 0) {
    atomicMin(ring.range, lo / bn);
    atomicMax(ring.range + 1, hi / bn + 1);
  }
  __syncthreads();
  const int cb_lo = ring.range[0], cb_hi = ring.range[1];

  const unsigned char* const dense = reinterpret_cast<const unsigned char*>(a.dense + (size_t)rep * a.k * G::D);
  if (threadIdx.x == 0 && cb_lo < cb_hi) band_issue<ROWB>(ring, dense, a.k, bn, cb_lo, 0, stage_bytes);

  for (int cb = cb_lo, it = 0; cb < cb_hi; ++cb, ++it) {
    if (threadIdx.x == 0 && cb + 1 < cb_hi) band_issue<ROWB>(ring, dense, a.k, bn, cb + 1, it + 1, stage_bytes);
    __syncwarp();
    const int st = it % kBandStages;
    mbar_wait(ring.full + st, (it / kBandStages) & 1);
    const unsigned char* const tile = ring.ring + (size_t)st * stage_bytes + sub * 16;
    const int blk_beg = cb * bn, blk_end = blk_beg + bn;
#pragma unroll
    for (int i = 0; i < RPW; ++i) {
      while (pos[i] < endp[i]) {   // warp-uniform
        const int c = cpre[i];
        const bool in = c < blk_end;
        const int cnt = __popc(__ballot_sync(kFullMask, in));
        if (cnt == 0) break;
        const unsigned myoff = in ? (unsigned)(c - blk_beg) * (unsigned)ROWB : 0u;
        const float myval = vpre[i];
        pos[i] += cnt;
        fetch(i);   // next visit's batch
#pragma unroll
        for (int g = 0; g < 8; ++g) {
          if (4 * g < cnt) {   // warp-uniform
            const unsigned o = __shfl_sync(kFullMask, myoff, 4 * g + q);
            const float v = __shfl_sync(kFullMask, myval, 4 * g + q);
            if (4 * g + q < cnt) {   // a quarter past the end of the visit multiplies nothing (no 0 * Inf)
#pragma unroll
              for (int cc = 0; cc < KC; ++cc) {
                float x[E];
                unpack_raw<T, E>(x, *reinterpret_cast<const uint4*>(tile + o + cc * 128));
#pragma unroll
                for (int e = 0; e < E; e += 2) ffma2(acc[i][cc][e], acc[i][cc][e + 1], v, v, x[e], x[e + 1]);
              }
            }
          }
        }
        if (cnt < 32) break;
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(ring.empty + st);
  }

  // ---- epilogue: sum the four quarters' partial rows, bias, ReLU, one 128-byte store per chunk ----
#pragma unroll
  for (int i = 0; i < RPW; ++i) {
    if (rows[i] < 0) continue;   // warp-uniform
    const float bias = a.bias ? __ldg(a.bias + rows[i]) : 0.f;
    T* const orow = a.out + ((size_t)rep * a.m + rows[i]) * G::D + sub * E;
#pragma unroll
    for (int cc = 0; cc < KC; ++cc) {
      float t[E];
#pragma unroll
      for (int e = 0; e < E; ++e) {
        float x = acc[i][cc][e];
        x += __shfl_xor_sync(kFullMask, x, 8);
        x += __shfl_xor_sync(kFullMask, x, 16);
        x += bias;
        t[e] = a.relu ? fmaxf(x, 0.f) : x;
      }
      if (q == (cc & 3)) store_elems<T, E>(orow + cc * 8 * E, t);
    }
  }
}

// ------------------------------------------------------------------------------------- host
// Mean nonzeros a warp finds per (row, block) visit; below ~8 the visits are mostly bookkeeping and
// the row kernels win.
inline double band_visit_population(int nnz, int m, int cols, int bn) {
  return m > 0 && cols > 0 ? (double)nnz / m * (double)min(bn, cols) / cols : 0.0;
}

template <typename T>
inline int band_kc(int d) {   // 16-byte-chunk rows per 128 bytes: 0 = not a band shape
  constexpr int chunk = 8 * Elem<T>::kPer16;
  return (d % chunk == 0 && d / chunk >= 1 && d / chunk <= 4) ? d / chunk : 0;
}

// Measured on B200 at the C4 attention shape (profiles/r02_band_ab.txt): the fp32 SDDMM band kernel
// beats the L1-gather rows kernel by 8 % (2.76 vs 2.99 ms), every other band variant loses to its
// L1-gather counterpart -- at 64 FMAs per nonzero the per-visit bookkeeping costs as many issue
// slots as the arithmetic.  So the heuristic takes the band kernel for fp32 SDDMM only; the knob
// (variant 6) forces it anywhere the shape allows.
template <typename T>
inline bool sddmm_band_applicable(const SddmmArgsT<T>& a, int replicas) {
  if (sizeof(T) != 4) return false;
  const int kc = band_kc<T>(a.k);
  if (kc == 0 || !aligned16(a.lhs) || !aligned16(a.rhs)) return false;
  if (a.m < 64 || a.n < 64) return false;
  const int bn = kBandStageBytes / (128 * kc);
  if (band_visit_population(a.nnz, a.m, a.n, bn) < 8.0) return false;
  return 2.0 * a.nnz * a.k * replicas >= 2.0e7;
}

template <typename T>
inline bool spmm_band_applicable(const SpmmArgsT<T>& a, int replicas) {
  if (true) return false;   // never by heuristic (see above)
  const int kc = band_kc<T>(a.n);
  if (kc == 0 || !aligned16(a.dense) || !aligned16(a.out)) return false;
  if (a.m < 64 || a.k < 64) return false;
  const int bn = kBandStageBytes / (128 * kc);
  if (band_visit_population(a.nnz, a.m, a.k, bn) < 8.0) return false;
  return 2.0 * a.nnz * a.n * replicas >= 2.0e7;
}

template <class Kernel>
inline int band_configure(Kernel kernel, size_t smem, bool* configured, const char* what) {
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 0 && dev < 64 && !configured[dev]) {
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return cuda_fail(e, what);
    configured[dev] = true;
  }
  return 0;
}

// Rows per warp for a WARPS-warp CTA: as many as keep the per-row register state (lhs slices /
// accumulators: KC * E floats per row and lane) within 64 registers (32 for a 32-warp CTA, whose
// threads get 64 registers in all), at most 8.
template <typename T, int KC, int WARPS>
constexpr int band_rpw() {
  constexpr int per_row = KC * Elem<T>::kPer16;
  constexpr int budget = WARPS > 16 ? 32 : 64;
  return per_row * 8 <= budget ? 8 : (per_row * 4 <= budget ? 4 : (per_row * 2 <= budget ? 2 : 1));
}
extern std::atomic<int> g_band_warps;   // benchmark knob "band_warps": 16 (default) or 32 warps per CTA

template <typename T, int KC, int WARPS>
inline int launch_sddmm_band_kc(const SddmmArgsT<T>& a, int replicas, cudaStream_t stream, const StagedGuard& guard) {
  using G = BandGeom<T, KC>;
  static bool configured[64] = {false};
  const size_t max_smem = G::smem_bytes(kBandStageBytes / G::ROWB);
  constexpr int RPW = band_rpw<T, KC, WARPS>();
  auto kernel = sddmm_band_kernel<T, KC, WARPS, RPW>;
  if (int rc = band_configure(kernel, max_smem, configured, "tsp_sddmm (band smem attribute)")) return rc;
  const int bn = G::block_rows(a.n);
  dim3 grid(ceil_div(a.m, WARPS * RPW), replicas);
  kernel<<<grid, WARPS * 32, G::smem_bytes(bn), stream>>>(a, guard, bn);
  return 0;
}

template <typename T>
inline int launch_sddmm_band(const SddmmArgsT<T>& a, int replicas, cudaStream_t stream, const StagedGuard& guard) {
  const bool wide = g_band_warps.load() == 32;
  switch (band_kc<T>(a.k)) {
    case 1: return wide ? launch_sddmm_band_kc<T, 1, 32>(a, replicas, stream, guard) : launch_sddmm_band_kc<T, 1, 16>(a, replicas, stream, guard);
    case 2: return wide ? launch_sddmm_band_kc<T, 2, 32>(a, replicas, stream, guard) : launch_sddmm_band_kc<T, 2, 16>(a, replicas, stream, guard);
    case 3: return launch_sddmm_band_kc<T, 3, 16>(a, replicas, stream, guard);
    default: return launch_sddmm_band_kc<T, 4, 16>(a, replicas, stream, guard);
  }
}

template <typename T, int KC, int WARPS>
inline int launch_spmm_band_kc(const SpmmArgsT<T>& a, int replicas, cudaStream_t stream, const StagedGuard& guard) {
  using G = BandGeom<T, KC>;
  static bool configured[64] = {false};
  const size_t max_smem = G::smem_bytes(kBandStageBytes / G::ROWB);
  constexpr int RPW = band_rpw<T, KC, WARPS>();
  auto kernel = spmm_band_kernel<T, KC, WARPS, RPW>;
  if (int rc = band_configure(kernel, max_smem, configured, "tsp_spmm (band smem attribute)")) return rc;
  const int bn = G::block_rows(a.k);
  dim3 grid(ceil_div(a.m, WARPS * RPW), replicas);
  kernel<<<grid, WARPS * 32, G::smem_bytes(bn), stream>>>(a, guard, bn);
  return 0;
}

template <typename T>
inline int launch_spmm_band(const SpmmArgsT<T>& a, int replicas, cudaStream_t stream, const StagedGuard& guard) {
  const bool wide = g_band_warps.load() == 32;
  switch (band_kc<T>(a.n)) {
    case 1: return wide ? launch_spmm_band_kc<T, 1, 32>(a, replicas, stream, guard) : launch_spmm_band_kc<T, 1, 16>(a, replicas, stream, guard);
    case 2: return wide ? launch_spmm_band_kc<T, 2, 32>(a, replicas, stream, guard) : launch_spmm_band_kc<T, 2, 16>(a, replicas, stream, guard);
    case 3: return launch_spmm_band_kc<T, 3, 16>(a, replicas, stream, guard);
    default: return launch_spmm_band_kc<T, 4, 16>(a, replicas, stream, guard);
  }
}

}  // namespace tsp
