// sddmm_tiled3.cuh -- SDDMM, dense register tiles with K-MAJOR operand tiles in shared memory.
//
// Same job as sddmm_tiled2_kernel (sddmm_tiled.cuh: every product of a [128 x 64] output tile, sampled at
// the mask through the bitmap view) with the operand arrangement of a classic SIMT SGEMM "TN" kernel:
//  * both operands lie K-contiguous in memory (lhs [m, k], rhs [n, k]); a thread fetches 16-byte pieces of
//    their rows with LDG.128 one k-step ahead and stores them TRANSPOSED, As[k][row] / Bs[k][col] (an XOR
//    swizzle of the row index by the k-chunk keeps both the 4-byte transposing stores and the 16-byte reads
//    free of bank conflicts without padding);
//  * in the inner loop a thread reads, per k, 8 lhs values (2 LDS.128) and 8 rhs values (2 LDS.128) and
//    issues 32 FFMA2: the two halves of an accumulator are two ADJACENT OUTPUT COLUMNS (the rhs pair comes
//    straight out of the 16-byte load), the lhs factor is FFMA2's scalar-broadcast operand -- 64 accumulator
//    registers for 64 outputs, no register shuffling.  sddmm_tiled2's even-k / odd-k pairing needs 128
//    accumulator registers and sits at 255 registers per thread;
//  * one __syncthreads per k-step (double-buffered tiles), tiles walked in the L2-aware order of raster_tile.
// Each output is ONE fp32 FMA chain in ascending k (the reference's order, sddmm/cuda_sddmm.cu.cc:117-139).
// cuBLAS's SIMT SGEMM for the same contraction (A @ B^T, 4096^3) takes 2.40 ms on this GPU
// (tools/ubench/sgemm_ceiling.py), sddmm_tiled2 2.76 ms, this kernel 2.83 ms: the synchronous LDG -> STS staging
// and the barrier per k-step cost what the leaner accumulators win.  Kept as a measured alternative behind the
// benchmark knob "sddmm_tiled3"; the dispatcher does not pick it.
#pragma once
#include "sddmm_tiled.cuh"

namespace tsp {

template <int BK_>
struct SddmmTiled3 {
  static constexpr int BM = 128, BN = 64, BK = BK_, THREADS = 128;
  static constexpr int KV = BK_ / 4;                         // 16-byte pieces per operand row and k-step
  static constexpr int SHIFT = KV == 4 ? 3 : 2;              // row-index swizzle: row ^ ((k / 4 % KV) << SHIFT)
  static constexpr int A_V4 = BM * KV / THREADS;             // LDG.128 per thread and k-step, lhs
  static constexpr int B_V4 = BN * KV / THREADS;             // ... rhs
  static constexpr size_t kSmemBytes = sizeof(float) * 2 * BK_ * (BM + BN) + BM * 8 + BM * 4 + 64 * 4;
};

template <int BK, int MINB>
__global__ void __launch_bounds__(128, MINB)
sddmm_tiled3_kernel(const SddmmArgs a, const StagedGuard guard, int col_blocks, int total_tiles,
                    const unsigned long long* __restrict__ masks, const int* __restrict__ bases) {
  using C = SddmmTiled3<BK>;
  constexpr int BM = C::BM, BN = C::BN, KV = C::KV, SHIFT = C::SHIFT, A_V4 = C::A_V4, B_V4 = C::B_V4;
  if (staged_declines(guard)) return;

  extern __shared__ __align__(16) unsigned char smem_t3[];
  float* const As = reinterpret_cast<float*>(smem_t3);                 // [2][BK][BM]
  float* const Bs = As + 2 * BK * BM;                                   // [2][BK][BN]
  unsigned long long* const rowmask = reinterpret_cast<unsigned long long*>(Bs + 2 * BK * BN);
  int* const rowbase = reinterpret_cast<int*>(rowmask + BM);
  int* const s_tile = rowbase + BM;                                     // [64]

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  // 8 x 4 threads per warp, lane bits as in sddmm_tiled2: every lane quad reads <= 2 distinct 16-byte pieces
  const int ty = (lane & 1) | ((lane >> 1) & 6), tx = ((lane >> 1) & 1) | ((lane >> 3) & 2);
  const int wr = warp >> 1, wc = warp & 1;          // 2 x 2 warps: 64 x 32 outputs each
  const int rep = blockIdx.y;
  const int ksteps1 = (a.k + BK - 1) / BK;
  const int nsteps = a.reduce_reps > 0 ? ksteps1 * a.reduce_reps : ksteps1;

  constexpr int SRB = 16, SCB = 16, kMaxTilesPerCta = 64;
  const int G = gridDim.x;
  const int ntiles = (int)blockIdx.x < total_tiles ? (total_tiles - (int)blockIdx.x + G - 1) / G : 0;
  if (ntiles == 0) return;
  if (tid < min(ntiles, kMaxTilesPerCta)) {
    int rb, cb;
    raster_tile((int)blockIdx.x + tid * G, total_tiles / col_blocks, col_blocks, SRB, SCB, rb, cb);
    s_tile[tid] = (rb << 16) | cb;
  }
  __syncthreads();

  float* __restrict__ out = a.out + (size_t)rep * a.nnz;
  // global -> register staging: piece f = tid + 128 * q of the step's [rows][KV] grid of 16-byte pieces
  const int ld_row = tid / KV, ld_kc = tid % KV;      // (+ 128 / KV rows per q)
  const int sw_st = ld_kc << SHIFT;

  for (int ti = 0; ti < ntiles; ++ti) {
    const int rb = s_tile[ti] >> 16, cb = s_tile[ti] & 0xffff;
    unsigned long long my_mask = 0ull;
    int my_base = 0;
    if (rb * BM + tid < a.m) {   // THREADS == BM
      const size_t e = (size_t)(rb * BM + tid) * col_blocks + cb;
      my_mask = __ldg(masks + e);
      my_base = __ldg(bases + e);
    }

    // Global addresses of the k-step's 16-byte pieces: two running pointers (advanced by BK per step, by the rest of
    // the replica when the replica loop folded into the k loop wraps), a constant stride between a thread's
    // pieces, and a per-tile bit per piece for rows / columns beyond the matrix -- recomputing
    // ((replica * rows + row) * k + kk) per piece cost ~300 integer instructions per step and warp, a third of
    // the warp's time (ncu source view).
    float4 ga[A_V4], gb[B_V4];
    unsigned ok_a = 0, ok_b = 0;
#pragma unroll
    for (int q = 0; q < A_V4; ++q) ok_a |= (rb * BM + ld_row + q * (128 / KV) < a.m ? 1u : 0u) << q;
#pragma unroll
    for (int q = 0; q < B_V4; ++q) ok_b |= (cb * BN + ld_row + q * (128 / KV) < a.n ? 1u : 0u) << q;
    const int rep0 = a.reduce_reps > 0 ? 0 : rep;
    const float* pa = a.lhs + ((size_t)rep0 * a.m + (rb * BM + ld_row)) * a.k + ld_kc * 4;
    const float* pb = a.rhs + ((size_t)rep0 * a.n + (cb * BN + ld_row)) * a.k + ld_kc * 4;
    const long long stride = (long long)(128 / KV) * a.k;
    const long long wrap_a = (long long)a.m * a.k - (long long)(ksteps1 - 1) * BK;
    const long long wrap_b = (long long)a.n * a.k - (long long)(ksteps1 - 1) * BK;
    int ks_next = 0;   // k-step (within the replica) the next load_global fetches
    auto load_global = [&]() {
      const bool kin = ks_next * BK + ld_kc * 4 < a.k;
#pragma unroll
      for (int q = 0; q < A_V4; ++q)
        ga[q] = (kin && ((ok_a >> q) & 1u)) ? ldg_f4(pa + q * stride) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int q = 0; q < B_V4; ++q)
        gb[q] = (kin && ((ok_b >> q) & 1u)) ? ldg_f4(pb + q * stride) : make_float4(0.f, 0.f, 0.f, 0.f);
      if (++ks_next == ksteps1) {
        ks_next = 0;
        pa += wrap_a;
        pb += wrap_b;
      } else {
        pa += BK;
        pb += BK;
      }
    };
    auto store_smem = [&](int buf) {
      float* const as = As + buf * BK * BM + (ld_kc * 4) * BM;
      float* const bs = Bs + buf * BK * BN + (ld_kc * 4) * BN;
#pragma unroll
      for (int q = 0; q < A_V4; ++q) {
        const int r = (ld_row + q * (128 / KV)) ^ sw_st;
        as[r] = ga[q].x; as[BM + r] = ga[q].y; as[2 * BM + r] = ga[q].z; as[3 * BM + r] = ga[q].w;
      }
#pragma unroll
      for (int q = 0; q < B_V4; ++q) {
        const int c = (ld_row + q * (128 / KV)) ^ sw_st;
        bs[c] = gb[q].x; bs[BN + c] = gb[q].y; bs[2 * BN + c] = gb[q].z; bs[3 * BN + c] = gb[q].w;
      }
    };

    unsigned long long acc[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = 0ull;

    load_global();
    store_smem(0);
    __syncthreads();
    const int ra0 = wr * 64 + ty * 4, cb0 = wc * 32 + tx * 4;
    for (int step = 0; step < nsteps; ++step) {
      if (step + 1 < nsteps) load_global();
      const float* const as = As + (step & 1) * BK * BM;
      const float* const bs = Bs + (step & 1) * BK * BN;
#pragma unroll
      for (int k = 0; k < BK; ++k) {
        const int sw = ((k >> 2) % KV) << SHIFT;
        const float4 a0 = *reinterpret_cast<const float4*>(as + k * BM + (ra0 ^ sw));
        const float4 a1 = *reinterpret_cast<const float4*>(as + k * BM + ((ra0 + 32) ^ sw));
        const ulonglong2 b0 = *reinterpret_cast<const ulonglong2*>(bs + k * BN + (cb0 ^ sw));
        const ulonglong2 b1 = *reinterpret_cast<const ulonglong2*>(bs + k * BN + ((cb0 + 16) ^ sw));
        const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const unsigned long long ap = pack2(av[i], av[i]);   // FFMA2's scalar-broadcast operand
          fma2(acc[i][0], ap, b0.x);
          fma2(acc[i][1], ap, b0.y);
          fma2(acc[i][2], ap, b1.x);
          fma2(acc[i][3], ap, b1.y);
        }
      }
      if (step + 1 < nsteps) store_smem((step + 1) & 1);
      __syncthreads();
    }

    rowmask[tid] = my_mask;
    rowbase[tid] = my_base;
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int rl = wr * 64 + (i >> 2) * 32 + ty * 4 + (i & 3);
      const unsigned long long mk = rowmask[rl];
      const int base = rowbase[rl];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int cl = wc * 32 + (j >> 2) * 16 + tx * 4 + (j & 3);
        if ((mk >> cl) & 1ull) {
          float lo, hi;
          unpack2(acc[i][j >> 1], lo, hi);
          out[base + __popcll(mk & ((1ull << cl) - 1ull))] = (j & 1) ? hi : lo;
        }
      }
    }
    __syncthreads();  // rowmask and the operand tiles are rewritten by the next tile
  }
}

extern std::atomic<int> g_sddmm_tiled3;   // benchmark knob "sddmm_tiled3": see launch_sddmm_tiled3

template <int BK, int MINB>
inline int launch_sddmm_tiled3_cfg(const SddmmArgs& a, int replicas, cudaStream_t stream, const StagedGuard& guard,
                                   const void* view) {
  using C = SddmmTiled3<BK>;
  static bool configured[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 0 && dev < 64 && !configured[dev]) {
    cudaError_t e = cudaFuncSetAttribute(sddmm_tiled3_kernel<BK, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)C::kSmemBytes);
    if (e != cudaSuccess) return cuda_fail(e, "tsp_sddmm_f32 (tiled3 smem attribute)");
    configured[dev] = true;
  }
  const int row_blocks = ceil_div(a.m, C::BM), col_blocks = ceil_div(a.n, C::BN);
  const long long total = (long long)row_blocks * col_blocks;
  const int slots = MINB * sm_count();
  const int grid_reps = a.reduce_reps > 0 ? 1 : replicas;
  int per_cta = (int)((total * grid_reps + slots - 1) / slots);
  per_cta = per_cta < 1 ? 1 : (per_cta > col_blocks ? col_blocks : per_cta);
  if (per_cta > 64) per_cta = 64;
  const int ctas = (int)((total + per_cta - 1) / per_cta);
  const size_t entries = (size_t)a.m * col_blocks;
  const unsigned long long* masks = static_cast<const unsigned long long*>(view);
  const int* bases = reinterpret_cast<const int*>(masks + entries);
  sddmm_tiled3_kernel<BK, MINB><<<dim3(ctas, grid_reps), 128, C::kSmemBytes, stream>>>(a, guard, col_blocks, (int)total,
                                                                                      masks, bases);
  TSP_CHECK_LAUNCH("tsp_sddmm_f32 (tiled3)");
  return 0;
}

// knob "sddmm_tiled3": 1 = 16-deep k-steps, 2 = 32-deep k-steps (both 2 CTAs per SM; a 3-CTA build at <= 168
// registers spills and is slower: 3.24 / 3.02 ms).  Measured at 4096^3: 3.09 / 2.83 ms against sddmm_tiled2's
// 2.76 ms -- an opt-in variant, not the default.
inline int launch_sddmm_tiled3(const SddmmArgs& a, int replicas, cudaStream_t stream, const StagedGuard& guard,
                               const void* view, int cfg) {
  if (cfg == 1) return launch_sddmm_tiled3_cfg<16, 2>(a, replicas, stream, guard, view);
  return launch_sddmm_tiled3_cfg<32, 2>(a, replicas, stream, guard, view);
}

}  // namespace tsp
