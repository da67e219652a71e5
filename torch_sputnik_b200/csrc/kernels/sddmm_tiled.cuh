// sddmm_tiled.cuh -- SDDMM, variant 2 ("tiled"): dense register-tiled products, sampled at the mask.
//
// out[j] = <lhs[row_of(j), :], rhs[col(j), :]>.  A gather SDDMM (the reference's design,
// sddmm/cuda_sddmm.cu.cc:34-165, and our variant 1) re-reads a K-long rhs row per nonzero:
// nnz*K*4 bytes through L2 (137 GB at 4096^3, d = 0.5) and one 16-byte load per 4 FMAs.  Once a
// mask is denser than ~20 % it is cheaper to compute EVERY product of a [128 x 64] output tile
// with an fp32 outer-product micro-kernel (1 byte of shared-memory traffic per FMA instead of
// 4, operands crossing L2 once per tile) and keep only the sampled ones:
//
//  * a CTA owns a run of consecutive [128 rows x 64 cols] tiles of one row block; per 32-deep
//    k-step warp 0 streams the 128 lhs rows and 64 rhs rows (128 B each, K-contiguous exactly as
//    they lie in memory) into a 4-stage ring with bulk async copies + mbarrier transactions;
//  * 256 threads, 8 x 4 outputs each (rows i*16+ty, cols j*16+tx -> conflict-free LDS.128 along
//    k); the FMA stream is fma.rn.f32x2 pairing EVEN/ODD k of the same output, so both packed
//    operands come straight out of the 16-byte loads and no register shuffling is needed.  An
//    output is (sum over even k) + (sum over odd k), each summed in ascending k;
//  * a pre-pass (csr_block_masks_kernel, one warp per row) turns the CSR pattern into a bitmap
//    view -- per (row, 64-column block) a 64-bit occupancy word + the flat index of its first
//    nonzero -- in stream-ordered scratch; the tile epilogue writes acc to
//    out[base + popc(mask below col)] for the occupied positions only.
//
// Precondition: ascending columns per row (checked on device, gather fallback otherwise),
// k % 4 == 0 and 16-byte aligned operands.
#pragma once
#include <cuda.h>  // CUtensorMap (types only; the encoder is fetched through the runtime)

#include "sddmm_gather.cuh"
#include "spmm_staged.cuh"

namespace tsp {

struct SddmmTiled {
  static constexpr int BM = 128, BN = 64, BK = 32;  // BK floats = 128 B = one swizzle-128B row
  static constexpr int STAGES = 4, THREADS = 256;
  static constexpr int kStageFloats = (BM + BN) * BK;
  static constexpr unsigned kStageBytes = kStageFloats * 4;
  static constexpr size_t kSmemBytes =
      sizeof(float) * STAGES * kStageFloats + BM * 8 + BM * 4 + 2 * STAGES * 8 + 1024;
};


// Bitmap view of the mask for the tiled kernel: for every (row, 64-column block) a 64-bit
// occupancy word and the flat index of the block's first nonzero.  One warp per row walks the
// row's (ascending) column indices 32 at a time; the same pass raises the unsorted flag.
// masks must be zero on entry.
__global__ void __launch_bounds__(256)
csr_block_masks_kernel(const int* __restrict__ row_offsets, const int* __restrict__ cols, int m,
                       int col_blocks, unsigned long long* __restrict__ masks, int* __restrict__ bases,
                       unsigned* __restrict__ flag) {
  const int row = (blockIdx.x * 256 + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (row >= m) return;
  const int beg = __ldg(row_offsets + row), end = __ldg(row_offsets + row + 1);
  bool bad = false;
  for (int j0 = beg; j0 < end; j0 += 32) {
    const int j = j0 + lane;
    const bool in = j < end;
    const int c = in ? __ldg(cols + j) : 0;
    if (in && j + 1 < end) bad |= c >= __ldg(cols + j + 1);
    const int cb = c >> 6;
    unsigned active = __ballot_sync(kFullMask, in);
    while (active) {
      const int leader = __ffs(active) - 1;
      const int cb0 = __shfl_sync(kFullMask, cb, leader);
      const bool sel = in && cb == cb0;
      const unsigned lo = __reduce_or_sync(kFullMask, (sel && (c & 63) < 32) ? 1u << (c & 31) : 0u);
      const unsigned hi = __reduce_or_sync(kFullMask, (sel && (c & 63) >= 32) ? 1u << (c & 31) : 0u);
      if (lane == leader) {
        const size_t e = (size_t)row * col_blocks + cb0;
        const unsigned long long old = masks[e];
        if (old == 0ull) bases[e] = j;  // first nonzero of this block (blocks arrive in order)
        masks[e] = old | ((unsigned long long)hi << 32) | lo;
      }
      active &= ~__ballot_sync(kFullMask, sel);
    }
  }
  if (flag != nullptr && __any_sync(kFullMask, bad) && lane == 0) *flag = 1u;
}

__global__ void __launch_bounds__(SddmmTiled::THREADS, 2)
sddmm_tiled_kernel(const SddmmArgs a, const StagedGuard guard, int tiles_per_cta, int col_blocks,
                   int total_tiles, const unsigned long long* __restrict__ masks,
                   const int* __restrict__ bases, const __grid_constant__ CUtensorMap tm_lhs,
                   const __grid_constant__ CUtensorMap tm_rhs) {
  using C = SddmmTiled;
  constexpr int BM = C::BM, BN = C::BN, BK = C::BK, STAGES = C::STAGES;
  if (staged_declines(guard)) return;

  extern __shared__ unsigned char smem_dyn[];
  // swizzle-128B tiles want 1024-byte aligned destinations
  unsigned char* const smem_raw = smem_dyn + ((1024u - (smem_u32(smem_dyn) & 1023u)) & 1023u);
  float* const ring = reinterpret_cast<float*>(smem_raw);
  unsigned long long* const rowmask = reinterpret_cast<unsigned long long*>(ring + STAGES * C::kStageFloats);
  int* const rowbase = reinterpret_cast<int*>(rowmask + BM);
  unsigned long long* const full = reinterpret_cast<unsigned long long*>(rowbase + BM);
  unsigned long long* const empty = full + STAGES;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  // 8 x 4 threads per warp (every half-warp touches 4 distinct 16-byte chunks per operand: 2
  // wavefronts per LDS.128), 2 x 4 warps: a warp owns 64 rows x 16 columns.
  // Lane -> (ty, tx) = (bits 0,2,3 ; bits 1,4): every aligned lane quad then reads at most two
  // distinct 16-byte chunks of either operand, which is what makes an LDS.128 cost 2 wavefronts
  // instead of 4 (tools/ubench/ubench_lds.cu patterns 10-16).
  const int ty = (lane & 1) | ((lane >> 1) & 6), tx = ((lane >> 1) & 1) | ((lane >> 3) & 2);
  const int wr = warp >> 2, wc = warp & 3;
  const int rep = blockIdx.y;
  const int K = a.k;
  const int ksteps = (K + BK - 1) / BK;
  const int t0 = blockIdx.x * tiles_per_cta;
  const int t1 = min(total_tiles, t0 + tiles_per_cta);
  if (t0 >= t1) return;
  const int Q = (t1 - t0) * ksteps;  // (tile, k-step) pairs this CTA consumes, in order

  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full + s, 1);
      mbar_init(empty + s, C::THREADS / 32);
    }
    mbar_fence_init();
  }
  __syncthreads();

  float* __restrict__ out = a.out + (size_t)rep * a.nnz;

  // Producer duty (one lane of warp 0): two TMA tile loads fill ring slot of pair q.  Rows past
  // m / n and the k tail are zero-filled by the TMA unit, so a stage always carries kStageBytes.
  auto issue = [&](int q) {
    const int st = q % STAGES, use = q / STAGES;
    if (use > 0) mbar_wait(empty + st, (use - 1) & 1);
    if (lane == 0) {
      const int t = t0 + q / ksteps, ks = q % ksteps;
      const int rb = t / col_blocks, cb = t % col_blocks;
      float* const sa = ring + st * C::kStageFloats;
      mbar_arrive_expect_tx(full + st, C::kStageBytes);
      tma_load_3d(sa, &tm_lhs, ks * BK, rb * BM, rep, full + st);
      tma_load_3d(sa + BM * BK, &tm_rhs, ks * BK, cb * BN, rep, full + st);
    }
    __syncwarp();
  };
  if (warp == 0)
    for (int q = 0; q < min(Q, STAGES - 1); ++q) issue(q);

  int q = 0;
  for (int t = t0; t < t1; ++t) {
    const int rb = t / col_blocks, cb = t % col_blocks;
    // Thread r < BM fetches (occupancy, first index) of row r for this tile now; the values are
    // not touched until the epilogue, so the load latency hides under the k loop.
    unsigned long long my_mask = 0ull;
    int my_base = 0;
    if (tid < BM && rb * BM + tid < a.m) {
      const size_t e = (size_t)(rb * BM + tid) * col_blocks + cb;
      my_mask = __ldg(masks + e);
      my_base = __ldg(bases + e);
    }

    unsigned long long acc[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = 0ull;

    for (int ks = 0; ks < ksteps; ++ks, ++q) {
      if (warp == 0 && q + STAGES - 1 < Q) issue(q + STAGES - 1);
      const int st = q % STAGES;
      mbar_wait(full + st, (q / STAGES) & 1);
      // swizzle-128B: 16-byte chunk c of row r lives at chunk (c ^ (r & 7)); our rows are
      // i*16+ty and j*16+tx, so the XOR term is a per-thread constant.
      const char* const sa = reinterpret_cast<const char*>(ring + st * C::kStageFloats) + (wr * 64 + ty) * 128;
      const char* const sb = reinterpret_cast<const char*>(ring + st * C::kStageFloats + BM * BK) + (wc * 16 + tx) * 128;
#pragma unroll
      for (int kk = 0; kk < BK / 4; ++kk) {
        const int ka = (kk ^ ty) << 4, kb0 = (kk ^ tx) << 4, kb1 = (kk ^ (4 + tx)) << 4;
        ulonglong2 bq[4];
#pragma unroll
        for (int j = 0; j < 4; ++j)
          bq[j] = *reinterpret_cast<const ulonglong2*>(sb + j * 4 * 128 + ((j & 1) ? kb1 : kb0));
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const ulonglong2 aq = *reinterpret_cast<const ulonglong2*>(sa + i * 8 * 128 + ka);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            fma2(acc[i][j], aq.x, bq[j].x);
            fma2(acc[i][j], aq.y, bq[j].y);
          }
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(empty + st);
    }

    if (tid < BM) {
      rowmask[tid] = my_mask;
      rowbase[tid] = my_base;
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int rl = wr * 64 + i * 8 + ty;
      const unsigned long long mk = rowmask[rl];
      const int base = rowbase[rl];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int cl = wc * 16 + j * 4 + tx;
        if ((mk >> cl) & 1ull) {
          float lo, hi;
          unpack2(acc[i][j], lo, hi);
          out[base + __popcll(mk & ((1ull << cl) - 1ull))] = lo + hi;
        }
      }
    }
    __syncthreads();  // rowmask is rewritten by the next tile
  }
}


// ---------------------------------------------------------------------------------------------
// v2 micro-kernel: 128 threads, 8 x 8 outputs per thread.
//
// What bounded v1 (ncu, profiles/r01_sddmm_tiled_d050_ncu_summary.txt): with a 16 x 16 thread
// layout a warp's rhs LDS.128 touches 16 distinct rows and costs 4 shared-memory wavefronts, the
// lhs one 2 (tools/ubench/ubench_lds.cu: an LDS.128 whose half-warps each touch <= 4 distinct
// 16-byte chunks costs 2 wavefronts, anything wider 4), so 12 loads = 32 wavefronts fed 64 FFMA2
// = 32 SM-cycles of FMA pipe: the LSU data pipe and the FMA pipe saturate together at ~58 %.
// Here a warp is 8 (rows) x 4 (cols) threads -- every half-warp touches exactly 4 distinct chunks
// for either operand -- and a thread owns 8 x 8 outputs: 16 LDS.128 (32 wavefronts) per 128 FFMA2
// (64 SM-cycles), i.e. the LSU pipe at 50 % when the FMA pipe is full.
struct SddmmTiled2 {
  static constexpr int BM = 128, BN = 64, BK = 32;
  static constexpr int STAGES = 4, THREADS = 128;   // 4 x 24 KB stages, 2 CTAs per SM
  static constexpr int kStageFloats = (BM + BN) * BK;
  static constexpr unsigned kStageBytes = kStageFloats * 4;
  static constexpr size_t kSmemBytes =
      sizeof(float) * STAGES * kStageFloats + BM * 8 + BM * 4 + 2 * STAGES * 8 + 1024;
};

__global__ void __launch_bounds__(SddmmTiled2::THREADS, 2)
sddmm_tiled2_kernel(const SddmmArgs a, const StagedGuard guard, int tiles_per_cta, int col_blocks,
                    int total_tiles, const unsigned long long* __restrict__ masks,
                    const int* __restrict__ bases, const __grid_constant__ CUtensorMap tm_lhs,
                    const __grid_constant__ CUtensorMap tm_rhs) {
  using C = SddmmTiled2;
  constexpr int BM = C::BM, BN = C::BN, BK = C::BK, STAGES = C::STAGES;
  if (staged_declines(guard)) return;

  extern __shared__ unsigned char smem_dyn[];
  unsigned char* const smem_raw = smem_dyn + ((1024u - (smem_u32(smem_dyn) & 1023u)) & 1023u);
  float* const ring = reinterpret_cast<float*>(smem_raw);
  unsigned long long* const rowmask = reinterpret_cast<unsigned long long*>(ring + STAGES * C::kStageFloats);
  int* const rowbase = reinterpret_cast<int*>(rowmask + BM);
  unsigned long long* const full = reinterpret_cast<unsigned long long*>(rowbase + BM);
  unsigned long long* const empty = full + STAGES;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  // 8 x 4 threads per warp; ty = lane bits 0,2,3 and tx = bits 1,4 (see v1 above)
  const int ty = (lane & 1) | ((lane >> 1) & 6), tx = ((lane >> 1) & 1) | ((lane >> 3) & 2);
  const int wr = warp >> 1, wc = warp & 1;          // 2 x 2 warps: 64 x 32 outputs each
  const int rep = blockIdx.y;
  const int ksteps1 = (a.k + BK - 1) / BK;                                   // k-steps of one replica
  const int ksteps = a.reduce_reps > 0 ? ksteps1 * a.reduce_reps : ksteps1;   // k-steps per tile
  // Tiles are dealt round-robin: CTA i takes tiles i, i + G, i + 2G, ... (G = gridDim.x) of the
  // rasterised sequence (raster_tile: bands of 16 row blocks, groups of 16 column blocks -- 256 tiles,
  // about one wave of CTAs).  CTAs advance roughly in step, so the CTAs in flight share 16 x 128 lhs
  // rows + 16 x 64 rhs rows (48 MB at k = 4096), the groups of a band reuse its lhs rows from L2, and
  // every operand byte comes from HBM about twice -- against every CTA streaming a private row block and
  // all of rhs at once when each owned a contiguous run of tiles (dram__bytes 1261 MB for 201 MB of
  // operands at density 0.5, profiles/r01_sddmm_tiled2_d050_ncu_summary.txt).
  // (The CTA's tile coordinates are worked out ONCE, by its first threads, into shared memory: the
  // kernel sits at 255 registers and the raster arithmetic inside the tile loop spilled.)
  constexpr int SRB = 16, SCB = 16, kMaxTilesPerCta = 64;
  __shared__ int s_tile[kMaxTilesPerCta];
  (void)tiles_per_cta;
  const int G = gridDim.x;
  const int ntiles = (int)blockIdx.x < total_tiles ? (total_tiles - (int)blockIdx.x + G - 1) / G : 0;
  if (ntiles == 0) return;
  const int Q = ntiles * ksteps;
  if ((int)threadIdx.x < min(ntiles, kMaxTilesPerCta)) {
    int rb, cb;
    raster_tile((int)blockIdx.x + (int)threadIdx.x * G, total_tiles / col_blocks, col_blocks, SRB, SCB, rb, cb);
    s_tile[threadIdx.x] = (rb << 16) | cb;
  }

  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full + s, 1);
      mbar_init(empty + s, C::THREADS / 32);
    }
    mbar_fence_init();
  }
  __syncthreads();

  float* __restrict__ out = a.out + (size_t)rep * a.nnz;

  auto issue = [&](int q) {   // one warp (q & 3 of the step it runs in), all lanes converge here
    const int st = q % STAGES, use = q / STAGES;
    if (use > 0) mbar_wait(empty + st, (use - 1) & 1);
    if (lane == 0) {
      const int kq = q % ksteps;
      const int ks = kq % ksteps1, rr = a.reduce_reps > 0 ? kq / ksteps1 : rep;
      const int packed = s_tile[q / ksteps];
      const int rb = packed >> 16, cb = packed & 0xffff;
      float* const sa = ring + st * C::kStageFloats;
      mbar_arrive_expect_tx(full + st, C::kStageBytes);
      tma_load_3d(sa, &tm_lhs, ks * BK, rb * BM, rr, full + st);
      tma_load_3d(sa + BM * BK, &tm_rhs, ks * BK, cb * BN, rr, full + st);
    }
    __syncwarp();
  };
  // Prefetch distance STAGES-2: the slot refilled at step q was consumed at step q-2, so the
  // producer warp does not have to wait for the slowest warp of the step it just left.
  constexpr int AHEAD = STAGES - 2;
  if (warp == 0)
    for (int q = 0; q < min(Q, AHEAD); ++q) issue(q);

  // Per-thread byte offsets inside a stage.  Rows i*8+ty have (row & 7) == ty; columns j*4+tx have
  // (col & 7) == (j & 1)*4 + tx: the swizzle XOR terms are per-thread constants.
  const int a_off = (wr * 64 + ty) * 128;
  const int b_off = BM * 128 + (wc * 32 + tx) * 128;
  const int xa = ty, xb0 = tx, xb1 = 4 + tx;

  int q = 0;
  for (int ti = 0; ti < ntiles; ++ti) {
    const int rb = s_tile[ti] >> 16, cb = s_tile[ti] & 0xffff;
    unsigned long long my_mask = 0ull;
    int my_base = 0;
    if (rb * BM + tid < a.m) {   // THREADS == BM
      const size_t e = (size_t)(rb * BM + tid) * col_blocks + cb;
      my_mask = __ldg(masks + e);
      my_base = __ldg(bases + e);
    }

    unsigned long long acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[i][j] = 0ull;

    for (int ks = 0; ks < ksteps; ++ks, ++q) {
      // The producer role rotates over the warps: with one fixed producer warp its extra work (the wait on
      // `empty`, the two TMA issues) made it the slowest, the other three ran ahead until they were blocked on
      // the slab it had not requested yet, and from then on took the full TMA latency at every slab (measured:
      // 63 % of the waits on `full` found the phase incomplete and lasted 715 cycles; rotating: 0.1 %).
      if (warp == (q & 3) && q + AHEAD < Q) issue(q + AHEAD);
      const int st = q % STAGES;
      mbar_wait(full + st, (q / STAGES) & 1);
      const char* const sbase = reinterpret_cast<const char*>(ring + st * C::kStageFloats);
      const char* const sa = sbase + a_off;
      const char* const sb = sbase + b_off;
      // One k-pair (even k, odd k) per step: 8 + 8 LDS.64 feed 64 FFMA2.  Narrow loads keep the
      // operand registers at 16 (+16 for the next step's lhs, requested one step ahead) next to
      // the 128 accumulator registers.
      auto a_addr = [&](int i, int kp) { return sa + i * 8 * 128 + (((kp >> 1) ^ xa) << 4) + (kp & 1) * 8; };
      auto step = [&](const unsigned long long (&av)[8], int kp) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const unsigned long long bq = *reinterpret_cast<const unsigned long long*>(
              sb + j * 4 * 128 + (((kp >> 1) ^ ((j & 1) ? xb1 : xb0)) << 4) + (kp & 1) * 8);
#pragma unroll
          for (int i = 0; i < 8; ++i) fma2(acc[i][j], av[i], bq);
        }
      };
      unsigned long long a0[8], a1[8];   // ping-pong lhs registers, each requested one step ahead
#pragma unroll
      for (int i = 0; i < 8; ++i) a0[i] = *reinterpret_cast<const unsigned long long*>(a_addr(i, 0));
#pragma unroll
      for (int kp = 0; kp < BK / 2; kp += 2) {
#pragma unroll
        for (int i = 0; i < 8; ++i) a1[i] = *reinterpret_cast<const unsigned long long*>(a_addr(i, kp + 1));
        step(a0, kp);
        if (kp + 2 < BK / 2) {
#pragma unroll
          for (int i = 0; i < 8; ++i) a0[i] = *reinterpret_cast<const unsigned long long*>(a_addr(i, kp + 2));
        }
        step(a1, kp + 1);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(empty + st);
    }

    rowmask[tid] = my_mask;
    rowbase[tid] = my_base;
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int rl = wr * 64 + i * 8 + ty;
      const unsigned long long mk = rowmask[rl];
      const int base = rowbase[rl];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int cl = wc * 32 + j * 4 + tx;
        if ((mk >> cl) & 1ull) {
          float lo, hi;
          unpack2(acc[i][j], lo, hi);
          out[base + __popcll(mk & ((1ull << cl) - 1ull))] = lo + hi;
        }
      }
    }
    __syncthreads();  // rowmask is rewritten by the next tile
  }
}


// cuTensorMapEncodeTiled through the runtime's driver-entry-point lookup (no libcuda link).
typedef CUresult (*TensorMapEncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                      const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                      const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                      CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
inline TensorMapEncodeFn tensor_map_encoder() {
  static TensorMapEncodeFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<TensorMapEncodeFn>(p);
  }
  return fn;
}

// [replicas][rows][k] fp32, K-contiguous; box = [1][box_rows][32 floats], 128-byte swizzle,
// out-of-range elements read as zero.
inline bool encode_rowmajor_map(CUtensorMap* tm, const float* base, int replicas, int rows, int k,
                                int box_rows) {
  TensorMapEncodeFn enc = tensor_map_encoder();
  if (!enc) return false;
  const cuuint64_t dims[3] = {(cuuint64_t)k, (cuuint64_t)rows, (cuuint64_t)replicas};
  const cuuint64_t strides[2] = {(cuuint64_t)k * 4, (cuuint64_t)rows * k * 4};
  const cuuint32_t box[3] = {32, (cuuint32_t)box_rows, 1};
  const cuuint32_t estr[3] = {1, 1, 1};
  return enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(base), dims, strides, box, estr,
             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

inline bool sddmm_tiled_applicable(const SddmmArgs& a, int replicas) {
  if (a.k % 4 != 0 || a.k < 128) return false;
  if (!aligned16(a.lhs) || !aligned16(a.rhs)) return false;
  if (a.m < 64 || a.n < 64) return false;
  return 2.0 * a.nnz * a.k * replicas >= 5.0e7;
}

// Bytes of the bitmap view of an m-row mask with `col_blocks` 64-column blocks: a 64-bit occupancy
// word + the flat index of the first nonzero per (row, block).
inline size_t sddmm_tiled_view_bytes(int m, int col_blocks) { return (size_t)m * col_blocks * 12 + 256; }

// Builds the bitmap view into `ws` (sddmm_tiled_view_bytes); flag_or_null != null also runs the
// sortedness check in the same pass.
inline int build_sddmm_tiled_view(const int* row_offsets, const int* column_indices, int m, int col_blocks,
                                  void* ws, unsigned* flag_or_null, cudaStream_t stream) {
  const size_t entries = (size_t)m * col_blocks;
  unsigned long long* masks = static_cast<unsigned long long*>(ws);
  int* bases = reinterpret_cast<int*>(masks + entries);
  cudaMemsetAsync(masks, 0, entries * 8, stream);
  csr_block_masks_kernel<<<ceil_div((long long)m * 32, 256), 256, 0, stream>>>(
      row_offsets, column_indices, m, col_blocks, masks, bases, flag_or_null);
  TSP_CHECK_LAUNCH("tsp_sddmm_f32 (block masks)");
  return 0;
}

// Returns 0 / an error code.  `view` is the bitmap view of the topology (from its plan).
inline int launch_sddmm_tiled(const SddmmArgs& a, int replicas, cudaStream_t stream, const StagedGuard& guard,
                              const void* view, bool v1 = false) {
  using C = SddmmTiled;
  static bool configured[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 0 && dev < 64 && !configured[dev]) {
    cudaError_t e = cudaFuncSetAttribute(sddmm_tiled_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)C::kSmemBytes);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(sddmm_tiled2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               (int)SddmmTiled2::kSmemBytes);
    if (e != cudaSuccess) return cuda_fail(e, "tsp_sddmm_f32 (tiled smem attribute)");
    configured[dev] = true;
  }
  const int row_blocks = ceil_div(a.m, C::BM), col_blocks = ceil_div(a.n, C::BN);
  const long long total = (long long)row_blocks * col_blocks;
  // One wave of 2 CTAs per SM when the replicas do not already fill the machine.
  const int slots = 2 * sm_count();
  const int grid_reps = a.reduce_reps > 0 ? 1 : replicas;
  int per_cta = (int)((total * grid_reps + slots - 1) / slots);
  per_cta = per_cta < 1 ? 1 : (per_cta > col_blocks ? col_blocks : per_cta);
  if (per_cta > 64) per_cta = 64;   // sddmm_tiled2_kernel keeps its tile list in a 64-entry shared array
  const int ctas = (int)((total + per_cta - 1) / per_cta);
  const size_t entries = (size_t)a.m * col_blocks;
  const unsigned long long* masks = static_cast<const unsigned long long*>(view);
  const int* bases = reinterpret_cast<const int*>(masks + entries);
  CUtensorMap tm_lhs, tm_rhs;
  if (!encode_rowmajor_map(&tm_lhs, a.lhs, replicas, a.m, a.k, C::BM) ||
      !encode_rowmajor_map(&tm_rhs, a.rhs, replicas, a.n, a.k, C::BN)) {
    set_error("tsp_sddmm_f32: cuTensorMapEncodeTiled failed");
    return 2;
  }
  dim3 grid(ctas, grid_reps);
  if (v1)
    sddmm_tiled_kernel<<<grid, C::THREADS, C::kSmemBytes, stream>>>(a, guard, per_cta, col_blocks,
                                                                    (int)total, masks, bases, tm_lhs, tm_rhs);
  else
    sddmm_tiled2_kernel<<<grid, SddmmTiled2::THREADS, SddmmTiled2::kSmemBytes, stream>>>(
        a, guard, per_cta, col_blocks, (int)total, masks, bases, tm_lhs, tm_rhs);
  TSP_CHECK_LAUNCH("tsp_sddmm_f32 (tiled)");
  return 0;
}

}  // namespace tsp
