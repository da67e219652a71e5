// sddmm_sampled.cuh -- SDDMM, variant 3 ("sampled"): operands staged in shared memory by TMA, only
// the sampled products computed.  The path for 2 % .. 20 % dense masks with a long inner dimension.
//
// out[j] = <lhs[row_of(j), :], rhs[col(j), :]>.  The gather SDDMM (variant 1, and the reference,
// sddmm/cuda_sddmm.cu.cc:34-165) pulls a K-long rhs row through L1/L2 for every nonzero; the tiled
// variant computes all BM x BN products and throws (1 - d) of them away.  Between the two:
//
//  * a CTA owns a [128 rows x BN cols] tile of the mask (BN sized from the density so the tile holds
//    ~1400 nonzeros) for the whole K loop; per 32-deep k-slab two TMA tensor-map loads bring the
//    128 lhs rows and BN rhs rows (128 B each) into a shared-memory ring (mbarrier complete_tx),
//    so every operand byte crosses L2 -> SM once per tile instead of once per nonzero;
//  * the tile's nonzeros are flattened (row-major) and dealt evenly to the 32 eight-lane groups
//    of the CTA, P slots each, as packed (rhs offset | lhs offset | new-row) words held in
//    REGISTERS for the whole K loop, next to P fp32 accumulators per lane;
//  * per slab and nonzero a group issues ONE LDS.128 (each lane its 16-byte k-slice of the rhs
//    row; the 8 lanes cover the 128-byte row, so a quarter-warp reads one full row: 4 wavefronts
//    per warp instruction, conflict-free, tools/ubench/ubench_lds.cu pattern 7) and 4 FFMA;
//    the lhs slice is reloaded only when the row changes (predicated LDS.128);
//  * after the K loop a transposing butterfly over the 8 lanes (7 shuffles per 8 nonzeros) leaves
//    one finished dot product per lane, written to out[] through the slot's flat nonzero index.
//
// Bound: 4 bytes of shared-memory bandwidth per FMA -> 32 FMA/clk/SM = 25 % of the fp32 pipe
// (18.6 TFLOP/s); there is no reuse to exploit below ~20 % density.
// Tiles holding more than 32*P nonzeros (skewed patterns) take extra rounds over the K loop.
//
// Precondition: ascending columns per row (device-checked; gather fallback otherwise), k % 4 == 0,
// 16-byte aligned operands.
#pragma once
#include <algorithm>
#include <type_traits>
#include "sddmm_tiled.cuh"

namespace tsp {

template <int P_, int STAGES_>
struct SddmmSampledCfg {
  static constexpr int BM = 128, BK = 32, THREADS = 256, GROUPS = 32;
  static constexpr int P = P_, STAGES = STAGES_;
  static constexpr int CAP = GROUPS * P_;
  static size_t smem_bytes(int bn) {
    return (size_t)STAGES_ * (BM + bn) * BK * 4 + (size_t)CAP * 8 + BM * 4 + (BM + 1) * 4 + BM * 4 + BM * 4 + 64 * 4 + 2 * GROUPS * 4 + 2 * STAGES_ * 8 + 256;
  }
};

// starts[row * (nb + 1) + b] = first flat index in row `row` whose column is >= b * bn
// (b == nb gives the row's end).  Needs ascending columns.
// One launch for both pre-passes of the sampled kernel: blocks [0, check_blocks) are
// csr_check_sorted_kernel, the rest csr_block_starts_kernel (small masks are bound by the number
// of launches per call, not by the work in them).
__global__ void __launch_bounds__(256)
csr_prepare_sampled_kernel(const int* __restrict__ row_offsets, const int* __restrict__ cols, int m, int nb, int bn,
                           int* __restrict__ starts, int check_blocks, unsigned* __restrict__ flag) {
  if ((int)blockIdx.x < check_blocks) {   // (check_blocks == 0: the plan already ran the check)
    const int warp = (blockIdx.x * 256 + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    const int nwarps = (check_blocks * 256) >> 5;
    bool bad = false;
    for (int row = warp; row < m; row += nwarps) {
      const int beg = __ldg(row_offsets + row), end = __ldg(row_offsets + row + 1);
      for (int j = beg + lane; j + 1 < end; j += 32) bad |= __ldg(cols + j) >= __ldg(cols + j + 1);
    }
    if (__any_sync(kFullMask, bad) && lane == 0) *flag = 1u;
    return;
  }
  const long long t = (long long)(blockIdx.x - check_blocks) * 256 + threadIdx.x;
  if (t >= (long long)m * (nb + 1)) return;
  const int row = (int)(t / (nb + 1)), b = (int)(t % (nb + 1));
  int lo = __ldg(row_offsets + row), hi = __ldg(row_offsets + row + 1);
  if (b == nb) {
    starts[t] = hi;
    return;
  }
  const int target = b * bn;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (__ldg(cols + mid) < target) lo = mid + 1; else hi = mid;
  }
  starts[t] = lo;
}

__device__ __forceinline__ float4 lds_f4(const unsigned char* p) {
  return *reinterpret_cast<const float4*>(p);
}
__device__ __forceinline__ float dot4(const float4& l, const float4& r, float acc) {
  acc = fmaf(l.x, r.x, acc);
  acc = fmaf(l.y, r.y, acc);
  acc = fmaf(l.z, r.z, acc);
  return fmaf(l.w, r.w, acc);
}

// NB_LO: batches (of 4 slots) executed unconditionally -- straight-line code the scheduler can
// software-pipeline across; the host derives it from the expected tile population minus 4 sigma.
// 16-bit storage: a row's 128-byte slab holds 64 elements and stays PACKED in registers; every pair of
// elements goes through the mixed-precision FMA (fma.rn.f32.f16 / .bf16, fp32 accumulator), so one LDS.128
// feeds 8 multiply-adds instead of 4 -- the shared-memory bound of this kernel doubles.
template <typename T>
__device__ __forceinline__ float dot_slab16(const uint4& l, const uint4& r, float acc) {
  acc = fma_pair_16<T>(l.x, r.x, acc);
  acc = fma_pair_16<T>(l.y, r.y, acc);
  acc = fma_pair_16<T>(l.z, r.z, acc);
  return fma_pair_16<T>(l.w, r.w, acc);
}
template <typename T>
__device__ __forceinline__ float dot_slab(const uint4& l, const uint4& r, float acc) {
  if constexpr (sizeof(T) == 4) {
    acc = fmaf(__uint_as_float(l.x), __uint_as_float(r.x), acc);
    acc = fmaf(__uint_as_float(l.y), __uint_as_float(r.y), acc);
    acc = fmaf(__uint_as_float(l.z), __uint_as_float(r.z), acc);
    return fmaf(__uint_as_float(l.w), __uint_as_float(r.w), acc);
  } else {
    return dot_slab16<T>(l, r, acc);
  }
}
__device__ __forceinline__ uint4 lds_u4(const unsigned char* p) { return *reinterpret_cast<const uint4*>(p); }

template <class Cfg, int NB_LO, typename T = float>
__global__ void __launch_bounds__(Cfg::THREADS, 2)
sddmm_sampled_kernel(const SddmmArgsT<T> a, const StagedGuard guard, int nb, int bn,
                     const int* __restrict__ starts, const __grid_constant__ CUtensorMap tm_lhs,
                     const __grid_constant__ CUtensorMap tm_rhs, const int ksplit) {
  constexpr int BM = Cfg::BM, P = Cfg::P, STAGES = Cfg::STAGES, CAP = Cfg::CAP;
  constexpr int BK = Cfg::BK * 4 / (int)sizeof(T);   // elements in a 128-byte slab row
  if (staged_declines(guard)) return;

  extern __shared__ unsigned char smem_dyn[];
  unsigned char* const smem_raw = smem_dyn + ((128u - (smem_u32(smem_dyn) & 127u)) & 127u);
  const int stage_bytes = (BM + bn) * 128;
  unsigned char* const ring = smem_raw;
  unsigned* const ent = reinterpret_cast<unsigned*>(ring + (size_t)STAGES * stage_bytes);
  int* const oidx = reinterpret_cast<int*>(ent + CAP);
  int* const row_start = oidx + CAP;       // [BM]   flat index of the row's first nonzero in the tile
  int* const row_off = row_start + BM;     // [BM+1] exclusive prefix of the per-row counts, each rounded up to even
  int* const row_cnt = row_off + BM + 1;   // [BM]   nonzeros of the row inside the tile
  int* const order = row_cnt + BM;         // [BM]   rows by descending count (aligned dealing)
  int* const quad_off = order + BM;        // [32]   first slot of a quad of rows inside its warp; [32..39] slots per warp
  unsigned* const fresh = reinterpret_cast<unsigned*>(quad_off + 64);   // [2*GROUPS] new-row bits per slot
  unsigned long long* const full = reinterpret_cast<unsigned long long*>(
      (reinterpret_cast<uintptr_t>(fresh + 2 * Cfg::GROUPS) + 7) & ~uintptr_t(7));
  unsigned long long* const empty = full + STAGES;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int grp = tid >> 3, sub = tid & 7;
  const int rb = blockIdx.x / nb, cb = blockIdx.x % nb;
  const int rep = blockIdx.y;
  const int ksteps1 = (a.k + BK - 1) / BK;                                   // k-slabs of one replica
  const int ksteps_all = a.reduce_reps > 0 ? ksteps1 * a.reduce_reps : ksteps1;   // slabs per tile round
  // split-K (small masks: fewer tiles than SMs): grid.z CTAs share a tile, each takes a contiguous
  // range of the slabs and writes its partial sums to slice (rep * ksplit + z) of `out`, which is
  // then scratch that reduce_replicas_kernel sums.
  const int kbeg = (int)((long long)ksteps_all * blockIdx.z / ksplit);
  const int ksteps = (int)((long long)ksteps_all * (blockIdx.z + 1) / ksplit) - kbeg;

  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full + s, 1);
      mbar_init(empty + s, Cfg::THREADS / 32);
    }
    mbar_fence_init();
  }

  // ---- per-row ranges of this tile and their exclusive prefix (one warp scan over 128 counts) ----
  if (tid < BM) {
    const int row = rb * BM + tid;
    int s = 0, e = 0;
    if (row < a.m) {
      const int* p = starts + (size_t)row * (nb + 1) + cb;
      s = __ldg(p);
      e = __ldg(p + 1);
    }
    row_start[tid] = s;
    row_cnt[tid] = e - s;
    // A row's slots start at an EVEN position (odd rows get one empty slot): a row can then change only at an
    // even slot, and the K loop carries one predicated lhs load per TWO slots.  Measured
    // (tools/ubench/ubench_lds_pred.cu): that load costs 4 cycles of the shared-memory pipe whenever any of the
    // warp's four groups takes it and 1 cycle when none does -- next to the 4 cycles of the slot's rhs load.
    row_off[tid + 1] = (e - s + 1) & ~1;
  }
  if (tid == 0) row_off[0] = 0;
  __syncthreads();
  if (warp == 0) {
    int c[4], sum = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      c[i] = row_off[1 + lane * 4 + i];
      sum += c[i];
    }
    int incl = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int v = __shfl_up_sync(kFullMask, incl, d);
      if (lane >= d) incl += v;
    }
    int run = incl - sum;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      run += c[i];
      row_off[1 + lane * 4 + i] = run;
    }
  }
  __syncthreads();
  const int total = row_off[BM];
  if (total == 0) return;

  // ---- aligned dealing (tiles that fit one round) ----
  // The four groups of a warp change rows at the SAME slots: the rows are ranked by their count, consecutive
  // quads of that ranking (counts within a quad differ by 0 - 1) go to one warp each, one row per group, padded
  // to the quad's longest row (rounded up to even); a warp takes four quads, dealt in snake order so the warps'
  // totals stay level.  The predicated lhs load of a slot pair then has all four quarter-warps active when
  // it fires (4 shared-memory cycles for four rows instead of 4 for one) and fires in 16 % of the pairs
  // instead of 50 %.  Tiles whose padded quads overflow a warp's P slots keep the flat dealing below.
  // (Worth its set-up -- the ranking -- from ~8 nonzeros per row: 4096^3 at 5 / 10 / 20 % 0.670 -> 0.641 /
  // 1.289 -> 1.250 / 2.53 -> 2.43 ms; at 2 %, 5 nonzeros per row, it costs 9 %.)
  bool aligned = false;
  if (total <= CAP && total >= 8 * BM) {
    if (tid < BM) {
      const int c = row_cnt[tid];
      int rank = 0;
      for (int j = 0; j < BM; ++j) {
        const int cj = row_cnt[j];
        rank += (cj > c || (cj == c && j < tid)) ? 1 : 0;
      }
      order[rank] = tid;
    }
    __syncthreads();
    if (tid < 8) {   // warp `tid`: its quads are tid, 15 - tid, 16 + tid, 31 - tid (snake over the ranking)
      const int qs[4] = {tid, 15 - tid, 16 + tid, 31 - tid};
      int run = 0;
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        quad_off[qs[r]] = run;
        run += (row_cnt[order[4 * qs[r]]] + 1) & ~1;   // the quad's longest row, even
      }
      quad_off[32 + tid] = run;
    }
    __syncthreads();
    int worst = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) worst = max(worst, quad_off[32 + w]);
    aligned = worst <= P;
  }

  const int* __restrict__ cols = a.column_indices;
  T* __restrict__ out = a.out + ((size_t)rep * ksplit + blockIdx.z) * a.nnz;
  const int c0 = cb * bn;
  int q = 0, q_base = 0;  // slabs consumed so far by this CTA (ring position carries across rounds)


  for (int lo = 0; lo < total; lo += CAP) {
    const int cnt_round = min(CAP, total - lo);
    const int per = ((cnt_round + Cfg::GROUPS - 1) / Cfg::GROUPS + 1) & ~1;  // slots in use per group (<= P), even
    const int nbatch = aligned ? (quad_off[32 + warp] + 3) >> 2 : (per + 3) >> 2;   // 4 slots per batch (per WARP when aligned)

    // ---- deal the round's nonzeros to (group, slot) ----
    // slot word: (rhs row offset / 16) << 16 | (lhs row offset / 16); both < 4096, so bits 12..15
    // are zero and `word >> 12` IS the rhs byte offset.  fresh[g] bit s: slot s starts a new row.
    for (int i = tid; i < CAP; i += Cfg::THREADS) {
      ent[i] = 0u;
      oidx[i] = -1;
    }
    if (tid < 2 * Cfg::GROUPS) fresh[tid] = 0u;
    __syncthreads();
    if (aligned) {
      for (int qd = warp; qd < 32; qd += Cfg::THREADS / 32) {
        const int w = qd < 8 ? qd : qd < 16 ? 15 - qd : qd < 24 ? qd - 16 : 31 - qd;
        const int base = quad_off[qd];
#pragma unroll
        for (int gq = 0; gq < 4; ++gq) {
          const int r = order[4 * qd + gq], cnt = row_cnt[r], s = row_start[r], g = 4 * w + gq;
          for (int i = lane; i < cnt; i += 32) {
            const unsigned col_local = (unsigned)(__ldg(cols + s + i) - c0);
            ent[g * P + base + i] = (col_local << 19) | ((unsigned)r << 3);
            oidx[g * P + base + i] = s + i;
          }
          if (lane == 0) {   // (rows without nonzeros start a segment too: the load then reads row r for nothing)
            if (cnt == 0) ent[g * P + base] = (unsigned)r << 3;
            atomicOr(fresh + 2 * g + (base >> 5), 1u << (base & 31));
          }
        }
      }
    } else
    for (int r = warp; r < BM; r += Cfg::THREADS / 32) {
      const int off = row_off[r], cnt = row_cnt[r], s = row_start[r];
      for (int i = lane; i < cnt; i += 32) {
        const int el = off + i - lo;
        if (el < 0 || el >= cnt_round) continue;
        const int g = el / per, slot = el - g * per;
        const unsigned col_local = (unsigned)(__ldg(cols + s + i) - c0);
        ent[g * P + slot] = (col_local << 19) | ((unsigned)r << 3);   // (col*8) << 16 | row*8
        oidx[g * P + slot] = s + i;
        if (i == 0 || slot == 0) atomicOr(fresh + 2 * g + (slot >> 5), 1u << (slot & 31));
      }
    }
    __syncthreads();

    const uint4* const my4 = reinterpret_cast<const uint4*>(ent + grp * P);
    const unsigned f_lo = fresh[2 * grp], f_hi = fresh[2 * grp + 1];
    float acc[P];
#pragma unroll
    for (int e = 0; e < P; ++e) acc[e] = 0.f;

    // ---- K loop: TMA ring, thread 0 is the producer ----
    auto issue = [&](int qq) {   // called by thread 0 only
      const int st = qq % STAGES, use = qq / STAGES;
      if (use > 0) mbar_wait(empty + st, (use - 1) & 1);
      unsigned char* const dst = ring + (size_t)st * stage_bytes;
      const int kq = kbeg + (qq - q_base);
      const int ks = kq % ksteps1, rr = a.reduce_reps > 0 ? kq / ksteps1 : rep;
      mbar_arrive_expect_tx(full + st, (unsigned)stage_bytes);
      tma_load_3d(dst, &tm_lhs, ks * BK, rb * BM, rr, full + st);
      tma_load_3d(dst + BM * 128, &tm_rhs, ks * BK, c0, rr, full + st);
    };
    const int q_end = q + ksteps;
    if (tid == 0)
      for (int qq = q; qq < min(q_end, q + STAGES - 1); ++qq) issue(qq);

    for (; q < q_end; ++q) {
      if (tid == 0 && q + STAGES - 1 < q_end) issue(q + STAGES - 1);
      const int st = q % STAGES;
      mbar_wait(full + st, (q / STAGES) & 1);
      // Plain shared loads (not asm): the mbarrier asm statements carry memory clobbers, so the
      // compiler keeps these between the wait and the arrive but is free to pipeline them.
      const unsigned char* const sl = ring + (size_t)st * stage_bytes + sub * 16;  // lhs rows
      const unsigned char* const sr = sl + BM * 128;                                // rhs rows
      uint4 l = make_uint4(0u, 0u, 0u, 0u);
      uint4 u = my4[0];
#pragma unroll
      for (int b = 0; b < P / 4; ++b) {
        if (b < NB_LO || b < nbatch) {
          const uint4 uc = u;
          if (b + 1 < P / 4) u = my4[(b + 1 < NB_LO || b + 1 < nbatch) ? b + 1 : 0];
          const uint4 r0 = lds_u4(sr + (uc.x >> 12)), r1 = lds_u4(sr + (uc.y >> 12));
          const uint4 r2 = lds_u4(sr + (uc.z >> 12)), r3 = lds_u4(sr + (uc.w >> 12));
          const unsigned fw = (4 * b < 32) ? f_lo : f_hi;
          if ((fw >> ((4 * b + 0) & 31)) & 1u) l = lds_u4(sl + ((uc.x & 0xfffu) << 4));
          acc[4 * b + 0] = dot_slab<T>(l, r0, acc[4 * b + 0]);
          acc[4 * b + 1] = dot_slab<T>(l, r1, acc[4 * b + 1]);   // (odd slot: same row as the slot before)
          if ((fw >> ((4 * b + 2) & 31)) & 1u) l = lds_u4(sl + ((uc.z & 0xfffu) << 4));
          acc[4 * b + 2] = dot_slab<T>(l, r2, acc[4 * b + 2]);
          acc[4 * b + 3] = dot_slab<T>(l, r3, acc[4 * b + 3]);
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(empty + st);
    }
    q_base = q;

    // ---- reduce the 8 k-slices of every slot, one finished value per lane ----
#pragma unroll
    for (int b = 0; b < P / 8; ++b) {
      if (2 * b < nbatch) {
        float v[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = acc[b * 8 + i];
        transpose_reduce<8>(v, sub);
        const int j = oidx[grp * P + b * 8 + sub];
        if (j >= 0) out[j] = Elem<T>::from_f(v[0]);
      }
    }
    __syncthreads();  // ent/oidx are rewritten by the next round
  }
}

// [replicas][rows][k] fp32, K-contiguous; box = [1][box_rows][32 floats], no swizzle.
inline bool encode_rowmajor_map_plain(CUtensorMap* tm, const float* base, int replicas, int rows, int k,
                                      int box_rows) {
  TensorMapEncodeFn enc = tensor_map_encoder();
  if (!enc) return false;
  const cuuint64_t dims[3] = {(cuuint64_t)k, (cuuint64_t)rows, (cuuint64_t)replicas};
  const cuuint64_t strides[2] = {(cuuint64_t)k * 4, (cuuint64_t)rows * k * 4};
  const cuuint32_t box[3] = {32, (cuuint32_t)box_rows, 1};
  const cuuint32_t estr[3] = {1, 1, 1};
  return enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(base), dims, strides, box, estr,
             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// the same for 16-bit elements: box = [1][box_rows][64 elements] (128 bytes per row again)
inline bool encode_rowmajor_map_plain_b16(CUtensorMap* tm, const void* base, bool bf16, int replicas, int rows, int k,
                                          int box_rows) {
  TensorMapEncodeFn enc = tensor_map_encoder();
  if (!enc) return false;
  const cuuint64_t dims[3] = {(cuuint64_t)k, (cuuint64_t)rows, (cuuint64_t)replicas};
  const cuuint64_t strides[2] = {(cuuint64_t)k * 2, (cuuint64_t)rows * k * 2};
  const cuuint32_t box[3] = {64, (cuuint32_t)box_rows, 1};
  const cuuint32_t estr[3] = {1, 1, 1};
  return enc(tm, bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, const_cast<void*>(base),
             dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
             CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

template <typename T>
inline bool encode_sampled_map(CUtensorMap* tm, const T* base, int replicas, int rows, int k, int box_rows) {
  if constexpr (sizeof(T) == 4) return encode_rowmajor_map_plain(tm, base, replicas, rows, k, box_rows);
  else return encode_rowmajor_map_plain_b16(tm, base, std::is_same<T, __nv_bfloat16>::value, replicas, rows, k, box_rows);
}

inline bool sddmm_sampled_applicable(const SddmmArgs& a, int replicas) {
  if (a.k % 4 != 0 || a.k < 256) return false;
  if (!aligned16(a.lhs) || !aligned16(a.rhs)) return false;
  if (a.m < 128 || a.n < 64) return false;
  return 2.0 * a.nnz * a.k * replicas >= 5.0e7;
}

// Split-K factor: when the mask has fewer tiles than the GPU has CTA slots (sparse-linear weight
// gradients of small layers: a 512 x 512 mask is 16 tiles), grid.z CTAs share a tile; every CTA
// keeps >= 8 slabs.
inline int sddmm_sampled_ksplit(long long tiles, int grid_reps, int ksteps_all) {
  if (grid_reps != 1 || tiles >= sm_count()) return 1;
  long long ks = std::min<long long>(2LL * sm_count() / tiles, ksteps_all / 8);
  return (int)std::max<long long>(1, std::min<long long>(ks, 64));
}

// Block-start view of an m-row mask for tile width bn: (ceil(n / bn) + 1) ints per row.
inline size_t sddmm_sampled_view_bytes(int m, int n, int bn) { return (size_t)m * (ceil_div(n, bn) + 1) * 4 + 256; }

// Builds the view into `ws`; flag_or_null != null also runs the sortedness check in the same launch.
inline int build_sddmm_sampled_view(const int* row_offsets, const int* column_indices, int m, int n, int bn,
                                    void* ws, unsigned* flag_or_null, cudaStream_t stream) {
  const int nb = ceil_div(n, bn);
  const int check_blocks = flag_or_null ? min(ceil_div(m, 8), 4 * sm_count()) : 0;
  csr_prepare_sampled_kernel<<<check_blocks + (int)ceil_div((long long)m * (nb + 1), 256), 256, 0, stream>>>(
      row_offsets, column_indices, m, nb, bn, static_cast<int*>(ws), check_blocks, flag_or_null);
  TSP_CHECK_LAUNCH("tsp_sddmm_f32 (block starts)");
  return 0;
}

template <class Cfg, int NB_LO, typename T>
inline int launch_sddmm_sampled_nb(const SddmmArgsT<T>& a, int replicas, cudaStream_t stream,
                                   const StagedGuard& guard, int bn, const int* starts) {
  constexpr int BKE = Cfg::BK * 4 / (int)sizeof(T);   // elements per slab row
  const int row_blocks = ceil_div(a.m, Cfg::BM), nb = ceil_div(a.n, bn);
  const size_t smem = Cfg::smem_bytes(bn);
  static bool configured[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 0 && dev < 64 && !configured[dev]) {
    cudaError_t e = cudaFuncSetAttribute(sddmm_sampled_kernel<Cfg, NB_LO, T>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Cfg::smem_bytes(256));
    if (e != cudaSuccess) return cuda_fail(e, "tsp_sddmm_f32 (sampled smem attribute)");
    configured[dev] = true;
  }
  CUtensorMap tm_lhs, tm_rhs;
  if (!encode_sampled_map<T>(&tm_lhs, a.lhs, replicas, a.m, a.k, Cfg::BM) ||
      !encode_sampled_map<T>(&tm_rhs, a.rhs, replicas, a.n, a.k, bn)) {
    set_error("tsp_sddmm_f32: cuTensorMapEncodeTiled failed");
    return 2;
  }
  const int grid_reps = a.reduce_reps > 0 ? 1 : replicas;
  // Split-K when the mask has fewer tiles than the GPU has CTA slots (sparse-linear weight
  // gradients of small layers: a 512 x 512 mask is 16 tiles): every CTA keeps >= 8 slabs.
  const int ksteps_all = ceil_div(a.k, BKE) * (a.reduce_reps > 0 ? a.reduce_reps : 1);
  const long long tiles = (long long)row_blocks * nb * grid_reps;
  // (16-bit outputs: no split-K -- the partial sums would have to stay fp32; the caller sends such shapes
  // through the widened fp32 path instead)
  const int ksplit = sizeof(T) == 4 ? sddmm_sampled_ksplit(tiles, grid_reps, ksteps_all) : 1;
  SddmmArgsT<T> b = a;
  T* partial = nullptr;
  if (ksplit > 1) {
    cudaError_t e = cudaMallocAsync(reinterpret_cast<void**>(&partial), sizeof(T) * (size_t)ksplit * a.nnz, stream);
    if (e != cudaSuccess) return cuda_fail(e, "tsp_sddmm_f32 (split-K scratch)");
    b.out = partial;
  }
  dim3 grid(row_blocks * nb, grid_reps, ksplit);
  sddmm_sampled_kernel<Cfg, NB_LO, T><<<grid, Cfg::THREADS, smem, stream>>>(b, guard, nb, bn, starts, tm_lhs, tm_rhs, ksplit);
  TSP_CHECK_LAUNCH("tsp_sddmm_f32 (sampled)");
  if constexpr (sizeof(T) == 4) if (ksplit > 1) {
    // (with unsorted columns the sampled kernel exits and this sums garbage; the guarded gather
    // fallback queued behind by the caller then rewrites every element of out)
    reduce_replicas_kernel<<<ceil_div(a.nnz, 256), 256, 0, stream>>>(partial, a.nnz, ksplit, a.out);
    TSP_CHECK_LAUNCH("tsp_sddmm_f32 (split-K sum)");
    cudaFreeAsync(partial, stream);
  }
  return 0;
}

template <class Cfg, typename T>
inline int launch_sddmm_sampled_cfg(const SddmmArgsT<T>& a, int replicas, cudaStream_t stream,
                                    const StagedGuard& guard, int bn, const int* starts) {
  // Batches every tile is expected to fill: (mu - 4 sigma) / 32 slots per group, 4 slots per batch.
  const double density = (double)a.nnz / ((double)a.m * (double)a.n);
  const double mu = (double)Cfg::BM * bn * density;
  const double lo = mu - 4.0 * sqrt(mu);
  const int nb_lo = lo > 0 ? (int)(lo / Cfg::GROUPS) / 4 : 0;
  if (nb_lo >= 12 && Cfg::P >= 48) return launch_sddmm_sampled_nb<Cfg, 12, T>(a, replicas, stream, guard, bn, starts);
  if (nb_lo >= 8) return launch_sddmm_sampled_nb<Cfg, 8, T>(a, replicas, stream, guard, bn, starts);
  if (nb_lo >= 4) return launch_sddmm_sampled_nb<Cfg, 4, T>(a, replicas, stream, guard, bn, starts);
  return launch_sddmm_sampled_nb<Cfg, 0, T>(a, replicas, stream, guard, bn, starts);
}

// Tile width: among the multiples of 8 columns (<= 256, <= n) whose expected nonzero count plus
// three standard deviations fits the 32*P register slots of a CTA, the one with the lowest
// modelled time = waves of 2 CTAs per SM x (slabs per CTA x per-slab cost + the per-tile set-up),
// where the per-slab cost is the larger of the shared-memory wavefronts (4 clk per slot per warp)
// and the operand staging traffic, and the set-up (row ranges, dealing the nonzeros to slots,
// filling the ring, the final butterfly and scatter) is ~6 us + 64 clk per slot -- it decides the
// width for small masks, where a CTA sees only a few slabs.
template <class Cfg>
inline int sddmm_sampled_tile_width(const SddmmArgs& a, int replicas) {
  const double density = (double)a.nnz / ((double)a.m * (double)a.n);
  const int row_blocks = ceil_div(a.m, Cfg::BM);
  const int slots = 2 * sm_count();
  const int ksteps_all = ceil_div(a.k, Cfg::BK) * (a.reduce_reps > 0 ? a.reduce_reps : 1);
  int best = 16;
  double best_cost = 1e300;
  const int bn_max = a.n / 8 * 8 < 256 ? a.n / 8 * 8 : 256;
  for (int bn = 16; bn <= bn_max; bn += 8) {
    const double mu = (double)Cfg::BM * bn * density;
    const double need = mu + 3.0 * sqrt(mu) + 1.0 + 0.5 * Cfg::BM;   // (+ half a slot per row: even-aligned rows)
    if (need > (double)Cfg::CAP && bn > 16) break;
    const int per = (int)ceil(need / Cfg::GROUPS);
    const int batches = (per + 3) / 4;
    const long long tiles = (long long)row_blocks * ceil_div(a.n, bn) * replicas;
    const int ksplit = sddmm_sampled_ksplit(tiles, replicas, ksteps_all);
    const double waves = (double)((tiles * ksplit + slots - 1) / slots);
    const double lsu = 64.0 * (4.0 * batches + 2.0), mem = 4.0 * (Cfg::BM + bn);
    const double slab = (lsu > mem ? lsu : mem) + 200.0;
    const double setup = 12000.0 + 64.0 * per;
    const double cost = waves * (slab * ceil_div(ksteps_all, ksplit) + setup);
    if (cost < best_cost) {
      best_cost = cost;
      best = bn;
    }
  }
  return best;
}

// Tile width the launcher below wants for this call (the key of the block-start view in the plan).
inline int sddmm_sampled_pick_bn(const SddmmArgs& a, int replicas, int bn_override = 0, int p_override = 0) {
  if (bn_override > 0) return bn_override;
  const int grid_reps = a.reduce_reps > 0 ? 1 : replicas;   // tiles are spread over this many grid.y slices
  if (p_override == 48) return sddmm_sampled_tile_width<SddmmSampledCfg<48, 2>>(a, grid_reps);
  return sddmm_sampled_tile_width<SddmmSampledCfg<56, 2>>(a, grid_reps);
}

// Returns 0 / an error code.  `starts` is the block-start view for width bn (from the plan).
template <typename T>
inline int launch_sddmm_sampled(const SddmmArgsT<T>& a, int replicas, cudaStream_t stream, const StagedGuard& guard,
                                int bn, const int* starts, int p_override = 0) {
  if constexpr (sizeof(T) == 4) {
    if (p_override == 48)
      return launch_sddmm_sampled_cfg<SddmmSampledCfg<48, 2>, T>(a, replicas, stream, guard, bn, starts);
  }
  return launch_sddmm_sampled_cfg<SddmmSampledCfg<56, 2>, T>(a, replicas, stream, guard, bn, starts);
}

// 16-bit storage: the shape the native sampled kernel takes (no split-K, no replica sum) and the
// "shape-only" fp32 arguments its tile-width model is asked with: a slab holds 64 elements, so the kernel
// sees half as many slabs as the fp32 one does for the same k.
template <typename T>
inline bool sddmm_sampled_b16_applicable(const SddmmArgsT<T>& a, int replicas, SddmmArgs* shape) {
  if (a.reduce_reps > 0 || a.k % 8 != 0 || a.k < 512) return false;
  if (!aligned16(a.lhs) || !aligned16(a.rhs)) return false;
  if (a.m < 128 || a.n < 64) return false;
  if (2.0 * a.nnz * a.k * replicas < 5.0e7) return false;
  shape->m = a.m; shape->n = a.n; shape->nnz = a.nnz; shape->k = a.k / 2;
  shape->lhs = shape->rhs = nullptr; shape->out = nullptr;
  shape->row_offsets = a.row_offsets; shape->column_indices = a.column_indices;
  const int bn = sddmm_sampled_pick_bn(*shape, replicas);
  const long long tiles = (long long)ceil_div(a.m, 128) * ceil_div(a.n, bn) * replicas;
  return sddmm_sampled_ksplit(tiles, replicas, ceil_div(a.k, 64)) == 1;
}

}  // namespace tsp
