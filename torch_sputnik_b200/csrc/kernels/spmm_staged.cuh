// spmm_staged.cuh -- CSR SpMM, variant 2 ("staged"): the fast path for sorted-column CSR.
//
// Why a second kernel.  A pure gather SpMM (the reference's design, cuda_spmm.cu.cc:224-238, and
// our variant 1) needs 4 bytes of L1/shared bandwidth per FMA: one LDG/LDS.128 of the dense
// operand feeds 4 FFMA.  An SM delivers 128 B/clk of L1/shared and issues 128 FMA/clk, so that
// design tops out at 25 % of the fp32 peak -- and lower still when the gathers miss L1 and go to
// L2 (about 42 B/clk/SM).  This kernel attacks both limits:
//
//  * the dense operand is STAGED: a dedicated producer warp streams [32 x NT] slabs of B into a
//    multi-stage shared-memory ring with bulk async copies (cp.async.bulk -> mbarrier
//    complete_tx, the TMA engine without a descriptor), so every B byte crosses L2->SM once per
//    row block, coalesced, and all gathers hit shared memory;
//  * rows are MERGED four at a time: the four rows' nonzeros that fall in the slab are scattered
//    into a dense [32 x 4] value tile (zeros elsewhere) plus a 32-bit occupancy mask; the warp
//    then walks the set bits, and for every occupied column loads the B row segment ONCE
//    (LDS.128 per lane) and one broadcast LDS.128 of the four values, and issues 4 x NT/32 FMAs.
//    That raises FMA per shared-memory wavefront from 1 : 1 (gather) to about R*d/(1-(1-d)^R)
//    useful FMAs per B load, R = 4: x1.16 at d = 0.1, x2.1 at d = 0.5 (DESIGN.md has the table);
//  * the FMA stream is packed fma.rn.f32x2 (SASS FFMA2): half the issue slots for the same
//    arithmetic and rounding, leaving issue room for the LDS and the bit-walk.
//
// Arithmetic: per output element, fp32 FMAs in ascending column order, exactly the reference's
// order; entries a row does not have contribute fma(0, b, acc) = acc for finite b.  A non-finite b
// would leak into the rows merged with the one that owns the column (0 * Inf = NaN), which the
// reference never does (cuda_spmm.cu.cc:224-238 reads only the dense rows a row names): the epilogue
// therefore tests every row's accumulators and recomputes a row that holds a non-finite value
// the way the gather kernel does, from its own nonzeros only (spmm_staged_redo_row).
// Precondition: column indices strictly ascending inside each row (what to_sparse_csr and the
// reference's generators produce).  tsp_spmm_f32 verifies it on the device in the same stream
// (csr_check_sorted_kernel) and falls back to the gather kernel if it does not hold.
#pragma once
#include "spmm_gather.cuh"

namespace tsp {

// *flag = 1 if some row has a column index not greater than its predecessor (*flag is 0 on entry).
__global__ void __launch_bounds__(256)
csr_check_sorted_kernel(const int* __restrict__ row_offsets, const int* __restrict__ cols, int m,
                        unsigned* __restrict__ flag) {
  const int warp = (blockIdx.x * 256 + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  const int nwarps = (gridDim.x * 256) >> 5;
  bool bad = false;
  for (int row = warp; row < m; row += nwarps) {
    const int beg = __ldg(row_offsets + row), end = __ldg(row_offsets + row + 1);
    for (int j = beg + lane; j + 1 < end; j += 32) bad |= __ldg(cols + j) >= __ldg(cols + j + 1);
  }
  if (__any_sync(kFullMask, bad) && lane == 0) *flag = 1u;
}

__device__ __forceinline__ unsigned long long pack2(float lo, float hi) {
  unsigned long long r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void unpack2(unsigned long long v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ void fma2(unsigned long long& d, unsigned long long a,
                                     unsigned long long b) {
  asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(d) : "l"(a), "l"(b));
}

template <int NV4_, int R_, int G_, int STAGES_, int WARPS_, int U_>
struct StagedCfg {
  static constexpr int R = R_;            // rows merged per group (4 or 8)
  static constexpr int LPR = 32 / R_;     // lanes (= candidate slots per batch) per row
  static constexpr int NV4 = NV4_;        // float4 chunks of B per lane
  static constexpr int NT = 128 * NV4_;   // output columns per CTA
  static constexpr int G = G_;            // R-row groups per warp
  static constexpr int KS = 32;           // slab depth == width of the occupancy mask
  static constexpr int STAGES = STAGES_;
  static constexpr int WARPS = WARPS_;    // compute warps (+1 producer warp)
  static constexpr int U = U_;            // prefetched candidates per row per slab = LPR*U
  static constexpr int RB = WARPS_ * G_ * R_;  // rows per CTA
  static constexpr int THREADS = (WARPS_ + 1) * 32;
  static constexpr size_t kSlabFloats = (size_t)KS * NT;
  // R values per occupied column + the column's byte offset
  static constexpr int kTileFloats = KS * R_ + KS;
  static constexpr size_t kSmemBytes =
      sizeof(float) * (STAGES_ * kSlabFloats + (size_t)WARPS_ * G_ * kTileFloats) + 2 * STAGES_ * 8 + 128;
};

// Rare path of the staged epilogue: one output row x this lane's NV4 float4 columns, recomputed in
// CSR order from the row's own nonzeros (global-memory gathers), bias / ReLU applied, stored.
template <int NV4>
__device__ __noinline__ void spmm_staged_redo_row(const float* __restrict__ dense, const float* __restrict__ vals,
                                                  const int* __restrict__ cols, int beg, int end, int n, int n0,
                                                  int lane, float bias, int relu, float* __restrict__ out) {
  float4 acc[NV4];
#pragma unroll
  for (int j = 0; j < NV4; ++j) acc[j] = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int p = beg; p < end; ++p) {
    const float v = __ldg(vals + p);
    const float* brow = dense + (size_t)__ldg(cols + p) * n + n0 + lane * 4;
#pragma unroll
    for (int j = 0; j < NV4; ++j) {
      if (n0 + lane * 4 + j * 128 >= n) continue;
      const float4 b = __ldg(reinterpret_cast<const float4*>(brow + j * 128));
      acc[j].x = fmaf(v, b.x, acc[j].x);
      acc[j].y = fmaf(v, b.y, acc[j].y);
      acc[j].z = fmaf(v, b.z, acc[j].z);
      acc[j].w = fmaf(v, b.w, acc[j].w);
    }
  }
#pragma unroll
  for (int j = 0; j < NV4; ++j) {
    if (n0 + lane * 4 + j * 128 >= n) continue;
    float4 o = acc[j];
    o.x += bias; o.y += bias; o.z += bias; o.w += bias;
    if (relu) {
      o.x = fmaxf(o.x, 0.f); o.y = fmaxf(o.y, 0.f); o.z = fmaxf(o.z, 0.f); o.w = fmaxf(o.w, 0.f);
    }
    *reinterpret_cast<float4*>(out + j * 128) = o;
  }
}

template <class Cfg>
__global__ void __launch_bounds__(Cfg::THREADS, (Cfg::kSmemBytes <= 110 * 1024) ? 2 : 1)
spmm_staged_kernel(const SpmmArgs a, const StagedGuard guard, const int row_blocks, const int col_tiles, const int srb,
                   const int scb) {
  constexpr int NV4 = Cfg::NV4, NT = Cfg::NT, G = Cfg::G, KS = Cfg::KS, STAGES = Cfg::STAGES,
                WARPS = Cfg::WARPS, U = Cfg::U, R = Cfg::R, LPR = Cfg::LPR;
  if (staged_declines(guard)) return;  // gather kernel takes over

  extern __shared__ __align__(128) unsigned char smem_raw[];
  float* const b_ring = reinterpret_cast<float*>(smem_raw);
  float* const a_tiles = b_ring + STAGES * Cfg::kSlabFloats;
  unsigned long long* const full = reinterpret_cast<unsigned long long*>(a_tiles + WARPS * G * Cfg::kTileFloats);
  unsigned long long* const empty = full + STAGES;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n = a.n, k = a.k;
  // The (row block, column tile) pairs are walked in the L2-aware order of raster_tile: the CTAs
  // resident together (srb x scb of them, one wave) share srb row blocks of A and scb column tiles of
  // the dense operand (16 MB at k = 4096, scb = 4), and the column groups of a band reuse its rows of A
  // from L2.  With row blocks as the fast grid dimension each column-tile pass re-streamed all of A and
  // B came from HBM once per row block wave (dram__bytes 561 MB for 201 MB of operands at density 0.5,
  // profiles/r01_spmm_staged_d050_ncu_summary.txt).
  // (Worked out once by thread 0 into shared memory and re-read where needed: the kernel sits at its
  // register budget, and two more live registers across the slab loop spilled -- 3.39 vs 3.15 ms.)
  __shared__ int s_where[2];
  if (threadIdx.x == 0) {
    int rb, nt;
    raster_tile((int)blockIdx.x, row_blocks, col_tiles, srb, scb, rb, nt);
    s_where[0] = rb;
    s_where[1] = nt * NT;
  }
  const volatile int* const where = s_where;
#define row_block (where[0])
#define n0 (where[1])
  const int rep = blockIdx.z;
  const int nslabs = (k + KS - 1) / KS;

  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full + s, 1);
      mbar_init(empty + s, WARPS);
    }
    mbar_fence_init();
  }
  __syncthreads();

  if (warp == WARPS) {
    // ---------------- producer warp: stream B slabs into the ring ----------------
    const float* __restrict__ dense = a.dense + (size_t)rep * k * n + n0;
    const unsigned row_bytes = (unsigned)min(NT, n - n0) * 4u;
    for (int s = 0; s < nslabs; ++s) {
      const int st = s % STAGES, use = s / STAGES;
      if (use > 0) mbar_wait(empty + st, (use - 1) & 1);
      const int rows = min(KS, k - s * KS);
      if (lane == 0) mbar_arrive_expect_tx(full + st, rows * row_bytes);
      __syncwarp();
      if (lane < rows)
        bulk_g2s(b_ring + st * Cfg::kSlabFloats + lane * NT, dense + (size_t)(s * KS + lane) * n,
                 row_bytes, full + st);
    }
    return;
  }

  // ---------------- compute warps ----------------
  const int r_lane = lane / LPR, t_lane = lane % LPR;  // lane <-> (row in group, candidate slot)
  const unsigned row_lanes = ((1u << LPR) - 1u) << (r_lane * LPR);
  const float* __restrict__ vals = a.values + (size_t)rep * a.values_stride;
  const int* __restrict__ cols = a.column_indices;
  constexpr int kTile = Cfg::kTileFloats;  // KS x 4 values + KS column offsets per 4-row group
  float* const at = a_tiles + warp * (G * kTile);

  int pos[G], endp[G];
#pragma unroll
  for (int g = 0; g < G; ++g) {
    const int item = row_block * Cfg::RB + (warp * G + g) * R + r_lane;
    pos[g] = endp[g] = 0;
    if (item < a.m) {
      const int row = a.row_indices ? __ldg(a.row_indices + item) : item;
      pos[g] = __ldg(a.row_offsets + row);
      endp[g] = __ldg(a.row_offsets + row + 1);
    }
  }

  unsigned long long acc[G][R][NV4 * 2];
#pragma unroll
  for (int g = 0; g < G; ++g)
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
      for (int j = 0; j < NV4 * 2; ++j) acc[g][r][j] = 0ull;

  // Candidates for the next slab: up to 8*U entries per row starting at the row cursor.
  int pc[G][U];
  float pv[G][U];
  auto prefetch = [&]() {
#pragma unroll
    for (int g = 0; g < G; ++g)
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int idx = pos[g] + u * LPR + t_lane;
        const bool in = idx < endp[g];
        pc[g][u] = in ? __ldg(cols + idx) : 0x7fffffff;
        pv[g][u] = in ? __ldg(vals + idx) : 0.f;
      }
  };
  prefetch();

  for (int s = 0; s < nslabs; ++s) {
    const int st = s % STAGES, use = s / STAGES;
    const int k0 = s * KS, k1 = k0 + KS;
    int cnts[G];

    // -- build this slab's COMPACT tiles from the prefetched candidates: entry i of a group holds
    //    the four rows' values at the group's i-th occupied column, klist[i] that column's byte
    //    offset into the B slab.
#pragma unroll
    for (int g = 0; g < G; ++g)
#pragma unroll
      for (int q = 0; q < R / 4; ++q)
        reinterpret_cast<float4*>(at + g * kTile)[lane + 32 * q] = make_float4(0.f, 0.f, 0.f, 0.f);
    __syncwarp();
#pragma unroll
    for (int g = 0; g < G; ++g) {
      float* const tile = at + g * kTile;
      int* const klist = reinterpret_cast<int*>(tile + KS * R);
      unsigned bits = 0;
      int taken = 0;
      bool take[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        take[u] = pc[g][u] < k1;
        if (take[u]) bits |= 1u << (pc[g][u] - k0);
        taken += __popc(__ballot_sync(kFullMask, take[u]) & row_lanes);
      }
      pos[g] += taken;
      if (__any_sync(kFullMask, taken == LPR * U)) {
        // Rare (U is sized from the density): a row has more than 8*U entries in this slab.
        // Fall back to a dense [KS x 4] tile for this slab and consume synchronously.
        unsigned mask = __reduce_or_sync(kFullMask, bits);
#pragma unroll
        for (int u = 0; u < U; ++u)
          if (take[u]) tile[(pc[g][u] - k0) * R + r_lane] = pv[g][u];
        bool more = taken == LPR * U;
        while (__any_sync(kFullMask, more)) {
          const int idx = pos[g] + t_lane;
          const bool in = more && idx < endp[g];
          const int c = in ? __ldg(cols + idx) : 0x7fffffff;
          const bool tk = c < k1;
          unsigned b2 = 0;
          if (tk) {
            tile[(c - k0) * R + r_lane] = __ldg(vals + idx);
            b2 = 1u << (c - k0);
          }
          mask |= __reduce_or_sync(kFullMask, b2);
          const int got = __popc(__ballot_sync(kFullMask, tk) & row_lanes);
          pos[g] += got;
          more = got == LPR;
        }
        __syncwarp();
        // compact in place: entry rank(j) <- slot j (rank(j) <= j, ascending j keeps it safe
        // because the whole tile is read into registers before it is rewritten).
        float4 mine[R / 4];
#pragma unroll
        for (int q = 0; q < R / 4; ++q) mine[q] = reinterpret_cast<const float4*>(tile)[lane * (R / 4) + q];
        __syncwarp();
        if ((mask >> lane) & 1u) {
          const int rk = __popc(mask & ((1u << lane) - 1u));
#pragma unroll
          for (int q = 0; q < R / 4; ++q) reinterpret_cast<float4*>(tile)[rk * (R / 4) + q] = mine[q];
          klist[rk] = lane * (NT * 4);
        }
        cnts[g] = __popc(mask);
      } else {
        const unsigned mask = __reduce_or_sync(kFullMask, bits);
#pragma unroll
        for (int u = 0; u < U; ++u)
          if (take[u]) {
            const int rk = __popc(mask & ((1u << (pc[g][u] - k0)) - 1u));
            tile[rk * R + r_lane] = pv[g][u];
          }
        if ((mask >> lane) & 1u) klist[__popc(mask & ((1u << lane) - 1u))] = lane * (NT * 4);
        cnts[g] = __popc(mask);
      }
    }
    __syncwarp();
    prefetch();  // for slab s+1; the loads fly under the FMA loop below

    mbar_wait(full + st, use & 1);
    const char* const bs = reinterpret_cast<const char*>(b_ring + st * Cfg::kSlabFloats + lane * 4);
#pragma unroll
    for (int g = 0; g < G; ++g) {
      const float* const tile = at + g * kTile;
      const int* const klist = reinterpret_cast<const int*>(tile + KS * R);
      const int cnt = cnts[g];
      auto column = [&](const int i, const int koff) {
        float4 av[R / 4];
#pragma unroll
        for (int q = 0; q < R / 4; ++q) av[q] = reinterpret_cast<const float4*>(tile)[i * (R / 4) + q];
        ulonglong2 bv[NV4];
#pragma unroll
        for (int j = 0; j < NV4; ++j)
          bv[j] = *reinterpret_cast<const ulonglong2*>(bs + koff + j * 512);
#pragma unroll
        for (int q = 0; q < R / 4; ++q) {
          const unsigned long long a2[4] = {pack2(av[q].x, av[q].x), pack2(av[q].y, av[q].y),
                                            pack2(av[q].z, av[q].z), pack2(av[q].w, av[q].w)};
          // (Skipping the rows that do not have this column -- a warp-uniform test of the broadcast
          // value -- was measured and is slower: as predicated FFMA2 it saves no issue slot, as a
          // real branch the BSSY/BRA/BSYNC per row cost more than the 4 FFMA2 they jump over:
          // 2.74 / 3.75 ms against 2.44 / 3.18 ms at density 0.25 / 0.5.)
#pragma unroll
          for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int j = 0; j < NV4; ++j) {
              fma2(acc[g][q * 4 + r][2 * j], a2[r], bv[j].x);
              fma2(acc[g][q * 4 + r][2 * j + 1], a2[r], bv[j].y);
            }
        }
      };
      int i = 0;
      for (; i + 2 <= cnt; i += 2) {
        const int2 ko = *reinterpret_cast<const int2*>(klist + i);
        column(i, ko.x);
        column(i + 1, ko.y);
      }
      if (i < cnt) column(i, klist[i]);
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(empty + st);
  }

  // ---------------- epilogue: bias, ReLU, coalesced float4 stores ----------------
#pragma unroll
  for (int g = 0; g < G; ++g)
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int item = row_block * Cfg::RB + (warp * G + g) * R + r;
      if (item >= a.m) continue;
      const int row = a.row_indices ? __ldg(a.row_indices + item) : item;
      // x - x is 0 for finite x and NaN otherwise; one sum per row finds any non-finite accumulator
      // (an Inf born of finite operands also lands here: the redo then reproduces it).
      float probe = 0.f;
#pragma unroll
      for (int j = 0; j < NV4 * 2; ++j) {
        float lo, hi;
        unpack2(acc[g][r][j], lo, hi);
        probe += (lo - lo) + (hi - hi);
      }
      const float bias = a.bias ? __ldg(a.bias + row) : 0.f;
      float* __restrict__ out = a.out + ((size_t)rep * a.m + row) * n + n0 + lane * 4;
      if (__any_sync(kFullMask, probe != 0.f)) {
        spmm_staged_redo_row<NV4>(a.dense + (size_t)rep * k * n, vals, cols, __ldg(a.row_offsets + row),
                                  __ldg(a.row_offsets + row + 1), n, n0, lane, bias, a.relu, out);
        continue;
      }
#pragma unroll
      for (int j = 0; j < NV4; ++j) {
        if (n0 + lane * 4 + j * 128 >= n) continue;
        float4 o;
        unpack2(acc[g][r][2 * j], o.x, o.y);
        unpack2(acc[g][r][2 * j + 1], o.z, o.w);
        o.x += bias; o.y += bias; o.z += bias; o.w += bias;
        if (a.relu) {
          o.x = fmaxf(o.x, 0.f); o.y = fmaxf(o.y, 0.f); o.z = fmaxf(o.z, 0.f); o.w = fmaxf(o.w, 0.f);
        }
        *reinterpret_cast<float4*>(out + j * 128) = o;
      }
    }
}

#undef row_block
#undef n0

// ------------------------------------------------------------------------------------ host
inline bool spmm_staged_applicable(const SpmmArgs& a, int replicas) {
  if (a.value_perm != nullptr) return false;   // the staged kernel sits at its register budget; permuted values go through the gather kernel
  if (a.n % 4 != 0 || a.n < 96) return false;
  if (!aligned16(a.dense) || !aligned16(a.out)) return false;
  if (a.k < 64 || a.nnz == 0) return false;
  // Small problems are launch-bound; staging cannot pay for its prologue there.
  return 2.0 * a.nnz * a.n * replicas >= 5.0e7;
}

extern std::atomic<int> g_spmm_scb;

template <class Cfg>
inline int launch_spmm_staged_cfg(const SpmmArgs& a, int replicas, const StagedGuard& guard,
                                  cudaStream_t stream) {
  static bool configured[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 0 && dev < 64 && !configured[dev]) {
    cudaError_t e = cudaFuncSetAttribute(spmm_staged_kernel<Cfg>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)Cfg::kSmemBytes);
    if (e != cudaSuccess) return cuda_fail(e, "tsp_spmm_f32 (staged smem attribute)");
    configured[dev] = true;
  }
  const int row_blocks = ceil_div(a.m, Cfg::RB), col_tiles = ceil_div(a.n, Cfg::NT);
  // one wave of CTAs per (band, column group): scb column tiles x srb row blocks
  const int resident = sm_count() * (Cfg::kSmemBytes <= 110 * 1024 ? 2 : 1);
  const int want_scb = g_spmm_scb.load() > 0 ? g_spmm_scb.load() : 4;   // benchmark knob "spmm_scb"
  const int scb = col_tiles < want_scb ? col_tiles : want_scb;
  const int srb = (resident + scb - 1) / scb;
  dim3 grid((unsigned)((long long)row_blocks * col_tiles), 1, replicas);
  spmm_staged_kernel<Cfg><<<grid, Cfg::THREADS, Cfg::kSmemBytes, stream>>>(a, guard, row_blocks, col_tiles, srb, scb);
  return 0;
}

// Returns TSP_OK / an error code.  The sortedness verdict comes from the topology's plan (plan.cuh):
// guard.flag == null when the host already knows the columns are sorted, else the plan's device
// word -- the caller then queues the guarded gather launch behind this one.
// rows_per_group: 0 = pick from the density, 4 or 8 = forced (benchmark knob).
inline int launch_spmm_staged(const SpmmArgs& a, int replicas, cudaStream_t stream, const StagedGuard& guard,
                              int rows_per_group = 0) {
  // Candidates prefetched per row per slab = LPR*U; sized so that a row's count in a 32-column
  // slab (mean avg = 32*density) almost never exceeds it.
  const double density = (double)a.nnz / ((double)a.m * (double)a.k);
  const double avg = 32.0 * density;
  const double want = avg + 3.0 * sqrt(avg) + 1.0;
  // Eight merged rows pay once nearly every column of a slab is occupied anyway.
  const int R = rows_per_group ? rows_per_group : (density >= 0.45 && a.n > 128 ? 8 : 4);
  int rc;
  if (R == 8) {
    if (want <= 24.0) rc = launch_spmm_staged_cfg<StagedCfg<2, 8, 1, 3, 7, 6>>(a, replicas, guard, stream);
    else rc = launch_spmm_staged_cfg<StagedCfg<2, 8, 1, 3, 7, 8>>(a, replicas, guard, stream);
  } else {
    int U = (int)ceil(want / 8.0);
    U = U < 1 ? 1 : (U > 4 ? 4 : U);
    // 256-column tiles (2 CTAs per SM) unless they leave the machine half empty: a column-sharded SpMM
    // (n = 512 per rank at 8 GPUs) has 74 x 2 = 148 such tiles for 296 CTA slots; 128-column tiles double
    // the CTA count at the same bytes per FMA.
    const long long wide_ctas = (long long)ceil_div(a.m, 56) * ceil_div(a.n, 256) * replicas;
    if (a.n > 128 && wide_ctas >= 2LL * sm_count()) {
      switch (U) {
        case 1: rc = launch_spmm_staged_cfg<StagedCfg<2, 4, 2, 3, 7, 1>>(a, replicas, guard, stream); break;
        case 2: rc = launch_spmm_staged_cfg<StagedCfg<2, 4, 2, 3, 7, 2>>(a, replicas, guard, stream); break;
        case 3: rc = launch_spmm_staged_cfg<StagedCfg<2, 4, 2, 3, 7, 3>>(a, replicas, guard, stream); break;
        default: rc = launch_spmm_staged_cfg<StagedCfg<2, 4, 2, 3, 7, 4>>(a, replicas, guard, stream); break;
      }
    } else {
      switch (U) {
        case 1: rc = launch_spmm_staged_cfg<StagedCfg<1, 4, 2, 4, 8, 1>>(a, replicas, guard, stream); break;
        case 2: rc = launch_spmm_staged_cfg<StagedCfg<1, 4, 2, 4, 8, 2>>(a, replicas, guard, stream); break;
        case 3: rc = launch_spmm_staged_cfg<StagedCfg<1, 4, 2, 4, 8, 3>>(a, replicas, guard, stream); break;
        default: rc = launch_spmm_staged_cfg<StagedCfg<1, 4, 2, 4, 8, 4>>(a, replicas, guard, stream); break;
      }
    }
  }
  if (rc != 0) return rc;
  TSP_CHECK_LAUNCH("tsp_spmm_f32 (staged)");
  return 0;
}

}  // namespace tsp
