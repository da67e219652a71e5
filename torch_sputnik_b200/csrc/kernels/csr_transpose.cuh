// csr_transpose.cuh -- stable CSR -> CSC (= CSR of the transpose), bit-exact, no atomics in the
// placement.
//
// Replaces cusparseCsr2cscEx2(..., CUSPARSE_CSR2CSC_ALG1) as called from
// src/transpose_cuda.cu:90-99.  ALG1 is a stable sort of the nonzeros by column, so inside
// every output row the entries keep ascending original-row order; an atomic scatter would not
// reproduce that.  Here: a stable LSD radix sort of the flat nonzero index keyed by column
// (ceil(log2 n) bits in ceil(bits/8) passes), each pass = per-tile digit histogram -> one
// exclusive scan of the [digit][tile] table -> stable in-tile ranking with warp match + ordered
// per-warp counters.  A last kernel turns the sorted order into values_t / column_indices_t /
// row_offsets_t / perm.
#pragma once
#include <atomic>
#include <mutex>

#include "common.cuh"

namespace tsp {

constexpr int kSortThreads = 256;
constexpr int kSortWarps = kSortThreads / 32;
constexpr int kSortItems = 16;                          // per thread
constexpr int kSortTile = kSortThreads * kSortItems;    // 4096 keys per CTA
constexpr int kSortWarpSeg = 32 * kSortItems;           // 512 consecutive keys per warp
constexpr int kMaxRadix = 256;

struct SortPlan {
  int bits, passes, digit_bits, radix, tiles;
};

inline SortPlan make_sort_plan(int n, int nnz) {
  SortPlan p;
  int bits = 1;
  while (bits < 31 && (1LL << bits) < (long long)n) ++bits;
  p.bits = bits;
  p.passes = (bits + 7) / 8;
  p.digit_bits = (bits + p.passes - 1) / p.passes;
  p.radix = 1 << p.digit_bits;
  p.tiles = nnz > 0 ? (nnz + kSortTile - 1) / kSortTile : 0;
  return p;
}

// hist[d * tiles + tile] = number of keys of `tile` whose digit is d.
__global__ void __launch_bounds__(kSortThreads)
radix_hist_kernel(const int* __restrict__ keys, int nnz, int shift, int radix, int tiles,
                  int* __restrict__ hist) {
  __shared__ int s_hist[kMaxRadix];
  for (int d = threadIdx.x; d < radix; d += kSortThreads) s_hist[d] = 0;
  __syncthreads();
  const int tile = blockIdx.x;
  const int base = tile * kSortTile;
  const int mask = radix - 1;
#pragma unroll 4
  for (int i = 0; i < kSortItems; ++i) {
    const int e = base + i * kSortThreads + threadIdx.x;
    if (e < nnz) atomicAdd(&s_hist[(__ldg(keys + e) >> shift) & mask], 1);
  }
  __syncthreads();
  for (int d = threadIdx.x; d < radix; d += kSortThreads) hist[(size_t)d * tiles + tile] = s_hist[d];
}

// In-place exclusive scan of `count` ints by one CTA (count = radix * tiles; a few 100k at most).
__global__ void __launch_bounds__(1024) exclusive_scan_kernel(int* __restrict__ data, int count) {
  __shared__ int s_warp[32];
  __shared__ int s_carry;
  const int tid = threadIdx.x;
  const int per = (count + 1023) / 1024;
  const int beg = min(count, tid * per), end = min(count, beg + per);
  int sum = 0;
  for (int i = beg; i < end; ++i) sum += data[i];
  // block exclusive scan of `sum`
  int incl = sum;
#pragma unroll
  for (int s = 1; s < 32; s <<= 1) {
    const int t = __shfl_up_sync(kFullMask, incl, s);
    if ((tid & 31) >= s) incl += t;
  }
  if ((tid & 31) == 31) s_warp[tid >> 5] = incl;
  __syncthreads();
  if (tid < 32) {
    int w = s_warp[tid];
    int wi = w;
#pragma unroll
    for (int s = 1; s < 32; s <<= 1) {
      const int t = __shfl_up_sync(kFullMask, wi, s);
      if (tid >= s) wi += t;
    }
    s_warp[tid] = wi - w;
    if (tid == 31) s_carry = wi;
  }
  __syncthreads();
  int run = s_warp[tid >> 5] + incl - sum;
  for (int i = beg; i < end; ++i) {
    const int v = data[i];
    data[i] = run;
    run += v;
  }
  (void)s_carry;
}

// Stable scatter of one radix pass.  Order inside a tile: warp-major, then round, then lane.
// payload_in == nullptr means the identity payload (first pass).
__global__ void __launch_bounds__(kSortThreads)
radix_scatter_kernel(const int* __restrict__ keys_in, const int* __restrict__ payload_in, int nnz,
                     int shift, int radix, int tiles, const int* __restrict__ hist_scanned,
                     int* __restrict__ keys_out, int* __restrict__ payload_out) {
  __shared__ int s_cnt[kSortWarps][kMaxRadix];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tile = blockIdx.x;
  const int mask = radix - 1;
  for (int d = lane; d < radix; d += 32) s_cnt[warp][d] = 0;
  __syncwarp();

  const int seg = tile * kSortTile + warp * kSortWarpSeg;
  int key[kSortItems], rank[kSortItems];
  const unsigned lt = (1u << lane) - 1u;
#pragma unroll
  for (int i = 0; i < kSortItems; ++i) {
    const int e = seg + i * 32 + lane;
    const bool in = e < nnz;
    key[i] = in ? __ldg(keys_in + e) : 0;
    const int d = in ? ((key[i] >> shift) & mask) : radix;  // radix = "no key" class
    const unsigned peers = __match_any_sync(kFullMask, d);
    int before = 0;
    if (in) before = s_cnt[warp][d];
    __syncwarp();
    if (in && (peers & lt) == 0) s_cnt[warp][d] = before + __popc(peers);
    __syncwarp();
    rank[i] = before + __popc(peers & lt);
  }
  __syncthreads();
  // Exclusive prefix over warps per digit, plus the global base of (digit, tile).
  for (int d = threadIdx.x; d < radix; d += kSortThreads) {
    int run = __ldg(hist_scanned + (size_t)d * tiles + tile);
#pragma unroll
    for (int w = 0; w < kSortWarps; ++w) {
      const int c = s_cnt[w][d];
      s_cnt[w][d] = run;
      run += c;
    }
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < kSortItems; ++i) {
    const int e = seg + i * 32 + lane;
    if (e < nnz) {
      const int d = (key[i] >> shift) & mask;
      const int pos = s_cnt[warp][d] + rank[i];
      keys_out[pos] = key[i];
      payload_out[pos] = payload_in ? __ldg(payload_in + e) : e;
    }
  }
}

struct TransposeFinishArgs {
  int m, n, nnz, replicas;
  const int* sorted_cols;   // [nnz] ascending
  const int* sorted_src;    // [nnz] original flat index of each sorted entry
  const int* row_offsets;   // [m+1] of the input
  const float* values;      // [replicas, nnz]
  float* values_t;          // [replicas, nnz]
  int* row_offsets_t;       // [n+1]
  int* column_indices_t;    // [nnz]
  int* perm;                // may be null
};

__device__ __forceinline__ int upper_row(const int* __restrict__ row_offsets, int m, int j) {
  int lo = 0, hi = m;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (__ldg(row_offsets + mid) <= j) lo = mid; else hi = mid;
  }
  return lo;
}

// One thread per sorted position p (plus one extra thread p == nnz closing the offsets).
__global__ void __launch_bounds__(256) transpose_finish_kernel(const TransposeFinishArgs a) {
  const int p = blockIdx.x * 256 + threadIdx.x;
  if (p > a.nnz) return;
  const int c = p < a.nnz ? __ldg(a.sorted_cols + p) : a.n;
  const int cprev = p > 0 ? __ldg(a.sorted_cols + p - 1) : -1;
  // Output rows cprev+1 .. c start at p (empty output rows collapse onto the same offset).
  for (int r = cprev + 1; r <= c; ++r) a.row_offsets_t[r] = p;
  if (p == a.nnz) return;
  const int src = __ldg(a.sorted_src + p);
  a.column_indices_t[p] = upper_row(a.row_offsets, a.m, src);
  if (a.perm) a.perm[p] = src;
  for (int r = 0; r < a.replicas; ++r)
    a.values_t[(size_t)r * a.nnz + p] = __ldg(a.values + (size_t)r * a.nnz + src);
}

// ------------------------------------------------------------------------------------------------
// Counting path (n <= kCountMaxCols): the input is already in row order, so a stable sort by column is
// ONE counting pass with radix = n -- three launches whatever n is:
//   count:  chunk s (a contiguous range of the flat nonzero index) counts its columns in shared memory
//           and writes row s of table[S][n];
//   offsets: per column, an exclusive prefix over the chunks (in place) and the column total; the totals
//           are scanned across column groups with a chained scan (ticket order, so a group only ever
//           waits for groups that are already running) -> row_offsets_t;
//   place:  chunk s keeps the next free output slot of every column in shared memory and walks its
//           range in flat order; inside a round the warps update the slots in warp order, so equal
//           columns keep ascending flat index = the stable order of cuSPARSE's ALG1.
constexpr int kCountThreads = 256;
constexpr int kCountWarps = kCountThreads / 32;
constexpr int kCountItems = 4;                                     // consecutive 32-entry groups per warp per round
constexpr int kCountRound = kCountThreads * kCountItems;           // 1024 entries per round
constexpr int kCountMaxCols = 49152;                               // n ints of shared memory per CTA
constexpr int kCountGroupCols = 256;                               // columns per CTA of the offsets kernel

struct CountPlan {
  bool use;
  int chunks;        // S
  int chunk_len;     // entries per chunk (multiple of kCountRound)
  int groups;        // column groups of the offsets kernel
};

// waves: cap on the chunk count in units of the SM count (0 = default).
inline CountPlan make_count_plan(int n, int nnz, int waves = 0) {
  CountPlan p{false, 0, 0, 0};
  if (n <= 0 || nnz <= 0 || n > kCountMaxCols) return p;
  if (waves <= 0) waves = 2;
  int chunks = min(ceil_div(nnz, 2 * kCountRound), waves * sm_count());
  // the table costs 4 * chunks * n bytes, written once and read twice, and every CTA clears / fills n
  // counters: a wide, nearly empty matrix (table larger than ~4 ints per nonzero) sorts faster by radix
  // passes (512 x 40000 at 2 %: 93 us radix, 112 - 148 us counting)
  if ((long long)chunks * n > 4LL * nnz) return p;
  p.chunk_len = ceil_div(ceil_div(nnz, chunks), kCountRound) * kCountRound;
  p.chunks = ceil_div(nnz, p.chunk_len);
  p.groups = ceil_div(n, kCountGroupCols);
  p.use = true;
  return p;
}

// ints of workspace: table[S][n], then the chained-scan words (ticket + one per group)
inline size_t count_plan_ints(const CountPlan& p, int n) {
  return (size_t)p.chunks * n + 64 + p.groups;
}

__global__ void __launch_bounds__(kCountThreads)
transpose_count_kernel(const int* __restrict__ cols, int n, int nnz, int chunk_len, int* __restrict__ table,
                       int* __restrict__ chain, int chain_words) {
  extern __shared__ int s_cnt[];
  if (blockIdx.x == 0)
    for (int i = threadIdx.x; i < chain_words; i += kCountThreads) chain[i] = 0;
  for (int c = threadIdx.x; c < n; c += kCountThreads) s_cnt[c] = 0;
  __syncthreads();
  const int beg = blockIdx.x * chunk_len, end = min(nnz, beg + chunk_len);
#pragma unroll 4
  for (int e = beg + threadIdx.x; e < end; e += kCountThreads) atomicAdd(&s_cnt[__ldg(cols + e)], 1);
  __syncthreads();
  int* __restrict__ row = table + (size_t)blockIdx.x * n;
  for (int c = threadIdx.x; c < n; c += kCountThreads) row[c] = s_cnt[c];
}

__global__ void __launch_bounds__(kCountGroupCols)
transpose_offsets_kernel(int n, int nnz, int chunks, int* __restrict__ table, int* __restrict__ chain,
                         int* __restrict__ row_offsets_t) {
  __shared__ int s_group, s_prev, s_warp[kCountGroupCols / 32];
  if (threadIdx.x == 0) s_group = atomicAdd(chain, 1);
  __syncthreads();
  const int group = s_group;
  const int c = group * kCountGroupCols + threadIdx.x;
  int run = 0;
  if (c < n) {
    int* __restrict__ col = table + c;
    for (int s0 = 0; s0 < chunks; s0 += 8) {
      int v[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = s0 + j < chunks ? col[(size_t)(s0 + j) * n] : 0;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        if (s0 + j < chunks) col[(size_t)(s0 + j) * n] = run;
        run += v[j];
      }
    }
  }
  // exclusive scan of the column totals inside the group
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int incl = run;
#pragma unroll
  for (int s = 1; s < 32; s <<= 1) {
    const int t = __shfl_up_sync(kFullMask, incl, s);
    if (lane >= s) incl += t;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  if (threadIdx.x == 0) {
    int total = 0;
#pragma unroll
    for (int w = 0; w < kCountGroupCols / 32; ++w) {
      const int t = s_warp[w];
      s_warp[w] = total;
      total += t;
    }
    // chained scan: word 64 + g holds (inclusive prefix of group g) + 1, 0 = not published yet
    volatile int* ready = chain + 64;
    int prev = 0;
    if (group > 0) {
      int v;
      while ((v = ready[group - 1]) == 0) {}
      prev = v - 1;
    }
    ready[group] = prev + total + 1;
    s_prev = prev;
  }
  __syncthreads();
  if (c < n) row_offsets_t[c] = s_prev + s_warp[warp] + incl - run;
  if (c == n - 1) row_offsets_t[n] = nnz;
}

struct TransposePlaceArgs {
  int m, n, nnz, replicas, chunk_len;
  const int* row_offsets;
  const int* column_indices;
  const float* values;          // [replicas, nnz]
  const int* table;             // [S][n] exclusive prefix over the chunks
  const int* row_offsets_t;     // [n+1]
  float* values_t;
  int* column_indices_t;
  int* perm;                    // may be null
};

__global__ void __launch_bounds__(kCountThreads) transpose_place_kernel(const TransposePlaceArgs a) {
  extern __shared__ int s_pos[];
  __shared__ int s_rows[2];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int beg = blockIdx.x * a.chunk_len, end = min(a.nnz, beg + a.chunk_len);
  {
    const int* __restrict__ trow = a.table + (size_t)blockIdx.x * a.n;
    for (int c = threadIdx.x; c < a.n; c += kCountThreads) s_pos[c] = __ldg(a.row_offsets_t + c) + __ldg(trow + c);
    if (threadIdx.x < 2) s_rows[threadIdx.x] = upper_row(a.row_offsets, a.m, threadIdx.x == 0 ? beg : end - 1);
  }
  __syncthreads();
  const int row_lo = s_rows[0], row_hi = s_rows[1] + 1;
  const unsigned lt = (1u << lane) - 1u;

  for (int base = beg; base < end; base += kCountRound) {
    int col[kCountItems], before[kCountItems];
    unsigned peers[kCountItems];
#pragma unroll
    for (int i = 0; i < kCountItems; ++i) {
      const int e = base + (warp * kCountItems + i) * 32 + lane;
      col[i] = e < end ? __ldg(a.column_indices + e) : -1;
      peers[i] = __match_any_sync(kFullMask, col[i]);
      before[i] = 0;
    }
    for (int w = 0; w < kCountWarps; ++w) {
      if (warp == w) {
#pragma unroll
        for (int i = 0; i < kCountItems; ++i) {
          if (col[i] >= 0 && (peers[i] & lt) == 0) {
            before[i] = s_pos[col[i]];
            s_pos[col[i]] = before[i] + __popc(peers[i]);
          }
          __syncwarp();
        }
      }
      __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < kCountItems; ++i) {
      const int first = __shfl_sync(kFullMask, before[i], __ffs(peers[i]) - 1);
      const int e = base + (warp * kCountItems + i) * 32 + lane;
      if (col[i] < 0) continue;
      const int pos = first + __popc(peers[i] & lt);
      int lo = row_lo, hi = row_hi;   // row of e: row_offsets[lo] <= e < row_offsets[hi]
      while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(a.row_offsets + mid) <= e) lo = mid; else hi = mid;
      }
      a.column_indices_t[pos] = lo;
      if (a.perm) a.perm[pos] = e;
      for (int r = 0; r < a.replicas; ++r)
        a.values_t[(size_t)r * a.nnz + pos] = __ldg(a.values + (size_t)r * a.nnz + e);
    }
  }
}

// V: float, or unsigned short for any 16-bit storage type (a pure move).
// masks.prefix != null: many masks in one launch -- replica blockIdx.y uses perm + prefix[mask] (indices LOCAL
// to the mask) and the ragged value layout of MaskTable; nnz is then the largest mask's count.
// A CTA moves a RUN of ITEMS * blockDim consecutive outputs.  The permutation of a transpose walks the
// output column by column, so the 32-byte sector holding values[row, c .. c+7] is wanted again one output
// row (= one column's worth of nonzeros) later: with a long run per CTA that second request finds the
// sector in this SM's L1 instead of fetching it from L2 once per element.
template <typename V, int ITEMS>
__global__ void __launch_bounds__(1024)
permute_values_kernel(int nnz, const V* __restrict__ values, const int* __restrict__ perm,
                      V* __restrict__ values_t, const MaskTable masks) {
  const MaskView mv = mask_view(masks, blockIdx.y, nnz);
  const int run = ITEMS * blockDim.x;
  const int p0 = blockIdx.x * run + threadIdx.x;
  if (p0 - (int)threadIdx.x >= mv.nnz) return;
  const V* __restrict__ src = values + (size_t)mv.values;
  V* __restrict__ dst = values_t + (size_t)mv.values;
  const int* __restrict__ pm = perm + mv.cols;
  int idx[ITEMS];
#pragma unroll
  for (int i = 0; i < ITEMS; ++i) {
    const int p = p0 + i * blockDim.x;
    idx[i] = p < mv.nnz ? __ldg(pm + p) : -1;
  }
  V v[ITEMS];
#pragma unroll
  for (int i = 0; i < ITEMS; ++i)
    if (idx[i] >= 0) v[i] = __ldg(src + idx[i]);
#pragma unroll
  for (int i = 0; i < ITEMS; ++i)
    if (idx[i] >= 0) dst[p0 + i * blockDim.x] = v[i];
}

extern std::atomic<int> g_permute_cfg;   // benchmark knob "permute": 100 * items + threads / 32 (0 = default)

template <typename V>
inline void launch_permute_values(int nnz, int replicas, const V* values, const int* perm, V* values_t,
                                  const MaskTable& masks, cudaStream_t stream) {
  int cfg = g_permute_cfg.load();
  if (cfg <= 0) cfg = nnz >= (1 << 16) ? 816 : 108;
  const int items = cfg / 100, threads = (cfg % 100) * 32;
  const dim3 grid(ceil_div(nnz, items * threads), replicas);
  switch (items) {
    case 1: permute_values_kernel<V, 1><<<grid, threads, 0, stream>>>(nnz, values, perm, values_t, masks); break;
    case 2: permute_values_kernel<V, 2><<<grid, threads, 0, stream>>>(nnz, values, perm, values_t, masks); break;
    case 4: permute_values_kernel<V, 4><<<grid, threads, 0, stream>>>(nnz, values, perm, values_t, masks); break;
    case 16: permute_values_kernel<V, 16><<<grid, threads, 0, stream>>>(nnz, values, perm, values_t, masks); break;
    default: permute_values_kernel<V, 8><<<grid, threads, 0, stream>>>(nnz, values, perm, values_t, masks); break;
  }
}

// ------------------------------------------------------------------------------------------------
// Tiled permutation view (plans): the gather above asks L2 for one 32-byte sector per 4-byte value --
// the request rate, not the bytes, is what bounds it (C4: 0.77 ms for 1.7 GB).  For a cached transpose
// the outputs are cut into tiles of kPermTile consecutive positions and, once per topology, each tile's
// elements are listed in ASCENDING SOURCE order (one stable radix pass over the tile id of the inverse
// permutation).  A tile of a transpose covers a block of columns, so in that order the sources form one
// short contiguous run per matrix row: neighbouring threads read neighbouring values, drop them into the
// tile's slot in shared memory, and the tile leaves as fully coalesced stores.
constexpr int kPermTileBits = 15;
constexpr int kPermTile = 1 << kPermTileBits;     // outputs per tile (128 KB of fp32 in shared memory)
constexpr int kPermThreads = 1024;
constexpr int kPermItems = kPermTile / kPermThreads;

__global__ void __launch_bounds__(256) inverse_perm_kernel(int nnz, const int* __restrict__ perm, int* __restrict__ inv) {
  const int p = blockIdx.x * 256 + threadIdx.x;
  if (p < nnz) inv[__ldg(perm + p)] = p;
}

// sorted_dst / sorted_src: the (destination, source) pairs ordered by (tile of destination, source)
__global__ void __launch_bounds__(256)
perm_view_finish_kernel(int nnz, const int* __restrict__ sorted_dst, const int* __restrict__ sorted_src,
                        int* __restrict__ view_src, unsigned short* __restrict__ view_loc) {
  const int j = blockIdx.x * 256 + threadIdx.x;
  if (j >= nnz) return;
  view_src[j] = __ldg(sorted_src + j);
  view_loc[j] = (unsigned short)(__ldg(sorted_dst + j) & (kPermTile - 1));
}

template <typename V>
__global__ void __launch_bounds__(kPermThreads, 1)
permute_tiled_kernel(int nnz, int replicas, int reps_per_cta, const V* __restrict__ values,
                     const int* __restrict__ view_src, const unsigned short* __restrict__ view_loc,
                     V* __restrict__ values_t) {
  extern __shared__ __align__(16) unsigned char s_raw[];
  V* const s_tile = reinterpret_cast<V*>(s_raw);
  const int t0 = blockIdx.x * kPermTile;
  const int cnt = min(kPermTile, nnz - t0);
  int src[kPermItems];
  unsigned loc[kPermItems / 2];
#pragma unroll
  for (int i = 0; i < kPermItems; i += 2) {
    const int j0 = i * kPermThreads + threadIdx.x, j1 = j0 + kPermThreads;
    src[i] = j0 < cnt ? __ldg(view_src + t0 + j0) : -1;
    src[i + 1] = j1 < cnt ? __ldg(view_src + t0 + j1) : -1;
    const unsigned l0 = j0 < cnt ? __ldg(view_loc + t0 + j0) : 0u;
    const unsigned l1 = j1 < cnt ? __ldg(view_loc + t0 + j1) : 0u;
    loc[i / 2] = l0 | (l1 << 16);
  }
  const int r0 = blockIdx.y * reps_per_cta, r1 = min(replicas, r0 + reps_per_cta);
  for (int r = r0; r < r1; ++r) {
    const V* __restrict__ vin = values + (size_t)r * nnz;
    V* __restrict__ vout = values_t + (size_t)r * nnz + t0;
#pragma unroll
    for (int b = 0; b < kPermItems; b += 8) {
      V v[8];
#pragma unroll
      for (int i = 0; i < 8; ++i)
        if (src[b + i] >= 0) v[i] = __ldg(vin + src[b + i]);
#pragma unroll
      for (int i = 0; i < 8; ++i)
        if (src[b + i] >= 0) s_tile[(loc[(b + i) / 2] >> (16 * ((b + i) & 1))) & 0xffffu] = v[i];
    }
    __syncthreads();
#pragma unroll 8
    for (int i = 0; i < kPermItems; ++i) {
      const int j = i * kPermThreads + threadIdx.x;
      if (j < cnt) vout[j] = s_tile[j];
    }
    __syncthreads();
  }
}

template <typename V>
inline int launch_permute_tiled(int nnz, int replicas, const V* values, const int* view_src,
                                const unsigned short* view_loc, V* values_t, cudaStream_t stream) {
  static std::once_flag once;
  std::call_once(once, [] {
    cudaFuncSetAttribute(permute_tiled_kernel<V>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         (int)(kPermTile * sizeof(V)));
  });
  const int tiles = ceil_div(nnz, kPermTile);
  int reps = (int)(((long long)tiles * replicas) / (4LL * sm_count()));
  reps = reps < 1 ? 1 : (reps > 16 ? 16 : reps);
  const dim3 grid(tiles, ceil_div(replicas, reps));
  permute_tiled_kernel<V><<<grid, kPermThreads, kPermTile * sizeof(V), stream>>>(nnz, replicas, reps, values, view_src,
                                                                                  view_loc, values_t);
  return 0;
}

__global__ void fill_int_kernel(int* __restrict__ dst, int count, int value) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < count) dst[i] = value;
}

}  // namespace tsp
