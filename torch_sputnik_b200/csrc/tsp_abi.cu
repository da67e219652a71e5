// tsp_abi.cu -- the C-ABI of libtsp.so (declared in include/tsp.h): argument validation,
// kernel-variant heuristics and launches.  Torch-free; the PyTorch shim (torch_shim.cpp) and
// any other host (ctypes, a C++ harness) call exactly these symbols.
#include "../../include/tsp.h"

#include <cmath>
#include <cstring>
#include <mutex>

#include "kernels/common.cuh"
#include "kernels/csr_softmax.cuh"
#include "kernels/csr_transpose.cuh"
#include "kernels/sddmm_gather.cuh"
#include "kernels/spmm_gather.cuh"
#include "kernels/spmm_staged.cuh"
#include "kernels/spmm_rowsplit.cuh"
#include "kernels/sddmm_sampled.cuh"
#include "kernels/sddmm_tiled3.cuh"
#include "kernels/sddmm_rows.cuh"
#include "kernels/attention_rows.cuh"
#include "kernels/band.cuh"
#include "kernels/plan.cuh"

namespace tsp {

static thread_local char t_error[512] = "";
std::atomic<unsigned long long> g_launches{0};
static std::atomic<int> g_spmm_variant{0};
std::atomic<int> g_spmm_warps{0};
std::atomic<int> g_sddmm_warps{0};
std::atomic<int> g_attn_warps{0};
std::atomic<int> g_band_warps{0};
std::atomic<int> g_transpose_variant{0};
std::atomic<int> g_permute_cfg{0};
std::atomic<int> g_permute_tiled{0};
std::atomic<int> g_sddmm_b16{0};
std::atomic<int> g_sddmm_tiled3{0};
// 16-bit SDDMM: native sampled kernel below this density, widened fp32 tiled kernel above
// (4096^3: 2.70 vs 2.96 ms at 35 %, 3.76 vs 2.96 at 50 %)
constexpr double kSddmmB16SampledMaxDensity = 0.4;
std::atomic<int> g_spmm_scb{0};
static std::atomic<int> g_sddmm_variant{0};
static std::atomic<int> g_spmm_rowsplit_max_n{0};   // fp32 operands up to this many columns take spmm_rowsplit
static std::atomic<int> g_sddmm_bn{0};
static std::atomic<int> g_sddmm_p{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(t_error, sizeof(t_error), fmt, ap);
  va_end(ap);
}

int sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int v = 0;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0)
      v = 148;
    cached[dev] = v;
  }
  return cached[dev];
}

static int invalid(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(t_error, sizeof(t_error), fmt, ap);
  va_end(ap);
  return TSP_ERR_INVALID_ARGUMENT;
}

}  // namespace tsp

using namespace tsp;

extern "C" {

int tsp_version(void) { return TSP_VERSION; }

const char* tsp_last_error(void) { return t_error; }

unsigned long long tsp_launch_count(void) { return g_launches.load(); }

int tsp_set_variant(const char* op, int variant) {
  if (op && strcmp(op, "spmm") == 0) return g_spmm_variant.exchange(variant);
  if (op && strcmp(op, "spmm_rowsplit_max_n") == 0) return g_spmm_rowsplit_max_n.exchange(variant);
  if (op && strcmp(op, "spmm_warps") == 0) return g_spmm_warps.exchange(variant);
  if (op && strcmp(op, "sddmm_warps") == 0) return g_sddmm_warps.exchange(variant);
  if (op && strcmp(op, "spmm_scb") == 0) return g_spmm_scb.exchange(variant);
  if (op && strcmp(op, "band_warps") == 0) return g_band_warps.exchange(variant);
  if (op && strcmp(op, "attn_warps") == 0) return g_attn_warps.exchange(variant);
  if (op && strcmp(op, "transpose") == 0) return g_transpose_variant.exchange(variant);
  if (op && strcmp(op, "permute") == 0) return g_permute_cfg.exchange(variant);
  if (op && strcmp(op, "permute_tiled") == 0) return g_permute_tiled.exchange(variant);
  if (op && strcmp(op, "sddmm_b16") == 0) return g_sddmm_b16.exchange(variant);
  if (op && strcmp(op, "sddmm_tiled3") == 0) return g_sddmm_tiled3.exchange(variant);
  if (op && strcmp(op, "sddmm") == 0) return g_sddmm_variant.exchange(variant);
  if (op && strcmp(op, "sddmm_bn") == 0) return g_sddmm_bn.exchange(variant);
  if (op && strcmp(op, "sddmm_p") == 0) return g_sddmm_p.exchange(variant);
  return -1;
}

// ------------------------------------------------------------------------------------ plans
static void ephemeral_plan(tsp_plan* p, int m, int n, int nnz, const int* row_offsets, const int* column_indices) {
  p->m = m; p->n = n; p->nnz = nnz;
  p->row_offsets = row_offsets;
  p->column_indices = column_indices;
  p->ephemeral = true;
}

int tsp_plan_create(int m, int n, int nnz, const int* row_offsets, const int* column_indices, void* stream_,
                    tsp_plan** plan_out) {
  if (!plan_out) return invalid("tsp_plan_create: null plan_out");
  *plan_out = nullptr;
  if (m < 0 || n < 0 || nnz < 0) return invalid("tsp_plan_create: negative dimension (m=%d n=%d nnz=%d)", m, n, nnz);
  if (!row_offsets || (nnz > 0 && !column_indices)) return invalid("tsp_plan_create: null pointer argument");
  tsp_plan* p = new tsp_plan();
  p->m = m; p->n = n; p->nnz = nnz;
  p->row_offsets = row_offsets;
  p->column_indices = column_indices;
  const int rc = plan_init_persistent(p, static_cast<cudaStream_t>(stream_));
  if (rc != 0) {
    plan_destroy(p);
    return rc;
  }
  *plan_out = p;
  return TSP_OK;
}

void tsp_plan_destroy(tsp_plan* plan) { plan_destroy(plan); }

int tsp_plan_sorted(tsp_plan* plan) { return plan ? plan_sorted_state(plan) : -1; }

// Tiled view of plan->perm for tsp_plan_permute_values_* (kernels/csr_transpose.cuh): the pairs
// (destination, source) in order of (tile of destination, source) = a stable radix sort of the inverse
// permutation by its tile bits.  Queued behind the transpose on the same stream; topologies smaller than
// two tiles keep the plain gather.
static int build_permutation_view(tsp_plan* plan, cudaStream_t stream) {
  const int nnz = plan->nnz;
  if (nnz < 2 * kPermTile) return TSP_OK;
  const SortPlan sp = make_sort_plan(ceil_div(nnz, kPermTile), nnz);
  const size_t seg = (sizeof(int) * (size_t)nnz + 255) & ~size_t(255);
  char* ws = nullptr;
  cudaError_t e = cudaMalloc(reinterpret_cast<void**>(&ws), 5 * seg + sizeof(int) * (size_t)sp.radix * sp.tiles + 256);
  if (e == cudaSuccess) e = cudaMalloc(reinterpret_cast<void**>(&plan->pv_src), sizeof(int) * (size_t)nnz);
  if (e == cudaSuccess) e = cudaMalloc(reinterpret_cast<void**>(&plan->pv_loc), sizeof(unsigned short) * (size_t)nnz);
  if (e != cudaSuccess) {   // no view: the gather still works
    (void)cudaGetLastError();
    cudaFree(ws); cudaFree(plan->pv_src); cudaFree(plan->pv_loc);
    plan->pv_src = nullptr; plan->pv_loc = nullptr;
    return TSP_OK;
  }
  int* inv = reinterpret_cast<int*>(ws);
  int* keys[2] = {reinterpret_cast<int*>(ws + seg), reinterpret_cast<int*>(ws + 2 * seg)};
  int* pay[2] = {reinterpret_cast<int*>(ws + 3 * seg), reinterpret_cast<int*>(ws + 4 * seg)};
  int* hist = reinterpret_cast<int*>(ws + 5 * seg);
  inverse_perm_kernel<<<ceil_div(nnz, 256), 256, 0, stream>>>(nnz, plan->perm, inv);
  const int* kin = inv;
  const int* pin = nullptr;
  for (int pass = 0; pass < sp.passes; ++pass) {
    const int shift = kPermTileBits + pass * sp.digit_bits;
    radix_hist_kernel<<<sp.tiles, kSortThreads, 0, stream>>>(kin, nnz, shift, sp.radix, sp.tiles, hist);
    exclusive_scan_kernel<<<1, 1024, 0, stream>>>(hist, sp.radix * sp.tiles);
    radix_scatter_kernel<<<sp.tiles, kSortThreads, 0, stream>>>(kin, pin, nnz, shift, sp.radix, sp.tiles, hist,
                                                              keys[pass & 1], pay[pass & 1]);
    kin = keys[pass & 1];
    pin = pay[pass & 1];
  }
  perm_view_finish_kernel<<<ceil_div(nnz, 256), 256, 0, stream>>>(nnz, kin, pin, plan->pv_src, plan->pv_loc);
  const cudaError_t le = cudaGetLastError();
  cudaFree(ws);   // waits for the kernels above
  if (le != cudaSuccess) return cuda_fail(le, "tsp_plan_transpose (permutation view)");
  return TSP_OK;
}

int tsp_plan_transpose(tsp_plan* plan, void* stream_, tsp_plan** transposed_out, const int** row_offsets_t,
                       const int** column_indices_t, const int** perm) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (!plan || plan->ephemeral) return invalid("tsp_plan_transpose: needs a plan from tsp_plan_create");
  std::lock_guard<std::mutex> lock(plan->mu);
  if (!plan->transposed) {
    if (stream_is_capturing(stream))
      return invalid("tsp_plan_transpose: the first transpose of a topology allocates; run it once before graph capture");
    const int m = plan->m, n = plan->n, nnz = plan->nnz;
    cudaError_t e = cudaMalloc(reinterpret_cast<void**>(&plan->ro_t), sizeof(int) * ((size_t)n + 1));
    if (e == cudaSuccess) e = cudaMalloc(reinterpret_cast<void**>(&plan->ci_t), sizeof(int) * (size_t)(nnz > 0 ? nnz : 1));
    if (e == cudaSuccess) e = cudaMalloc(reinterpret_cast<void**>(&plan->perm), sizeof(int) * (size_t)(nnz > 0 ? nnz : 1));
    void* ws = nullptr;
    const size_t ws_bytes = tsp_csr_transpose_workspace(m, n, nnz);
    if (e == cudaSuccess) e = cudaMalloc(&ws, ws_bytes + 256);
    if (e != cudaSuccess) {
      cudaFree(plan->ro_t); cudaFree(plan->ci_t); cudaFree(plan->perm); cudaFree(ws);
      plan->ro_t = plan->ci_t = plan->perm = nullptr;
      return cuda_fail(e, "tsp_plan_transpose (buffers)");
    }
    int rc = tsp_csr_transpose_f32(m, n, nnz, 0, nullptr, plan->row_offsets, plan->column_indices, nullptr,
                                   plan->ro_t, plan->ci_t, plan->perm, ws, ws_bytes + 256, stream_);
    if (rc == TSP_OK) {
      tsp_plan* t = nullptr;
      rc = tsp_plan_create(n, m, nnz, plan->ro_t, plan->ci_t, stream_, &t);
      plan->transposed = t;
    }
    if (rc == TSP_OK) rc = build_permutation_view(plan, stream);
    if (rc == TSP_OK) {
      plan->t_stream = stream;
      if (cudaEventCreateWithFlags(&plan->ev_t, cudaEventDisableTiming) != cudaSuccess ||
          cudaEventRecord(plan->ev_t, stream) != cudaSuccess) {
        (void)cudaGetLastError();
        cudaStreamSynchronize(stream);
        plan->t_settled = true;
      }
    }
    cudaFree(ws);   // waits for the sort kernels that use it
    if (rc != TSP_OK) return rc;
  }
  if (!plan->t_settled && stream != plan->t_stream) {
    if (stream_is_capturing(stream)) {
      return invalid("tsp_plan_transpose: the transposed topology was built on another stream and is not known to be "
                     "complete; synchronise once before capturing a CUDA graph");
    } else if (cudaEventQuery(plan->ev_t) == cudaSuccess) {
      plan->t_settled = true;
    } else {
      (void)cudaGetLastError();
      cudaStreamWaitEvent(stream, plan->ev_t, 0);
    }
  }
  if (transposed_out) *transposed_out = plan->transposed;
  if (row_offsets_t) *row_offsets_t = plan->ro_t;
  if (column_indices_t) *column_indices_t = plan->ci_t;
  if (perm) *perm = plan->perm;
  return TSP_OK;
}

}  // extern "C"

// ---------------------------------------------------------------- band kernels (band.cuh)
// Convention of the staged attempts:  0: the kernel wrote the result for certain (columns known
// sorted);  1: launched under the plan's device guard -- the caller queues the guarded fallback
// (guard_flag is set in the arguments);  -1: not launched;  > 1: error code + 1.
template <typename T>
static int try_sddmm_band(tsp_plan* plan, SddmmArgsT<T>& a, int replicas, cudaStream_t stream) {
  StagedGuard guard;
  int rc = 0;
  const int gs = plan_guard(plan, stream, &guard, &rc);
  if (rc != 0) return rc + 1;
  if (gs == kGuardUnsorted) return -1;
  if ((rc = plan_flush_check(plan, stream)) != 0) return rc + 1;
  if ((rc = launch_sddmm_band<T>(a, replicas, stream, guard)) != 0) return rc + 1;
  g_launches.fetch_add(1, std::memory_order_relaxed);
  if (cudaError_t e = cudaGetLastError(); e != cudaSuccess) return cuda_fail(e, "tsp_sddmm (band)") + 1;
  if (gs == kGuardSorted) return 0;
  a.guard_flag = guard.flag;
  return 1;
}

template <typename T>
static int try_spmm_band(tsp_plan* plan, SpmmArgsT<T>& a, int replicas, cudaStream_t stream) {
  StagedGuard guard;
  int rc = 0;
  const int gs = plan_guard(plan, stream, &guard, &rc);
  if (rc != 0) return rc + 1;
  if (gs == kGuardUnsorted) return -1;
  if ((rc = plan_flush_check(plan, stream)) != 0) return rc + 1;
  if ((rc = launch_spmm_band<T>(a, replicas, stream, guard)) != 0) return rc + 1;
  g_launches.fetch_add(1, std::memory_order_relaxed);
  if (cudaError_t e = cudaGetLastError(); e != cudaSuccess) return cuda_fail(e, "tsp_spmm (band)") + 1;
  if (gs == kGuardSorted) return 0;
  a.guard_flag = guard.flag;
  return 1;
}

extern "C" {

// ------------------------------------------------------------------------------------ SpMM
// The gather kernel with the widest vector the operand layout allows.
static void launch_spmm_gather_any(const SpmmArgs& a, int replicas, cudaStream_t stream) {
  const int n = a.n;
  const bool vec4_ok = (n % 4 == 0) && aligned16(a.dense) && aligned16(a.out);
  const bool vec2_ok = (n % 2 == 0) && ((reinterpret_cast<uintptr_t>(a.dense) & 7u) == 0) &&
                       ((reinterpret_cast<uintptr_t>(a.out) & 7u) == 0);
  if (vec4_ok && n > 128) launch_spmm_gather<4, 2>(a, replicas, stream);
  else if (vec4_ok && n > 64) launch_spmm_gather<4, 1>(a, replicas, stream);
  else if (vec2_ok && n > 64) launch_spmm_gather<2, 2>(a, replicas, stream);
  else if (vec2_ok && n > 32) launch_spmm_gather<2, 1>(a, replicas, stream);
  else if (n > 32) launch_spmm_gather<1, 2>(a, replicas, stream);
  else launch_spmm_gather<1, 1>(a, replicas, stream);
}

static int spmm_core(tsp_plan* plan, int n, int replicas, int values_shared, const int* row_indices,
                     const float* values, const int* value_perm_or_null, const float* dense,
                     const float* bias_or_null, int relu, float* out, cudaStream_t stream) {
  const int m = plan->m, k = plan->n, nnz = plan->nnz;
  if (n < 0 || replicas < 0)
    return invalid("tsp_spmm_f32: negative dimension (n=%d replicas=%d)", n, replicas);
  if (replicas > 65535) return invalid("tsp_spmm_f32: replicas=%d exceeds 65535", replicas);
  if (m == 0 || n == 0 || replicas == 0) return TSP_OK;
  if (!plan->row_offsets || !out || (nnz > 0 && (!values || !plan->column_indices || !dense)))
    return invalid("tsp_spmm_f32: null pointer argument");

  SpmmArgs a;
  a.m = m; a.k = k; a.n = n; a.nnz = nnz;
  a.values_stride = values_shared ? 0 : (long long)nnz;
  a.row_indices = row_indices;
  a.values = values;
  a.row_offsets = plan->row_offsets;
  a.column_indices = plan->column_indices;
  a.dense = dense;
  a.bias = bias_or_null;
  a.relu = relu;
  a.out = out;
  a.value_perm = value_perm_or_null;

  const int variant = g_spmm_variant.load();
  // variant 6: the band kernel wherever its shape allows; 0: where it is expected to win; 7: never.
  if ((variant == 6 && band_kc<float>(n) != 0 && aligned16(dense) && aligned16(out)) ||
      (variant == 0 && spmm_band_applicable(a, replicas))) {
    const int st = try_spmm_band<float>(plan, a, replicas, stream);
    if (st > 1) return st - 1;
    if (st == 0) return TSP_OK;
    // st == 1: the gather launch below is the guarded fallback; st == -1: it is the kernel
  }
  const bool vec4_ok = (n % 4 == 0) && aligned16(dense) && aligned16(out);

  // variant: 0 = heuristic, 1 = gather, 2 = staged (rows per group from density), 3 = staged with
  // 4-row groups, 4 = staged with 8-row groups.
  const double density = (double)nnz / ((double)m * (double)k);
  // Below ~22 % density (measured crossover on B200, profiles/) rows of a group share almost no columns, merging cannot raise the FMA
  // per shared-memory wavefront ratio, and the L1/L2 gather kernel has less fixed cost per slab.
  const bool staged_ok = variant != 1 && a.guard_flag == nullptr && spmm_staged_applicable(a, replicas) &&
                         (variant != 0 || density >= 0.22) && variant < 5;
  if (staged_ok) {
    StagedGuard guard;
    int rc = 0;
    const int gs = plan_guard(plan, stream, &guard, &rc);
    if (rc != 0) return rc;
    if (gs != kGuardUnsorted) {
      if ((rc = plan_flush_check(plan, stream)) != 0) return rc;
      rc = launch_spmm_staged(a, replicas, stream, guard, variant == 3 ? 4 : (variant == 4 ? 8 : 0));
      if (rc != 0) return rc;
      if (gs == kGuardSorted) return TSP_OK;
      // Verdict not on the host yet: the same arguments go through the gather kernel, which exits
      // at once unless the check raised the plan's flag.
      a.guard_flag = guard.flag;
      launch_spmm_gather<4, 2>(a, replicas, stream);
      TSP_CHECK_LAUNCH("tsp_spmm_f32 (gather fallback)");
      return TSP_OK;
    }
  }
  // variant 5 (and the heuristic for narrow 16-byte-aligned operands): the row-split kernel, whose
  // G groups of lanes take G nonzeros at once instead of leaving 32 - n/4 lanes idle.
  if ((variant == 5 || (variant == 0 && g_spmm_rowsplit_max_n.load() >= n && vec4_ok)) && spmm_rowsplit_applicable(a)) {
    launch_spmm_rowsplit<float>(a, replicas, stream);
    TSP_CHECK_LAUNCH("tsp_spmm_f32 (rowsplit)");
    return TSP_OK;
  }
  launch_spmm_gather_any(a, replicas, stream);
  TSP_CHECK_LAUNCH("tsp_spmm_f32 (gather)");
  return TSP_OK;
}

int tsp_plan_spmm_f32(tsp_plan* plan, int n, int replicas, int values_shared, const int* row_indices,
                      const float* values, const int* value_perm_or_null, const float* dense,
                      const float* bias_or_null, int relu, float* out, void* stream) {
  if (!plan) return invalid("tsp_plan_spmm_f32: null plan");
  return spmm_core(plan, n, replicas, values_shared, row_indices, values, value_perm_or_null, dense,
                   bias_or_null, relu, out, static_cast<cudaStream_t>(stream));
}

int tsp_spmm_f32(int m, int k, int n, int nnz, int replicas, int values_shared,
                 const int* row_indices, const float* values, const int* row_offsets,
                 const int* column_indices, const float* dense, const float* bias_or_null,
                 int relu, float* out, void* stream_) {
  return tsp_spmm_perm_f32(m, k, n, nnz, replicas, values_shared, row_indices, values, nullptr, row_offsets,
                           column_indices, dense, bias_or_null, relu, out, stream_);
}

int tsp_spmm_perm_f32(int m, int k, int n, int nnz, int replicas, int values_shared,
                      const int* row_indices, const float* values, const int* value_perm_or_null,
                      const int* row_offsets, const int* column_indices, const float* dense,
                      const float* bias_or_null, int relu, float* out, void* stream_) {
  if (m < 0 || k < 0 || n < 0 || nnz < 0 || replicas < 0)
    return invalid("tsp_spmm_f32: negative dimension (m=%d k=%d n=%d nnz=%d replicas=%d)", m, k, n,
                   nnz, replicas);
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  tsp_plan eph;
  ephemeral_plan(&eph, m, k, nnz, row_offsets, column_indices);
  const int rc = spmm_core(&eph, n, replicas, values_shared, row_indices, values, value_perm_or_null, dense,
                           bias_or_null, relu, out, stream);
  plan_release_ephemeral(&eph, stream);
  return rc;
}

// ----------------------------------------------------------------------------------- SDDMM
static void launch_sddmm_gather_any(const SddmmArgs& a, int replicas, cudaStream_t stream) {
  const int k = a.k;
  const bool vec4_ok = (k % 4 == 0) && aligned16(a.lhs) && aligned16(a.rhs);
  if (vec4_ok) {
    if (k <= 32) launch_sddmm_gather<4, 8>(a, replicas, stream);
    else if (k <= 64) launch_sddmm_gather<4, 16>(a, replicas, stream);
    else launch_sddmm_gather<4, 32>(a, replicas, stream);
  } else {
    if (k <= 8) launch_sddmm_gather<1, 8>(a, replicas, stream);
    else if (k <= 16) launch_sddmm_gather<1, 16>(a, replicas, stream);
    else launch_sddmm_gather<1, 32>(a, replicas, stream);
  }
}

// The staged SDDMM kernels (tiled: dense register tiles; sampled: TMA-staged operands, only the
// sampled products).  Returns  0: a staged kernel wrote out[] for certain (columns known sorted);
// 1: a staged kernel was launched under the plan's device guard -- the caller queues the guarded
// gather launch behind it (a.guard_flag is set);  -1: no staged kernel launched;  > 1: error code + 1.
static int sddmm_staged(tsp_plan* plan, SddmmArgs& a, int replicas, cudaStream_t stream, bool want_tiled,
                        bool tiled_v1, bool want_sampled) {
  const bool tiled = want_tiled && sddmm_tiled_applicable(a, replicas);
  const bool sampled = !tiled && want_sampled && sddmm_sampled_applicable(a, replicas);
  if (!tiled && !sampled) return -1;
  StagedGuard guard;
  int rc = 0;
  const int gs = plan_guard(plan, stream, &guard, &rc);
  if (rc != 0) return rc + 1;
  if (gs == kGuardUnsorted) return -1;
  const int bn = sampled ? sddmm_sampled_pick_bn(a, replicas, g_sddmm_bn.load(), g_sddmm_p.load()) : 0;
  const void* view = plan_view(plan, tiled ? kBitmapKey : bn, stream, &rc);
  if (rc != 0) return rc + 1;
  if (!view) return -1;
  if ((rc = plan_flush_check(plan, stream)) != 0) return rc + 1;
  const int t3 = g_sddmm_tiled3.load();
  rc = tiled ? (t3 > 0 && !tiled_v1 ? launch_sddmm_tiled3(a, replicas, stream, guard, view, t3)
                                    : launch_sddmm_tiled(a, replicas, stream, guard, view, tiled_v1))
             : launch_sddmm_sampled(a, replicas, stream, guard, bn, static_cast<const int*>(view), g_sddmm_p.load());
  if (rc != 0) return rc + 1;
  if (gs == kGuardSorted) return 0;
  a.guard_flag = guard.flag;
  return 1;
}

static int sddmm_core(tsp_plan* plan, int k, int replicas, const int* row_indices, const float* lhs,
                      const float* rhs, float* out, cudaStream_t stream, bool reduce) {
  (void)row_indices;  // the flat-nonzero decomposition is balanced without a row swizzle
  const int m = plan->m, n = plan->n, nnz = plan->nnz;
  if (k < 0 || replicas < 0) return invalid("tsp_sddmm_f32: negative dimension (k=%d replicas=%d)", k, replicas);
  if (replicas > 65535) return invalid("tsp_sddmm_f32: replicas=%d exceeds 65535", replicas);
  if (nnz == 0 || replicas == 0) return TSP_OK;
  if (!plan->row_offsets || !plan->column_indices || !out || (k > 0 && (!lhs || !rhs)))
    return invalid("tsp_sddmm_f32: null pointer argument");

  SddmmArgs a;
  a.m = m; a.k = k; a.n = n; a.nnz = nnz;
  a.row_offsets = plan->row_offsets;
  a.column_indices = plan->column_indices;
  a.lhs = lhs;
  a.rhs = rhs;
  a.out = out;
  const double density = (double)nnz / ((double)m * (double)n);
  const int variant = g_sddmm_variant.load();
  if (reduce) {
    // out[j] = sum over replicas: only the staged kernels fold the replica loop into their k loop;
    // anything else computes [replicas, nnz] into stream-ordered scratch and sums it.
    a.reduce_reps = replicas;
    // variant 1: no staged kernel (benchmark knob)
    const int st = variant == 1 ? -1 : sddmm_staged(plan, a, replicas, stream, density >= 0.22, false, density >= 0.015);
    if (st > 1) return st - 1;
    if (st == 0) return TSP_OK;
    if (st == 1) {
      // The staged kernel wrote out[] unless the columns turn out unsorted; in that case the
      // guarded gather kernel (which exits at once otherwise) redoes the work, with the replica
      // loop inside it.
      launch_sddmm_gather_any(a, replicas, stream);
      TSP_CHECK_LAUNCH("tsp_sddmm_reduce_f32 (gather fallback)");
      return TSP_OK;
    }
    // no staged kernel applies: [replicas, nnz] into stream-ordered scratch, then the sum
    float* scratch = nullptr;
    cudaError_t e = cudaMallocAsync(reinterpret_cast<void**>(&scratch), sizeof(float) * (size_t)replicas * nnz, stream);
    if (e != cudaSuccess) return cuda_fail(e, "tsp_sddmm_reduce_f32 (scratch)");
    int rc2 = sddmm_core(plan, k, replicas, row_indices, lhs, rhs, scratch, stream, false);
    if (rc2 == TSP_OK) {
      reduce_replicas_kernel<<<ceil_div(nnz, 256), 256, 0, stream>>>(scratch, nnz, replicas, out);
      g_launches.fetch_add(1, std::memory_order_relaxed);
      if (cudaGetLastError() != cudaSuccess) rc2 = TSP_ERR_CUDA;
    }
    cudaFreeAsync(scratch, stream);
    return rc2;
  }

  // variant: 0 = heuristic, 1 = gather, 2 = tiled (dense register tiles, sampled), 3 = sampled
  // (TMA-staged operands, only the sampled products), 4 = tiled v1,
  // 5 = rows (k <= 128).
  const bool want_tiled = variant == 2 || variant == 4 || (variant == 0 && density >= 0.22);   // 4 = tiled v1
  const bool want_sampled = variant == 3 || (variant == 0 && density >= 0.015 && density < 0.22);
  // 6 = band (short inner dimension, operand blocks staged in shared memory); 0: where it should win
  if ((variant == 6 && band_kc<float>(k) != 0 && aligned16(lhs) && aligned16(rhs)) ||
      (variant == 0 && sddmm_band_applicable(a, replicas))) {
    const int st = try_sddmm_band<float>(plan, a, replicas, stream);
    if (st > 1) return st - 1;
    if (st == 0) return TSP_OK;
  }
  // 5 = rows (short inner dimension, k <= 128): no sortedness requirement, no fallback needed.
  if ((variant == 5 || variant == 6 || (variant == 0 && (k < 128 || density < 0.22))) && sddmm_rows_applicable(a)) {
    launch_sddmm_rows(a, replicas, stream);
    TSP_CHECK_LAUNCH("tsp_sddmm_f32 (rows)");
    return TSP_OK;
  }
  const int st = a.guard_flag != nullptr ? -1 : sddmm_staged(plan, a, replicas, stream, want_tiled, variant == 4, want_sampled);
  if (st > 1) return st - 1;
  if (st == 0) return TSP_OK;
  launch_sddmm_gather_any(a, replicas, stream);   // st == 1: guarded fallback; st == -1: the kernel
  TSP_CHECK_LAUNCH("tsp_sddmm_f32 (gather)");
  return TSP_OK;
}

int tsp_plan_sddmm_f32(tsp_plan* plan, int k, int replicas, int reduce, const int* row_indices,
                       const float* lhs, const float* rhs, float* out, void* stream) {
  if (!plan) return invalid("tsp_plan_sddmm_f32: null plan");
  return sddmm_core(plan, k, replicas, row_indices, lhs, rhs, out, static_cast<cudaStream_t>(stream),
                    reduce != 0 && replicas > 1);
}

static int sddmm_unplanned(int m, int k, int n, int nnz, int replicas, const int* row_indices,
                           const int* row_offsets, const int* column_indices, const float* lhs,
                           const float* rhs, float* out, void* stream_, bool reduce) {
  if (m < 0 || k < 0 || n < 0 || nnz < 0 || replicas < 0)
    return invalid("tsp_sddmm_f32: negative dimension (m=%d k=%d n=%d nnz=%d replicas=%d)", m, k,
                   n, nnz, replicas);
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  tsp_plan eph;
  ephemeral_plan(&eph, m, n, nnz, row_offsets, column_indices);
  const int rc = sddmm_core(&eph, k, replicas, row_indices, lhs, rhs, out, stream, reduce);
  plan_release_ephemeral(&eph, stream);
  return rc;
}

int tsp_sddmm_f32(int m, int k, int n, int nnz, int replicas, const int* row_indices,
                  const int* row_offsets, const int* column_indices, const float* lhs,
                  const float* rhs, float* out, void* stream) {
  return sddmm_unplanned(m, k, n, nnz, replicas, row_indices, row_offsets, column_indices, lhs, rhs, out, stream, false);
}

int tsp_sddmm_reduce_f32(int m, int k, int n, int nnz, int replicas, const int* row_indices,
                         const int* row_offsets, const int* column_indices, const float* lhs,
                         const float* rhs, float* out, void* stream) {
  return sddmm_unplanned(m, k, n, nnz, replicas, row_indices, row_offsets, column_indices, lhs, rhs, out, stream, true);
}

// --------------------------------------------------------------------------------- softmax
static int pick_softmax_reg(int m, int nnz) {
  // Registers per lane so that (nearly) every row fits: mean row length + 4.5 sigma of a binomial
  // row population; longer rows take the three-pass path of the same kernel.
  const double avg = m > 0 ? (double)nnz / m : 0.0;
  const double want = avg + 4.5 * sqrt(avg) + 1.0 + 7.0;   // + up to 7 slots of vector alignment slack
  if (want <= 32 * 4) return 4;
  if (want <= 32 * 8) return 8;
  if (want <= 32 * 16) return 16;
  return 32;
}

}  // extern "C"

template <typename T>
static int softmax_any(const char* name, int m, int nnz, int replicas, float scale, const void* values,
                       const int* row_indices, const int* row_offsets, void* out, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (m < 0 || nnz < 0 || replicas < 0) return invalid("%s: negative dimension", name);
  if (replicas > 65535) return invalid("%s: replicas=%d exceeds 65535", name, replicas);
  if (m == 0 || nnz == 0 || replicas == 0) return TSP_OK;
  if (!values || !row_offsets || !out) return invalid("%s: null pointer argument", name);
  SoftmaxArgsT<T> a;
  a.m = m; a.nnz = nnz; a.scale = scale;
  a.row_indices = row_indices;
  a.row_offsets = row_offsets;
  a.x = static_cast<const T*>(values); a.dy = nullptr; a.out = static_cast<T*>(out);
  constexpr int WARPS = 8;
  dim3 grid(ceil_div(m, WARPS * kSoftmaxRowsPerWarp), replicas);
  int reg = pick_softmax_reg(m, nnz);
  if (sizeof(T) == 2 && reg < 8) reg = 8;   // one 16-byte vector = 8 values per lane
  switch (reg) {
    case 4: csr_softmax_kernel<4, WARPS, T><<<grid, WARPS * 32, 0, stream>>>(a); break;
    case 8: csr_softmax_kernel<8, WARPS, T><<<grid, WARPS * 32, 0, stream>>>(a); break;
    case 16: csr_softmax_kernel<16, WARPS, T><<<grid, WARPS * 32, 0, stream>>>(a); break;
    default: csr_softmax_kernel<32, WARPS, T><<<grid, WARPS * 32, 0, stream>>>(a); break;
  }
  TSP_CHECK_LAUNCH(name);
  return TSP_OK;
}

extern "C" {

int tsp_csr_softmax_f32(int m, int nnz, int replicas, float scale, const float* values,
                        const int* row_indices, const int* row_offsets, float* out,
                        void* stream_) {
  return softmax_any<float>("tsp_csr_softmax_f32", m, nnz, replicas, scale, values, row_indices, row_offsets, out, stream_);
}

int tsp_sparse_attention_supported(int n, int d) {
  return d >= 32 && d <= 128 && d % 32 == 0 &&
         (unsigned long long)n * (unsigned long long)d * 4ull < (1ull << 31);
}

int tsp_sparse_attention_supported_b16(int n, int d) {
  return d >= 64 && d <= 256 && d % 64 == 0 &&
         (unsigned long long)n * (unsigned long long)d * 2ull < (1ull << 31);
}

}  // extern "C"

template <typename T>
static int sparse_attention_any(const char* name, int m, int n, int d, int nnz, int replicas, float scale,
                                const int* row_offsets, const int* column_indices, const void* q,
                                const void* k, const void* v, void* out, float* lse_or_null, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (m < 0 || n < 0 || d < 0 || nnz < 0 || replicas < 0) return invalid("%s: negative dimension", name);
  if (replicas > 65535) return invalid("%s: replicas=%d exceeds 65535", name, replicas);
  if (m == 0 || replicas == 0) return TSP_OK;
  if (!(sizeof(T) == 4 ? tsp_sparse_attention_supported(n, d) : tsp_sparse_attention_supported_b16(n, d)))
    return invalid("%s: head dimension %d (n=%d) not supported; use the three-op chain", name, d, n);
  if (!row_offsets || !q || !k || !v || !out || (nnz > 0 && !column_indices))
    return invalid("%s: null pointer argument", name);
  AttnArgsT<T> a;
  a.m = m; a.n = n; a.d = d; a.nnz = nnz;
  a.row_offsets = row_offsets; a.column_indices = column_indices;
  a.q = static_cast<const T*>(q); a.k = static_cast<const T*>(k); a.v = static_cast<const T*>(v);
  a.out = static_cast<T*>(out); a.lse = lse_or_null; a.scale = scale;
  if (!attention_rows_applicable(a)) return invalid("%s: q, k, v and out must be 16-byte aligned", name);
  launch_attention_rows<T>(a, replicas, stream);
  TSP_CHECK_LAUNCH(name);
  return TSP_OK;
}

extern "C" {

int tsp_sparse_attention_f32(int m, int n, int d, int nnz, int replicas, float scale,
                             const int* row_indices, const int* row_offsets,
                             const int* column_indices, const float* q, const float* k,
                             const float* v, float* out, float* lse_or_null, void* stream) {
  (void)row_indices;   // rows are taken in order: neighbouring rows share k / v rows in L1
  return sparse_attention_any<float>("tsp_sparse_attention_f32", m, n, d, nnz, replicas, scale, row_offsets,
                                     column_indices, q, k, v, out, lse_or_null, stream);
}

int tsp_sparse_attention_b16(int bf16, int m, int n, int d, int nnz, int replicas, float scale,
                             const int* row_indices, const int* row_offsets, const int* column_indices,
                             const void* q, const void* k, const void* v, void* out, float* lse_or_null,
                             void* stream) {
  (void)row_indices;
  return bf16 ? sparse_attention_any<__nv_bfloat16>("tsp_sparse_attention_b16", m, n, d, nnz, replicas, scale,
                                                    row_offsets, column_indices, q, k, v, out, lse_or_null, stream)
              : sparse_attention_any<__half>("tsp_sparse_attention_b16", m, n, d, nnz, replicas, scale, row_offsets,
                                             column_indices, q, k, v, out, lse_or_null, stream);
}

int tsp_csr_softmax_backward_f32(int m, int nnz, int replicas, float scale, const float* y,
                                 const float* dy, const int* row_indices, const int* row_offsets,
                                 float* dx, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (m < 0 || nnz < 0 || replicas < 0)
    return invalid("tsp_csr_softmax_backward_f32: negative dimension");
  if (replicas > 65535)
    return invalid("tsp_csr_softmax_backward_f32: replicas=%d exceeds 65535", replicas);
  if (m == 0 || nnz == 0 || replicas == 0) return TSP_OK;
  if (!y || !dy || !row_offsets || !dx)
    return invalid("tsp_csr_softmax_backward_f32: null pointer argument");
  SoftmaxArgs a;
  a.m = m; a.nnz = nnz; a.scale = scale;
  a.row_indices = row_indices;
  a.row_offsets = row_offsets;
  a.x = y; a.dy = dy; a.out = dx;
  constexpr int WARPS = 8;
  dim3 grid(ceil_div(m, WARPS), replicas);
  switch (pick_softmax_reg(m, nnz)) {
    case 4: csr_softmax_bwd_kernel<4, WARPS><<<grid, WARPS * 32, 0, stream>>>(a); break;
    case 8: csr_softmax_bwd_kernel<8, WARPS><<<grid, WARPS * 32, 0, stream>>>(a); break;
    default: csr_softmax_bwd_kernel<16, WARPS><<<grid, WARPS * 32, 0, stream>>>(a); break;
  }
  TSP_CHECK_LAUNCH("tsp_csr_softmax_backward_f32");
  return TSP_OK;
}

// ------------------------------------------------------------------------------- transpose
static size_t align256(size_t x) { return (x + 255) & ~size_t(255); }

size_t tsp_csr_transpose_workspace(int m, int n, int nnz) {
  (void)m;
  if (nnz < 0 || n < 0) return 0;
  const int variant = g_transpose_variant.load();
  const CountPlan c = make_count_plan(n, nnz, variant);
  if (c.use && variant != 1) return align256(sizeof(int) * count_plan_ints(c, n)) + 256;
  const SortPlan p = make_sort_plan(n, nnz);
  // keys A/B, payload A/B, [radix][tiles] histogram
  return 4 * align256(sizeof(int) * (size_t)nnz) + align256(sizeof(int) * (size_t)p.radix * p.tiles) +
         256;
}

// benchmark knob "transpose" (g_transpose_variant): 1 = radix path always, k >= 2 = at most k * SMs chunks

// Counting path: count -> offsets -> place (kernels/csr_transpose.cuh).
static int transpose_counting(const CountPlan& c, int m, int n, int nnz, int replicas, const float* values,
                              const int* row_offsets, const int* column_indices, float* values_t,
                              int* row_offsets_t, int* column_indices_t, int* perm_or_null, void* workspace,
                              cudaStream_t stream) {
  int* table = static_cast<int*>(workspace);
  int* chain = table + (size_t)c.chunks * n;
  const size_t smem = sizeof(int) * (size_t)n;
  if (smem > 48 * 1024) {
    static std::once_flag once;
    std::call_once(once, [] {
      cudaFuncSetAttribute(transpose_count_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           (int)(sizeof(int) * kCountMaxCols));
      cudaFuncSetAttribute(transpose_place_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           (int)(sizeof(int) * kCountMaxCols));
    });
  }
  transpose_count_kernel<<<c.chunks, kCountThreads, smem, stream>>>(column_indices, n, nnz, c.chunk_len, table, chain,
                                                                    64 + c.groups);
  TSP_CHECK_LAUNCH("tsp_csr_transpose_f32 (count)");
  transpose_offsets_kernel<<<c.groups, kCountGroupCols, 0, stream>>>(n, nnz, c.chunks, table, chain, row_offsets_t);
  TSP_CHECK_LAUNCH("tsp_csr_transpose_f32 (offsets)");
  TransposePlaceArgs a;
  a.m = m; a.n = n; a.nnz = nnz; a.replicas = replicas; a.chunk_len = c.chunk_len;
  a.row_offsets = row_offsets; a.column_indices = column_indices; a.values = values;
  a.table = table; a.row_offsets_t = row_offsets_t;
  a.values_t = values_t; a.column_indices_t = column_indices_t; a.perm = perm_or_null;
  transpose_place_kernel<<<c.chunks, kCountThreads, smem, stream>>>(a);
  TSP_CHECK_LAUNCH("tsp_csr_transpose_f32 (place)");
  return TSP_OK;
}

int tsp_csr_transpose_f32(int m, int n, int nnz, int replicas, const float* values,
                          const int* row_offsets, const int* column_indices, float* values_t,
                          int* row_offsets_t, int* column_indices_t, int* perm_or_null,
                          void* workspace, size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (m < 0 || n < 0 || nnz < 0 || replicas < 0)
    return invalid("tsp_csr_transpose_f32: negative dimension");
  if (!row_offsets_t) return invalid("tsp_csr_transpose_f32: null row_offsets_t");
  if (nnz > 0 && (!row_offsets || !column_indices || !column_indices_t || (replicas > 0 && (!values || !values_t))))
    return invalid("tsp_csr_transpose_f32: null pointer argument");
  if (nnz == 0) {
    fill_int_kernel<<<ceil_div(n + 1, 256), 256, 0, stream>>>(row_offsets_t, n + 1, 0);
    TSP_CHECK_LAUNCH("tsp_csr_transpose_f32 (empty)");
    return TSP_OK;
  }
  if (workspace_bytes < tsp_csr_transpose_workspace(m, n, nnz) || !workspace) {
    set_error("tsp_csr_transpose_f32: workspace of %zu bytes is smaller than the %zu required",
              workspace_bytes, tsp_csr_transpose_workspace(m, n, nnz));
    return TSP_ERR_WORKSPACE;
  }
  if ((reinterpret_cast<uintptr_t>(workspace) & 255u) != 0)
    return invalid("tsp_csr_transpose_f32: workspace must be 256-byte aligned");

  const int variant = g_transpose_variant.load();
  const CountPlan cp = make_count_plan(n, nnz, variant);
  if (cp.use && variant != 1)
    return transpose_counting(cp, m, n, nnz, replicas, values, row_offsets, column_indices, values_t, row_offsets_t,
                              column_indices_t, perm_or_null, workspace, stream);
  const SortPlan p = make_sort_plan(n, nnz);
  char* ws = static_cast<char*>(workspace);
  const size_t seg = align256(sizeof(int) * (size_t)nnz);
  int* keys[2] = {reinterpret_cast<int*>(ws), reinterpret_cast<int*>(ws + seg)};
  int* pay[2] = {reinterpret_cast<int*>(ws + 2 * seg), reinterpret_cast<int*>(ws + 3 * seg)};
  int* hist = reinterpret_cast<int*>(ws + 4 * seg);

  const int* kin = column_indices;
  const int* pin = nullptr;
  for (int pass = 0; pass < p.passes; ++pass) {
    const int shift = pass * p.digit_bits;
    radix_hist_kernel<<<p.tiles, kSortThreads, 0, stream>>>(kin, nnz, shift, p.radix, p.tiles, hist);
    TSP_CHECK_LAUNCH("tsp_csr_transpose_f32 (hist)");
    exclusive_scan_kernel<<<1, 1024, 0, stream>>>(hist, p.radix * p.tiles);
    TSP_CHECK_LAUNCH("tsp_csr_transpose_f32 (scan)");
    radix_scatter_kernel<<<p.tiles, kSortThreads, 0, stream>>>(kin, pin, nnz, shift, p.radix,
                                                              p.tiles, hist, keys[pass & 1],
                                                              pay[pass & 1]);
    TSP_CHECK_LAUNCH("tsp_csr_transpose_f32 (scatter)");
    kin = keys[pass & 1];
    pin = pay[pass & 1];
  }

  TransposeFinishArgs f;
  f.m = m; f.n = n; f.nnz = nnz; f.replicas = replicas;
  f.sorted_cols = kin;
  f.sorted_src = pin;
  f.row_offsets = row_offsets;
  f.values = values;
  f.values_t = values_t;
  f.row_offsets_t = row_offsets_t;
  f.column_indices_t = column_indices_t;
  f.perm = perm_or_null;
  transpose_finish_kernel<<<ceil_div((long long)nnz + 1, 256), 256, 0, stream>>>(f);
  TSP_CHECK_LAUNCH("tsp_csr_transpose_f32 (finish)");
  return TSP_OK;
}

// ---- csr_transpose without any allocation by anyone: the scratch is a file-scope device array ----
// (no counterpart in the reference; added for callers that cannot call
// cudaMalloc: CUDA-graph capture, tight serving loops, harnesses that forbid it).  Every GPU that loads the
// library holds these TSP_NOALLOC_SCRATCH_BYTES until the process exits.  All calls share the array and it keeps
// the last call's contents: the kernels write every word of it they later read (the count table, the chained-scan
// words, the radix histograms and key / payload buffers), so nothing has to start at zero; two calls whose
// kernels overlap on the GPU (different streams, no ordering) race on it.
static __device__ __align__(256) unsigned char g_noalloc_scratch[TSP_NOALLOC_SCRATCH_BYTES];

// Device address of g_noalloc_scratch on the CURRENT device (in host code the array's name is only a host-side
// stand-in): looked up once per device with that device current.
static int noalloc_scratch_address(void** out) {
  static void* cached[64] = {nullptr};
  static std::mutex mu;
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return cuda_fail(e, "tsp_csr_transpose_noalloc_f32 (cudaGetDevice)");
  if (dev < 0 || dev >= 64) return invalid("tsp_csr_transpose_noalloc_f32: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (cached[dev] == nullptr) {
    void* ptr = nullptr;
    e = cudaGetSymbolAddress(&ptr, g_noalloc_scratch);
    if (e != cudaSuccess || ptr == nullptr)
      return cuda_fail(e != cudaSuccess ? e : cudaErrorInvalidSymbol, "tsp_csr_transpose_noalloc_f32 (cudaGetSymbolAddress)");
    cached[dev] = ptr;
  }
  *out = cached[dev];
  return TSP_OK;
}

int tsp_csr_transpose_noalloc_f32(int m, int n, int nnz, int replicas, const float* values,
                                  const int* row_offsets, const int* column_indices, float* values_t,
                                  int* row_offsets_t, int* column_indices_t, int* perm_or_null, void* stream_) {
  if (m < 0 || n < 0 || nnz < 0 || replicas < 0)
    return invalid("tsp_csr_transpose_noalloc_f32: negative dimension");
  const size_t need = tsp_csr_transpose_workspace(m, n, nnz);
  if (need > (size_t)TSP_NOALLOC_SCRATCH_BYTES) {
    set_error("tsp_csr_transpose_noalloc_f32: this problem (m=%d n=%d nnz=%d) needs %zu bytes of scratch, more than "
              "the %zu-byte static array; use tsp_csr_transpose_f32 with a caller-supplied workspace",
              m, n, nnz, need, (size_t)TSP_NOALLOC_SCRATCH_BYTES);
    return TSP_ERR_WORKSPACE;
  }
  void* scratch = nullptr;
  const int rc = noalloc_scratch_address(&scratch);
  if (rc != TSP_OK) return rc;
  return tsp_csr_transpose_f32(m, n, nnz, replicas, values, row_offsets, column_indices, values_t, row_offsets_t,
                               column_indices_t, perm_or_null, scratch, (size_t)TSP_NOALLOC_SCRATCH_BYTES, stream_);
}

int tsp_permute_values_f32(int nnz, int replicas, const float* values, const int* perm,
                           float* values_t, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (nnz < 0 || replicas < 0) return invalid("tsp_permute_values_f32: negative dimension");
  if (replicas > 65535) return invalid("tsp_permute_values_f32: replicas=%d exceeds 65535", replicas);
  if (nnz == 0 || replicas == 0) return TSP_OK;
  if (!values || !perm || !values_t) return invalid("tsp_permute_values_f32: null pointer argument");
  launch_permute_values<float>(nnz, replicas, values, perm, values_t, MaskTable{}, stream);
  TSP_CHECK_LAUNCH("tsp_permute_values_f32");
  return TSP_OK;
}

int tsp_permute_values_b16(int nnz, int replicas, const void* values, const int* perm, void* values_t,
                           void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (nnz < 0 || replicas < 0) return invalid("tsp_permute_values_b16: negative dimension");
  if (replicas > 65535) return invalid("tsp_permute_values_b16: replicas=%d exceeds 65535", replicas);
  if (nnz == 0 || replicas == 0) return TSP_OK;
  if (!values || !perm || !values_t) return invalid("tsp_permute_values_b16: null pointer argument");
  launch_permute_values<unsigned short>(nnz, replicas, static_cast<const unsigned short*>(values), perm,
                                        static_cast<unsigned short*>(values_t), MaskTable{}, stream);
  TSP_CHECK_LAUNCH("tsp_permute_values_b16");
  return TSP_OK;
}

// values_t[r, p] = values[r, perm[p]] with the permutation of the plan's cached transpose: the tiled view
// when the plan has one and there is enough to move, else the plain gather.
// benchmark knob "permute_tiled" (g_permute_tiled): 1 = never, 2 = whenever a view exists
}  // extern "C"

template <typename V>
static int plan_permute_values(const char* name, tsp_plan* plan, int replicas, const V* values, V* values_t,
                               void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (!plan || plan->ephemeral) return invalid("%s: needs a plan from tsp_plan_create", name);
  if (replicas < 0) return invalid("%s: negative replicas", name);
  if (replicas > 65535) return invalid("%s: replicas=%d exceeds 65535", name, replicas);
  const int* perm = nullptr;
  const int rc = tsp_plan_transpose(plan, stream_, nullptr, nullptr, nullptr, &perm);
  if (rc != TSP_OK) return rc;
  const int nnz = plan->nnz;
  if (nnz == 0 || replicas == 0) return TSP_OK;
  if (!values || !values_t) return invalid("%s: null pointer argument", name);
  const int knob = g_permute_tiled.load();
  const bool tiled = plan->pv_src != nullptr && knob != 1 &&
                     (knob == 2 || (long long)nnz * replicas >= (1LL << 21));
  if (tiled) launch_permute_tiled<V>(nnz, replicas, values, plan->pv_src, plan->pv_loc, values_t, stream);
  else launch_permute_values<V>(nnz, replicas, values, perm, values_t, MaskTable{}, stream);
  TSP_CHECK_LAUNCH(name);
  return TSP_OK;
}

extern "C" {

int tsp_plan_permute_values_f32(tsp_plan* plan, int replicas, const float* values, float* values_t, void* stream) {
  return plan_permute_values<float>("tsp_plan_permute_values_f32", plan, replicas, values, values_t, stream);
}

int tsp_plan_permute_values_b16(tsp_plan* plan, int replicas, const void* values, void* values_t, void* stream) {
  return plan_permute_values<unsigned short>("tsp_plan_permute_values_b16", plan, replicas,
                                             static_cast<const unsigned short*>(values),
                                             static_cast<unsigned short*>(values_t), stream);
}

// ----------------------------------------------------------------- many masks in one launch
static int many_check(const char* name, int b, int heads, int m, const int* prefix, const int* row_offsets) {
  if (b < 1 || heads < 1 || m < 0) return invalid("%s: bad mask count / heads / m (b=%d heads=%d m=%d)", name, b, heads, m);
  if ((long long)b * heads > 65535) return invalid("%s: b * heads = %lld exceeds 65535", name, (long long)b * heads);
  if (!prefix || !row_offsets) return invalid("%s: null pointer argument", name);
  return TSP_OK;
}

int tsp_sddmm_many_f32(int b, int heads, int m, int k, int n, int max_nnz, const int* nnz_prefix,
                       const int* row_offsets, const int* column_indices, const float* lhs, const float* rhs,
                       float* out, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = many_check("tsp_sddmm_many_f32", b, heads, m, nnz_prefix, row_offsets)) return rc;
  if (k < 0 || n < 0 || max_nnz < 0) return invalid("tsp_sddmm_many_f32: negative dimension");
  if (max_nnz == 0 || m == 0) return TSP_OK;
  if (!column_indices || !out || (k > 0 && (!lhs || !rhs))) return invalid("tsp_sddmm_many_f32: null pointer argument");
  SddmmArgs a;
  a.m = m; a.k = k; a.n = n; a.nnz = max_nnz;
  a.row_offsets = row_offsets;
  a.column_indices = column_indices;
  a.lhs = lhs; a.rhs = rhs; a.out = out;
  a.masks.prefix = nnz_prefix;
  a.masks.heads = heads;
  if (sddmm_rows_applicable(a)) launch_sddmm_rows(a, b * heads, stream);
  else launch_sddmm_gather_any(a, b * heads, stream);
  TSP_CHECK_LAUNCH("tsp_sddmm_many_f32");
  return TSP_OK;
}

int tsp_spmm_many_f32(int b, int heads, int m, int k, int n, int max_nnz, const int* nnz_prefix,
                      const int* row_indices, const float* values, const int* row_offsets,
                      const int* column_indices, const float* dense, float* out, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = many_check("tsp_spmm_many_f32", b, heads, m, nnz_prefix, row_offsets)) return rc;
  if (k < 0 || n < 0 || max_nnz < 0) return invalid("tsp_spmm_many_f32: negative dimension");
  if (m == 0 || n == 0) return TSP_OK;
  if (!out || (max_nnz > 0 && (!values || !column_indices || !dense))) return invalid("tsp_spmm_many_f32: null pointer argument");
  SpmmArgs a;
  a.m = m; a.k = k; a.n = n; a.nnz = max_nnz;
  a.values_stride = 0;
  a.row_indices = row_indices;
  a.values = values;
  a.row_offsets = row_offsets;
  a.column_indices = column_indices;
  a.dense = dense;
  a.bias = nullptr;
  a.relu = 0;
  a.out = out;
  a.masks.prefix = nnz_prefix;
  a.masks.heads = heads;
  launch_spmm_gather_any(a, b * heads, stream);
  TSP_CHECK_LAUNCH("tsp_spmm_many_f32");
  return TSP_OK;
}

int tsp_csr_softmax_many_f32(int b, int heads, int m, int max_nnz, const int* nnz_prefix, float scale,
                             const float* values, const int* row_offsets, float* out, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = many_check("tsp_csr_softmax_many_f32", b, heads, m, nnz_prefix, row_offsets)) return rc;
  if (max_nnz <= 0 || m == 0) return TSP_OK;
  if (!values || !out) return invalid("tsp_csr_softmax_many_f32: null pointer argument");
  SoftmaxArgs a;
  a.m = m; a.nnz = max_nnz; a.scale = scale;
  a.row_indices = nullptr;
  a.row_offsets = row_offsets;
  a.x = values; a.dy = nullptr; a.out = out;
  a.masks.prefix = nnz_prefix;
  a.masks.heads = heads;
  constexpr int WARPS = 8;
  dim3 grid(ceil_div(m, WARPS * kSoftmaxRowsPerWarp), b * heads);
  switch (pick_softmax_reg(m, max_nnz)) {
    case 4: csr_softmax_kernel<4, WARPS, float><<<grid, WARPS * 32, 0, stream>>>(a); break;
    case 8: csr_softmax_kernel<8, WARPS, float><<<grid, WARPS * 32, 0, stream>>>(a); break;
    case 16: csr_softmax_kernel<16, WARPS, float><<<grid, WARPS * 32, 0, stream>>>(a); break;
    default: csr_softmax_kernel<32, WARPS, float><<<grid, WARPS * 32, 0, stream>>>(a); break;
  }
  TSP_CHECK_LAUNCH("tsp_csr_softmax_many_f32");
  return TSP_OK;
}

int tsp_permute_values_many_f32(int b, int heads, int max_nnz, const int* nnz_prefix, const float* values,
                                const int* perm, float* values_t, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (b < 1 || heads < 1 || (long long)b * heads > 65535) return invalid("tsp_permute_values_many_f32: bad b / heads");
  if (max_nnz <= 0) return TSP_OK;
  if (!nnz_prefix || !values || !perm || !values_t) return invalid("tsp_permute_values_many_f32: null pointer argument");
  MaskTable t;
  t.prefix = nnz_prefix;
  t.heads = heads;
  launch_permute_values<float>(max_nnz, b * heads, values, perm, values_t, t, stream);
  TSP_CHECK_LAUNCH("tsp_permute_values_many_f32");
  return TSP_OK;
}

// ------------------------------------------------------------ 16-bit storage, fp32 arithmetic
// (all gather-type kernels: no topology views, so no plan is involved)
}  // extern "C"

// ---- 16-bit storage through the fp32 staged kernels -------------------------------------------------
// Where the fp32 path would take a kernel that stages dense tiles in shared memory (SpMM at density >=
// 0.22, SDDMM with a long inner dimension), arithmetic -- not operand bytes -- is the bound, and the
// 16-bit gather kernels lose to it (fp16 SDDMM at 4096^3, density 0.5: 9.9 ms gathered against 2.9 ms
// tiled in fp32).  There the 16-bit operands are widened once into stream-ordered fp32 scratch (an
// HBM-bound pass of ~30 us per 64 MB), the fp32 kernel runs, and the result is rounded once on the way
// back: the same contract (fp32 products and sums, one rounding of the result), never slower than the
// fp32 call by more than the two conversion passes.
template <typename T>
__global__ void __launch_bounds__(256) widen_kernel(const T* __restrict__ in, float* __restrict__ out, size_t count) {
  constexpr int E = 8;
  const size_t i = ((size_t)blockIdx.x * 256 + threadIdx.x) * E;
  if (i + E <= count && ((reinterpret_cast<uintptr_t>(in + i) & 15u) == 0)) {
    float v[E];
    unpack_raw<T, E>(v, ld_stream_u4(in + i));
    st_stream_u4(out + i, make_uint4(__float_as_uint(v[0]), __float_as_uint(v[1]), __float_as_uint(v[2]), __float_as_uint(v[3])));
    st_stream_u4(out + i + 4, make_uint4(__float_as_uint(v[4]), __float_as_uint(v[5]), __float_as_uint(v[6]), __float_as_uint(v[7])));
  } else {
    for (size_t j = i; j < count && j < i + E; ++j) out[j] = Elem<T>::to_f(in[j]);
  }
}
template <typename T>
__global__ void __launch_bounds__(256) narrow_kernel(const float* __restrict__ in, T* __restrict__ out, size_t count) {
  constexpr int E = 8;
  const size_t i = ((size_t)blockIdx.x * 256 + threadIdx.x) * E;
  if (i + E <= count && ((reinterpret_cast<uintptr_t>(out + i) & 15u) == 0)) {
    const uint4 a = ld_stream_u4(in + i), b = ld_stream_u4(in + i + 4);
    st_stream_u4(out + i, make_uint4(pack_pair<T>(__uint_as_float(a.x), __uint_as_float(a.y)),
                                     pack_pair<T>(__uint_as_float(a.z), __uint_as_float(a.w)),
                                     pack_pair<T>(__uint_as_float(b.x), __uint_as_float(b.y)),
                                     pack_pair<T>(__uint_as_float(b.z), __uint_as_float(b.w))));
  } else {
    for (size_t j = i; j < count && j < i + E; ++j) out[j] = Elem<T>::from_f(in[j]);
  }
}
template <typename T>
static void widen(const T* in, float* out, size_t count, cudaStream_t stream) {
  widen_kernel<T><<<(unsigned)((count + 2047) / 2048), 256, 0, stream>>>(in, out, count);
  g_launches.fetch_add(1, std::memory_order_relaxed);
}
template <typename T>
static void narrow(const float* in, T* out, size_t count, cudaStream_t stream) {
  narrow_kernel<T><<<(unsigned)((count + 2047) / 2048), 256, 0, stream>>>(in, out, count);
  g_launches.fetch_add(1, std::memory_order_relaxed);
}
constexpr size_t kWidenScratchLimit = size_t(2) << 30;   // fp32 scratch per call the widening path may take

// The widening scratch is hundreds of MB per call: it comes from a pool of this library's own whose
// release threshold keeps the memory across synchronisation points (the device's default pool hands
// everything back at every synchronise, which made each call pay for fresh physical pages: +0.25 ms).
static cudaError_t widen_scratch_alloc(void** p, size_t bytes, cudaStream_t stream) {
  static cudaMemPool_t pools[64] = {nullptr};
  static std::mutex mu;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return cudaMallocAsync(p, bytes, stream);
  {
    std::lock_guard<std::mutex> lock(mu);
    if (pools[dev] == nullptr) {
      cudaMemPoolProps props = {};
      props.allocType = cudaMemAllocationTypePinned;
      props.handleTypes = cudaMemHandleTypeNone;
      props.location.type = cudaMemLocationTypeDevice;
      props.location.id = dev;
      cudaMemPool_t pool = nullptr;
      if (cudaMemPoolCreate(&pool, &props) != cudaSuccess) {
        (void)cudaGetLastError();
        return cudaMallocAsync(p, bytes, stream);
      }
      unsigned long long keep = kWidenScratchLimit;
      cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
      pools[dev] = pool;
    }
  }
  return cudaMallocFromPoolAsync(p, bytes, pools[dev], stream);
}

static int spmm_core(tsp_plan* plan, int n, int replicas, int values_shared, const int* row_indices,
                     const float* values, const int* value_perm_or_null, const float* dense,
                     const float* bias_or_null, int relu, float* out, cudaStream_t stream);
static int sddmm_core(tsp_plan* plan, int k, int replicas, const int* row_indices, const float* lhs,
                      const float* rhs, float* out, cudaStream_t stream, bool reduce);

template <typename T>
static int spmm_b16(const char* name, tsp_plan* plan, int n, int replicas, int values_shared,
                    const int* row_indices, const void* values, const int* value_perm_or_null,
                    const void* dense, const float* bias_or_null, int relu, void* out, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const int m = plan->m, k = plan->n, nnz = plan->nnz;
  const int* row_offsets = plan->row_offsets;
  const int* column_indices = plan->column_indices;
  if (n < 0 || replicas < 0) return invalid("%s: negative dimension (n=%d replicas=%d)", name, n, replicas);
  if (replicas > 65535) return invalid("%s: replicas=%d exceeds 65535", name, replicas);
  if (m == 0 || n == 0 || replicas == 0) return TSP_OK;
  if (!row_offsets || !out || (nnz > 0 && (!values || !column_indices || !dense)))
    return invalid("%s: null pointer argument", name);
  SpmmArgsT<T> a;
  a.m = m; a.k = k; a.n = n; a.nnz = nnz;
  a.values_stride = values_shared ? 0 : (long long)nnz;
  a.row_indices = row_indices;
  a.values = static_cast<const T*>(values);
  a.row_offsets = row_offsets;
  a.column_indices = column_indices;
  a.dense = static_cast<const T*>(dense);
  a.bias = bias_or_null;
  a.relu = relu;
  a.out = static_cast<T*>(out);
  a.value_perm = value_perm_or_null;
  if (((reinterpret_cast<uintptr_t>(dense) | reinterpret_cast<uintptr_t>(out) | reinterpret_cast<uintptr_t>(values)) & 1u) != 0)
    return invalid("%s: operands must be 2-byte aligned", name);
  if (!spmm_rowsplit_applicable(a)) return invalid("%s: dense operand too large for 32-bit row offsets", name);
  const int variant = g_spmm_variant.load();
  if ((variant == 6 && band_kc<T>(n) != 0 && aligned16(dense) && aligned16(out)) ||
      (variant == 0 && spmm_band_applicable(a, replicas))) {
    const int st = try_spmm_band<T>(plan, a, replicas, stream);
    if (st > 1) return st - 1;
    if (st == 0) return TSP_OK;
  }
  {
    // the shapes the fp32 path gives to the staged (merged-row, FFMA2) kernel: widen, run it, narrow
    SpmmArgs f;
    f.m = m; f.k = k; f.n = n; f.nnz = nnz;
    f.dense = reinterpret_cast<const float*>(uintptr_t(16)); f.out = reinterpret_cast<float*>(uintptr_t(16));   // (alignment probe only)
    f.value_perm = value_perm_or_null;
    const size_t nv = (size_t)nnz * (values_shared ? 1 : replicas), nd = (size_t)k * n * replicas, no = (size_t)m * n * replicas;
    // (crossover measured at 4096^3 fp16: row-split 2.10 / 4.15 ms at density 0.25 / 0.5, widened 2.63 / 3.46 ms)
    if (variant == 0 && a.guard_flag == nullptr && (double)nnz / ((double)m * k) >= 0.35 && spmm_staged_applicable(f, replicas) &&
        plan_sorted_state(plan, stream, true) != 0 && (nv + nd + no) * 4 <= kWidenScratchLimit) {
      float* scratch = nullptr;
      cudaError_t e = widen_scratch_alloc(reinterpret_cast<void**>(&scratch), (nv + nd + no) * 4 + 768, stream);
      if (e != cudaSuccess) return cuda_fail(e, name);
      float* v32 = scratch;
      float* d32 = v32 + ((nv + 63) & ~size_t(63));
      float* o32 = d32 + ((nd + 63) & ~size_t(63));
      widen<T>(a.values, v32, nv, stream);
      widen<T>(a.dense, d32, nd, stream);
      int rc = spmm_core(plan, n, replicas, values_shared, row_indices, v32, nullptr, d32, bias_or_null, relu, o32, stream);
      if (rc == TSP_OK) {
        narrow<T>(o32, a.out, no, stream);
        if (cudaGetLastError() != cudaSuccess) rc = TSP_ERR_CUDA;
      }
      cudaFreeAsync(scratch, stream);
      return rc;
    }
  }
  launch_spmm_rowsplit<T>(a, replicas, stream);   // (guarded fallback when a band launch is in front of it)
  TSP_CHECK_LAUNCH(name);
  return TSP_OK;
}

template <typename T>
static int sddmm_b16(const char* name, tsp_plan* plan, int k, int replicas, const void* lhs, const void* rhs,
                     void* out, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const int m = plan->m, n = plan->n, nnz = plan->nnz;
  const int* row_offsets = plan->row_offsets;
  const int* column_indices = plan->column_indices;
  if (k < 0 || replicas < 0) return invalid("%s: negative dimension (k=%d replicas=%d)", name, k, replicas);
  if (replicas > 65535) return invalid("%s: replicas=%d exceeds 65535", name, replicas);
  if (nnz == 0 || replicas == 0) return TSP_OK;
  if (!row_offsets || !column_indices || !out || (k > 0 && (!lhs || !rhs)))
    return invalid("%s: null pointer argument", name);
  SddmmArgsT<T> a;
  a.m = m; a.k = k; a.n = n; a.nnz = nnz;
  a.row_offsets = row_offsets;
  a.column_indices = column_indices;
  a.lhs = static_cast<const T*>(lhs);
  a.rhs = static_cast<const T*>(rhs);
  a.out = static_cast<T*>(out);
  const int variant = g_sddmm_variant.load();
  if ((variant == 6 && band_kc<T>(k) != 0 && aligned16(lhs) && aligned16(rhs)) ||
      (variant == 0 && sddmm_band_applicable(a, replicas))) {
    const int st = try_sddmm_band<T>(plan, a, replicas, stream);
    if (st > 1) return st - 1;
    if (st == 0) return TSP_OK;
  }
  if (variant != 1 && sddmm_rows_applicable(a)) {
    launch_sddmm_rows<T>(a, replicas, stream);
    TSP_CHECK_LAUNCH(name);
    return TSP_OK;
  }
  {
    // Below the density where dense register tiles win: the sampled kernel on the 16-bit operands themselves
    // (TMA-staged 128-byte slabs of 64 elements, packed operands into the mixed-precision FMA).
    SddmmArgs shape;
    const double density = (double)nnz / ((double)m * (double)n);
    const int b16_variant = g_sddmm_b16.load();   // benchmark knob "sddmm_b16": 1 = never, 2 = at any density
    if ((variant == 0 || variant == 3) && b16_variant != 1 && density >= 0.015 &&
        (density < kSddmmB16SampledMaxDensity || b16_variant == 2 || variant == 3) &&
        sddmm_sampled_b16_applicable(a, replicas, &shape)) {
      StagedGuard guard;
      int rc = 0;
      const int gs = plan_guard(plan, stream, &guard, &rc);
      if (rc != 0) return rc;
      if (gs != kGuardUnsorted) {
        const int bn = sddmm_sampled_pick_bn(shape, replicas);
        const void* view = plan_view(plan, bn, stream, &rc);
        if (rc != 0) return rc;
        if (view) {
          if ((rc = plan_flush_check(plan, stream)) != 0) return rc;
          rc = launch_sddmm_sampled<T>(a, replicas, stream, guard, bn, static_cast<const int*>(view));
          if (rc != 0) return rc;
          if (gs == kGuardSorted) return TSP_OK;
          a.guard_flag = guard.flag;   // columns not known sorted: the guarded gather kernel below redoes the work
        }
      }
    }
  }
  if (a.guard_flag == nullptr) {
    // the shapes the fp32 path gives to the TMA-staged kernels (sampled / tiled): widen, run them, narrow
    SddmmArgs f;
    f.m = m; f.k = k; f.n = n; f.nnz = nnz;
    f.lhs = reinterpret_cast<const float*>(uintptr_t(16)); f.rhs = f.lhs;   // (alignment probe only)
    const double density = (double)nnz / ((double)m * (double)n);
    const size_t nl = (size_t)m * k * replicas, nr = (size_t)n * k * replicas, no = (size_t)nnz * replicas;
    const bool staged = (density >= 0.22 && sddmm_tiled_applicable(f, replicas)) ||
                        (density >= 0.015 && sddmm_sampled_applicable(f, replicas));
    // (columns known unsorted: no staged kernel would run on the widened copy either -- straight to the gather)
    if (variant == 0 && a.guard_flag == nullptr && staged && plan_sorted_state(plan, stream, true) != 0 &&
        (nl + nr + no) * 4 <= kWidenScratchLimit) {
      float* scratch = nullptr;
      cudaError_t e = widen_scratch_alloc(reinterpret_cast<void**>(&scratch), (nl + nr + no) * 4 + 768, stream);
      if (e != cudaSuccess) return cuda_fail(e, name);
      float* l32 = scratch;
      float* r32 = l32 + ((nl + 63) & ~size_t(63));
      float* o32 = r32 + ((nr + 63) & ~size_t(63));
      widen<T>(a.lhs, l32, nl, stream);
      widen<T>(a.rhs, r32, nr, stream);
      int rc = sddmm_core(plan, k, replicas, nullptr, l32, r32, o32, stream, false);
      if (rc == TSP_OK) {
        narrow<T>(o32, a.out, no, stream);
        if (cudaGetLastError() != cudaSuccess) rc = TSP_ERR_CUDA;
      }
      cudaFreeAsync(scratch, stream);
      return rc;
    }
  }
  if (k % 8 == 0 && aligned16(lhs) && aligned16(rhs)) {
    if (k <= 64) launch_sddmm_gather<8, 8, T>(a, replicas, stream);
    else if (k <= 128) launch_sddmm_gather<8, 16, T>(a, replicas, stream);
    else launch_sddmm_gather<8, 32, T>(a, replicas, stream);
  } else {
    if (k <= 8) launch_sddmm_gather<1, 8, T>(a, replicas, stream);
    else if (k <= 16) launch_sddmm_gather<1, 16, T>(a, replicas, stream);
    else launch_sddmm_gather<1, 32, T>(a, replicas, stream);
  }
  TSP_CHECK_LAUNCH(name);
  return TSP_OK;
}

extern "C" {

#define TSP_B16_UNPLANNED_SPMM(NAME, TYPE)                                                                     \
  int NAME(int m, int k, int n, int nnz, int replicas, int values_shared, const int* row_indices,              \
           const void* values, const int* value_perm_or_null, const int* row_offsets,                          \
           const int* column_indices, const void* dense, const float* bias_or_null, int relu, void* out,       \
           void* stream) {                                                                                     \
    if (m < 0 || k < 0 || n < 0 || nnz < 0 || replicas < 0)                                                    \
      return invalid(#NAME ": negative dimension (m=%d k=%d n=%d nnz=%d replicas=%d)", m, k, n, nnz, replicas); \
    tsp_plan eph;                                                                                              \
    ephemeral_plan(&eph, m, k, nnz, row_offsets, column_indices);                                              \
    const int rc = spmm_b16<TYPE>(#NAME, &eph, n, replicas, values_shared, row_indices, values,                \
                                  value_perm_or_null, dense, bias_or_null, relu, out, stream);                 \
    plan_release_ephemeral(&eph, static_cast<cudaStream_t>(stream));                                           \
    return rc;                                                                                                 \
  }
TSP_B16_UNPLANNED_SPMM(tsp_spmm_f16, __half)
TSP_B16_UNPLANNED_SPMM(tsp_spmm_bf16, __nv_bfloat16)

#define TSP_B16_UNPLANNED_SDDMM(NAME, TYPE)                                                                    \
  int NAME(int m, int k, int n, int nnz, int replicas, const int* row_indices, const int* row_offsets,         \
           const int* column_indices, const void* lhs, const void* rhs, void* out, void* stream) {             \
    (void)row_indices;                                                                                         \
    if (m < 0 || k < 0 || n < 0 || nnz < 0 || replicas < 0)                                                    \
      return invalid(#NAME ": negative dimension (m=%d k=%d n=%d nnz=%d replicas=%d)", m, k, n, nnz, replicas); \
    tsp_plan eph;                                                                                              \
    ephemeral_plan(&eph, m, n, nnz, row_offsets, column_indices);                                              \
    const int rc = sddmm_b16<TYPE>(#NAME, &eph, k, replicas, lhs, rhs, out, stream);                           \
    plan_release_ephemeral(&eph, static_cast<cudaStream_t>(stream));                                           \
    return rc;                                                                                                 \
  }
TSP_B16_UNPLANNED_SDDMM(tsp_sddmm_f16, __half)
TSP_B16_UNPLANNED_SDDMM(tsp_sddmm_bf16, __nv_bfloat16)

int tsp_plan_spmm_b16(tsp_plan* plan, int bf16, int n, int replicas, int values_shared, const int* row_indices,
                      const void* values, const int* value_perm_or_null, const void* dense,
                      const float* bias_or_null, int relu, void* out, void* stream) {
  if (!plan) return invalid("tsp_plan_spmm_b16: null plan");
  return bf16 ? spmm_b16<__nv_bfloat16>("tsp_plan_spmm_b16", plan, n, replicas, values_shared, row_indices, values,
                                        value_perm_or_null, dense, bias_or_null, relu, out, stream)
              : spmm_b16<__half>("tsp_plan_spmm_b16", plan, n, replicas, values_shared, row_indices, values,
                                 value_perm_or_null, dense, bias_or_null, relu, out, stream);
}

int tsp_plan_sddmm_b16(tsp_plan* plan, int bf16, int k, int replicas, const int* row_indices, const void* lhs,
                       const void* rhs, void* out, void* stream) {
  (void)row_indices;
  if (!plan) return invalid("tsp_plan_sddmm_b16: null plan");
  return bf16 ? sddmm_b16<__nv_bfloat16>("tsp_plan_sddmm_b16", plan, k, replicas, lhs, rhs, out, stream)
              : sddmm_b16<__half>("tsp_plan_sddmm_b16", plan, k, replicas, lhs, rhs, out, stream);
}

int tsp_csr_softmax_f16(int m, int nnz, int replicas, float scale, const void* values, const int* row_indices,
                        const int* row_offsets, void* out, void* stream) {
  return softmax_any<__half>("tsp_csr_softmax_f16", m, nnz, replicas, scale, values, row_indices, row_offsets, out, stream);
}
int tsp_csr_softmax_bf16(int m, int nnz, int replicas, float scale, const void* values, const int* row_indices,
                         const int* row_offsets, void* out, void* stream) {
  return softmax_any<__nv_bfloat16>("tsp_csr_softmax_bf16", m, nnz, replicas, scale, values, row_indices, row_offsets, out, stream);
}

}  // extern "C"
