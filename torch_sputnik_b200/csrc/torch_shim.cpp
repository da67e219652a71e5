// torch_shim.cpp -- the `torch_sputnik` Python module, rebuilt as a thin shim over libtsp.so.
//
// Same module name, op names, positional signatures, return shapes and dtypes as the reference
// binding (src/sputnik.cpp:36-42 of mabdullahsoyturk/Torch-Sputnik), so the reference's
// autograd Functions (modules/spmm.py, sddmm.py, sparse_linear.py, sparse_attention.py) run
// unchanged.  What lives here: shape bookkeeping and the reference's quirks, argument checks
// (TORCH_CHECK -> RuntimeError; the reference uses bare asserts and aborts the process through
// glog on a CUDA error, include/error_check.h:5-17), output allocation with torch::empty (every
// element is written by the kernels, so the reference's torch::zeros pre-clear is dropped), a
// device guard, and the current-stream lookup.  All arithmetic is behind the C-ABI of
// include/tsp.h; there is no CPU or library fallback.
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>
#include <cuda_runtime_api.h>
#include <torch/extension.h>

#include <algorithm>
#include <cstdint>
#include <map>
#include <memory>
#include <mutex>
#include <tuple>
#include <vector>

#include "../../include/tsp.h"

namespace {

#define TSP_CHECK_CUDA_F32(t, name)                                                          \
  TORCH_CHECK((t).is_cuda(), "torch_sputnik: ", name, " must be a CUDA tensor");             \
  TORCH_CHECK((t).scalar_type() == torch::kFloat32, "torch_sputnik: ", name,                 \
              " must be float32, got ", (t).scalar_type());                                  \
  TORCH_CHECK((t).is_contiguous(), "torch_sputnik: ", name, " must be contiguous")

#define TSP_CHECK_CUDA_I32(t, name)                                                          \
  TORCH_CHECK((t).is_cuda(), "torch_sputnik: ", name, " must be a CUDA tensor");             \
  TORCH_CHECK((t).scalar_type() == torch::kInt32, "torch_sputnik: ", name,                   \
              " must be int32, got ", (t).scalar_type());                                    \
  TORCH_CHECK((t).is_contiguous(), "torch_sputnik: ", name, " must be contiguous")

void check_rc(int rc, const char* op) {
  TORCH_CHECK(rc == TSP_OK, "torch_sputnik.", op, " failed (", rc, "): ", tsp_last_error());
}

// A float32 CUDA operand, made contiguous.  The reference assumes contiguity and reads garbage
// otherwise: its own callers hand the ops transposed views in backward (the gradient of
// `.transpose(1,2).contiguous()` in modules/sparse_attention.py:108 reaches
// SparseLinearFunction.backward, modules/sparse_linear.py:44,60, as a non-contiguous view), so the
// shim copies instead of failing.
torch::Tensor f32_operand(const torch::Tensor& t, const char* name) {
  TORCH_CHECK(t.is_cuda(), "torch_sputnik: ", name, " must be a CUDA tensor");
  TORCH_CHECK(t.scalar_type() == torch::kFloat32, "torch_sputnik: ", name, " must be float32, got ",
              t.scalar_type());
  return t.contiguous();
}

// modules/sparse_linear.py:5-7,57 hands an int64 row_indices to left_spmm in backward; the
// reference's data_ptr<int>() would throw.  Accept it and narrow.
torch::Tensor as_i32_indices(const torch::Tensor& t, const char* name) {
  TORCH_CHECK(t.is_cuda(), "torch_sputnik: ", name, " must be a CUDA tensor");
  if (t.scalar_type() == torch::kInt64) return t.to(torch::kInt32).contiguous();
  TORCH_CHECK(t.scalar_type() == torch::kInt32, "torch_sputnik: ", name,
              " must be int32 (or int64), got ", t.scalar_type());
  return t.contiguous();
}

// fp32, or 16-bit storage (fp16 / bf16: the a13 entry points of include/tsp.h), made contiguous.
torch::Tensor value_operand(const torch::Tensor& t, const char* name) {
  TORCH_CHECK(t.is_cuda(), "torch_sputnik: ", name, " must be a CUDA tensor");
  const auto st = t.scalar_type();
  TORCH_CHECK(st == torch::kFloat32 || st == torch::kFloat16 || st == torch::kBFloat16, "torch_sputnik: ", name,
              " must be float32, float16 or bfloat16, got ", st);
  return t.contiguous();
}

void* current_stream() { return static_cast<void*>(at::cuda::getCurrentCUDAStream().stream()); }

bool current_stream_capturing() {
  cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(at::cuda::getCurrentCUDAStream().stream(), &st) != cudaSuccess) {
    (void)cudaGetLastError();
    return false;
  }
  return st != cudaStreamCaptureStatusNone;
}

void check_topology(int64_t m, const torch::Tensor& row_indices, const torch::Tensor& row_offsets,
                    const torch::Tensor& column_indices) {
  TORCH_CHECK(row_indices.dim() == 1, "torch_sputnik: row_indices must have 1 dimension");
  TORCH_CHECK(row_offsets.dim() == 1, "torch_sputnik: row_offsets must have 1 dimension");
  TORCH_CHECK(column_indices.dim() == 1, "torch_sputnik: column_indices must have 1 dimension");
  TORCH_CHECK(row_indices.size(0) == m, "torch_sputnik: expected ", m, " row_indices, got ",
              row_indices.size(0));
  TORCH_CHECK(row_offsets.size(0) == m + 1, "torch_sputnik: expected ", m + 1,
              " row_offsets, got ", row_offsets.size(0));
  TSP_CHECK_CUDA_I32(row_offsets, "row_offsets");
  TSP_CHECK_CUDA_I32(column_indices, "column_indices");
}

// ------------------------------------------------------------------------------- plan cache
// One tsp_plan per topology the process has seen ("optional transpose-plan cache
// ... mutex-guarded and keyed on (data_ptr, numel, _version)").  The reference's callers keep their
// topology tensors for the life of the module (modules/sparse_attention.py:49-52,
// modules/sparse_linear.py:78-83) and pass them to every call, so the plan is found by address.
// An entry HOLDS the two index tensors: their storage cannot be freed -- and its address handed
// to a different topology by the caching allocator -- while the entry lives, and an in-place edit
// bumps _version, which is compared on every hit.  LRU-bounded.
struct PlanHolder {
  tsp_plan* plan = nullptr;
  ~PlanHolder() { tsp_plan_destroy(plan); }
};

struct PlanEntry {
  std::shared_ptr<PlanHolder> holder;   // owner of `plan` (the PARENT's holder for a transposed child)
  tsp_plan* plan = nullptr;
  torch::Tensor row_offsets, column_indices;   // keep the storage (and so the address) alive
  uint32_t ro_version = 0, ci_version = 0;
  // transposed topology, filled on demand: views of the plan's own device arrays
  torch::Tensor ro_t, ci_t, perm, ri_t;
  int64_t nnz = 0;
  bool verdict_known = false;
  uint64_t tick = 0;
};

using PlanKey = std::tuple<const void*, const void*, int64_t, int64_t, int64_t>;

struct PlanCache {
  std::mutex mu;
  std::map<PlanKey, std::shared_ptr<PlanEntry>> entries;
  // perm array of a planned transpose -> owner of the plan it belongs to (the array is the plan's own device
  // memory, so while the owner lives the address cannot mean anything else)
  std::map<const void*, std::weak_ptr<PlanHolder>> by_perm;
  uint64_t tick = 0, hits = 0, misses = 0, evictions = 0;
  int policy = 1;        // 0: never plan (every call takes the unplanned C-ABI entry), 1: plan on first sight
  size_t capacity = 256;
};

PlanCache& plan_cache() {
  static PlanCache* c = new PlanCache();   // never destroyed: no CUDA calls during interpreter teardown
  return *c;
}

void evict_locked(PlanCache& c) {
  while (c.entries.size() > c.capacity) {
    auto victim = c.entries.begin();
    for (auto it = c.entries.begin(); it != c.entries.end(); ++it)
      if (it->second->tick < victim->second->tick) victim = it;
    c.entries.erase(victim);
    ++c.evictions;
  }
}

// The plan of the m x n topology at (ro_ptr, ci_ptr), slices of the two tensors; nullptr when
// planning is off or a new plan cannot be made right now (stream under CUDA-graph capture).
std::shared_ptr<PlanEntry> get_plan(int64_t m, int64_t n, int64_t nnz, const torch::Tensor& row_offsets,
                                    const int* ro_ptr, const torch::Tensor& column_indices, const int* ci_ptr) {
  PlanCache& c = plan_cache();
  const PlanKey key{ro_ptr, ci_ptr, m, n, nnz};
  std::lock_guard<std::mutex> lock(c.mu);
  if (c.policy == 0) return nullptr;
  auto it = c.entries.find(key);
  if (it != c.entries.end()) {
    PlanEntry& e = *it->second;
    if (e.ro_version == row_offsets._version() && e.ci_version == column_indices._version() &&
        e.row_offsets.data_ptr() == row_offsets.data_ptr() &&
        e.column_indices.data_ptr() == column_indices.data_ptr()) {
      e.tick = ++c.tick;
      ++c.hits;
      // Under CUDA-graph capture only a plan whose verdict already reached the host is used (finding it out
      // would take an event query, which invalidates a capture); other calls take the unplanned entries.
      if (!e.verdict_known) {
        if (current_stream_capturing()) return nullptr;
        e.verdict_known = tsp_plan_sorted(e.plan) >= 0;
      }
      return it->second;
    }
    c.entries.erase(it);   // edited in place (or another view of the storage): plan again
  }
  ++c.misses;
  if (current_stream_capturing()) return nullptr;
  auto e = std::make_shared<PlanEntry>();
  e->holder = std::make_shared<PlanHolder>();
  check_rc(tsp_plan_create((int)m, (int)n, (int)nnz, ro_ptr, ci_ptr, current_stream(), &e->holder->plan),
           "plan_create");
  e->plan = e->holder->plan;
  e->row_offsets = row_offsets;
  e->column_indices = column_indices;
  e->ro_version = row_offsets._version();
  e->ci_version = column_indices._version();
  e->nnz = nnz;
  e->tick = ++c.tick;
  c.entries[key] = e;
  evict_locked(c);
  return e;
}

std::shared_ptr<PlanEntry> get_plan(int64_t m, int64_t n, const torch::Tensor& row_offsets,
                                    const torch::Tensor& column_indices) {
  return get_plan(m, n, column_indices.size(0), row_offsets, row_offsets.data_ptr<int>(), column_indices,
                  column_indices.data_ptr<int>());
}

// Row swizzle of a topology: heaviest row first (tests/sparse_matrix.py:22), int32.
torch::Tensor heaviest_first(const torch::Tensor& row_offsets) {
  const int64_t rows = row_offsets.size(0) - 1;
  torch::Tensor lens = row_offsets.slice(0, 1, rows + 1) - row_offsets.slice(0, 0, rows);
  return torch::argsort(lens, /*stable=*/true, /*dim=*/0, /*descending=*/true).to(torch::kInt32);
}

// Fills e.ro_t / e.ci_t / e.perm / e.ri_t (views of the plan's device arrays, kept alive by the
// holder) and registers the child plan under the transposed arrays' addresses, so an
// `spmm(k, m, values_t, row_indices_t, row_offsets_t, column_indices_t, ...)` that follows
// (modules/spmm.py:67-73) finds it.
void ensure_transposed(const std::shared_ptr<PlanEntry>& e, int64_t m, int64_t n, const torch::Device& device) {
  PlanCache& c = plan_cache();
  {
    std::lock_guard<std::mutex> lock(c.mu);
    if (e->ro_t.defined()) {
      TORCH_CHECK(e->ro_t._version() == 0 && e->ci_t._version() == 0 && e->perm._version() == 0,
                  "torch_sputnik.csr_transpose: the transposed index tensors of this topology are shared between "
                  "calls and were modified in place; clone() them before editing");
      return;
    }
  }
  tsp_plan* child = nullptr;
  const int *ro_t = nullptr, *ci_t = nullptr, *perm = nullptr;
  check_rc(tsp_plan_transpose(e->plan, current_stream(), &child, &ro_t, &ci_t, &perm), "plan_transpose");
  auto iopt = torch::TensorOptions().dtype(torch::kInt32).device(device);
  std::shared_ptr<PlanHolder> keep = e->holder;
  auto deleter = [keep](void*) {};
  const int64_t count = e->nnz;   // (a many-mask slice has fewer entries than the tensor it lives in)
  torch::Tensor ro_t_t = torch::from_blob(const_cast<int*>(ro_t), {n + 1}, deleter, iopt);
  torch::Tensor ci_t_t = torch::from_blob(const_cast<int*>(ci_t), {count}, deleter, iopt);
  torch::Tensor perm_t = torch::from_blob(const_cast<int*>(perm), {count}, deleter, iopt);
  torch::Tensor ri_t = heaviest_first(ro_t_t);
  std::lock_guard<std::mutex> lock(c.mu);
  if (e->ro_t.defined()) return;
  e->ro_t = ro_t_t;
  e->ci_t = ci_t_t;
  e->perm = perm_t;
  e->ri_t = ri_t;
  auto ce = std::make_shared<PlanEntry>();
  ce->holder = e->holder;
  ce->plan = child;
  ce->row_offsets = ro_t_t;
  ce->column_indices = ci_t_t;
  ce->nnz = count;
  ce->tick = ++c.tick;
  c.entries[PlanKey{ro_t, ci_t, n, m, count}] = ce;
  for (auto it = c.by_perm.begin(); it != c.by_perm.end();)
    it = it->second.expired() ? c.by_perm.erase(it) : std::next(it);
  c.by_perm[perm] = e->holder;
  evict_locked(c);
}

// Shared body of spmm / left_spmm / spmm_bias_relu.
torch::Tensor spmm_impl(int64_t m, int64_t k, const torch::Tensor& values_in,
                        const torch::Tensor& row_indices_in, const torch::Tensor& row_offsets,
                        const torch::Tensor& column_indices, const torch::Tensor& dense_in,
                        const c10::optional<torch::Tensor>& bias, bool relu, bool left,
                        const char* op, const c10::optional<torch::Tensor>& value_perm = c10::nullopt) {
  TORCH_CHECK(values_in.dim() == 1 || values_in.dim() == 2, "torch_sputnik.", op,
              ": values must have 1 or 2 dimensions");
  TORCH_CHECK(dense_in.dim() == 2 || dense_in.dim() == 3, "torch_sputnik.", op,
              ": dense must have 2 or 3 dimensions");
  torch::Tensor values = value_operand(values_in, "values");
  torch::Tensor dense = value_operand(dense_in, "dense");
  TORCH_CHECK(values.scalar_type() == dense.scalar_type(), "torch_sputnik.", op, ": values (", values.scalar_type(),
              ") and dense (", dense.scalar_type(), ") must have the same dtype");
  const auto dtype = values.scalar_type();
  c10::cuda::CUDAGuard guard(values.device());
  torch::Tensor row_indices = as_i32_indices(row_indices_in, "row_indices");
  check_topology(m, row_indices, row_offsets, column_indices);

  const int64_t nnz = column_indices.size(0);
  const int dim_offset = dense.dim() - 2;
  const int64_t replicas = dim_offset == 1 ? dense.size(0) : 1;
  const int64_t n = dense.size(dim_offset + 1);
  TORCH_CHECK(dense.size(dim_offset) == k, "torch_sputnik.", op,
              ": inner dimensions must match (dense has ", dense.size(dim_offset), " rows, k=", k, ")");
  TORCH_CHECK(values.size(-1) == nnz, "torch_sputnik.", op,
              ": number of values must equal number of column indices");
  bool shared = left || values.dim() == 1;
  if (!shared)
    TORCH_CHECK(values.size(0) == replicas, "torch_sputnik.", op,
                ": first dims of values and dense must match");
  if (left) TORCH_CHECK(values.dim() == 1, "torch_sputnik.left_spmm: values must be 1-D");

  const float* bias_ptr = nullptr;
  torch::Tensor bias_t;
  if (bias.has_value()) {
    bias_t = f32_operand(*bias, "bias");
    TORCH_CHECK(bias_t.dim() == 1 && bias_t.size(0) == m, "torch_sputnik.", op,
                ": bias must have m elements (it is indexed by output row)");
    bias_ptr = bias_t.data_ptr<float>();
  }

  const int* perm_ptr = nullptr;
  torch::Tensor perm_t;
  if (value_perm.has_value()) {
    perm_t = *value_perm;
    TSP_CHECK_CUDA_I32(perm_t, "value_perm");
    TORCH_CHECK(perm_t.dim() == 1 && perm_t.size(0) == nnz, "torch_sputnik.", op, ": value_perm must have nnz elements");
    perm_ptr = perm_t.data_ptr<int>();
  }

  auto options = torch::TensorOptions().dtype(dtype).device(values.device());
  // Reference quirk: `spmm` returns 2-D whenever the replica count is 1 (src/spmm_cuda.cu:46);
  // `left_spmm` is always 3-D (src/left_replicated_spmm.cu:30).
  torch::Tensor out = (left || replicas != 1) ? torch::empty({replicas, m, n}, options)
                                              : torch::empty({m, n}, options);
  if (dtype != torch::kFloat32) {
    // 16-bit storage, fp32 arithmetic, result in the storage type (cuda_spmm.h:75-127)
    if (auto e = get_plan(m, k, row_offsets, column_indices)) {
      check_rc(tsp_plan_spmm_b16(e->plan, dtype == torch::kBFloat16 ? 1 : 0, (int)n, (int)replicas, shared ? 1 : 0,
                                 row_indices.data_ptr<int>(), values.data_ptr(), perm_ptr, dense.data_ptr(), bias_ptr,
                                 relu ? 1 : 0, out.data_ptr(), current_stream()),
               op);
      return out;
    }
    auto fn = dtype == torch::kFloat16 ? tsp_spmm_f16 : tsp_spmm_bf16;
    check_rc(fn((int)m, (int)k, (int)n, (int)nnz, (int)replicas, shared ? 1 : 0, row_indices.data_ptr<int>(),
                values.data_ptr(), perm_ptr, row_offsets.data_ptr<int>(), column_indices.data_ptr<int>(),
                dense.data_ptr(), bias_ptr, relu ? 1 : 0, out.data_ptr(), current_stream()),
             op);
    return out;
  }
  if (auto e = get_plan(m, k, row_offsets, column_indices)) {
    check_rc(tsp_plan_spmm_f32(e->plan, (int)n, (int)replicas, shared ? 1 : 0, row_indices.data_ptr<int>(),
                               values.data_ptr<float>(), perm_ptr, dense.data_ptr<float>(), bias_ptr,
                               relu ? 1 : 0, out.data_ptr<float>(), current_stream()),
             op);
    return out;
  }
  check_rc(tsp_spmm_perm_f32((int)m, (int)k, (int)n, (int)nnz, (int)replicas, shared ? 1 : 0,
                             row_indices.data_ptr<int>(), values.data_ptr<float>(), perm_ptr,
                             row_offsets.data_ptr<int>(), column_indices.data_ptr<int>(),
                             dense.data_ptr<float>(), bias_ptr, relu ? 1 : 0, out.data_ptr<float>(),
                             current_stream()),
           op);
  return out;
}

}  // namespace

// src/sputnik.cpp:4-9 -> src/spmm_cuda.cu:9-60
torch::Tensor spmm(int m, int k, torch::Tensor values, torch::Tensor row_indices,
                   torch::Tensor row_offsets, torch::Tensor column_indices, torch::Tensor dense) {
  return spmm_impl(m, k, values, row_indices, row_offsets, column_indices, dense, c10::nullopt,
                   false, false, "spmm");
}

// src/sputnik.cpp:11-16 -> src/left_replicated_spmm.cu:8-44
torch::Tensor left_spmm(int m, int k, torch::Tensor values, torch::Tensor row_indices,
                        torch::Tensor row_offsets, torch::Tensor column_indices,
                        torch::Tensor dense) {
  return spmm_impl(m, k, values, row_indices, row_offsets, column_indices, dense, c10::nullopt,
                   false, true, "left_spmm");
}

// The op tests/test_spmm_bias_relu.py:37 calls (`spmm_bias`) but the reference never bound:
// sputnik::CudaSpmmBiasRelu, out = max(A @ dense + bias[:, None], 0).
torch::Tensor spmm_bias_relu(int m, int k, torch::Tensor values, torch::Tensor row_indices,
                             torch::Tensor row_offsets, torch::Tensor column_indices,
                             torch::Tensor bias, torch::Tensor dense) {
  return spmm_impl(m, k, values, row_indices, row_offsets, column_indices, dense, bias, true,
                   false, "spmm_bias_relu");
}

static torch::Tensor sddmm_impl(int m, int n, const torch::Tensor& row_indices_in,
                                const torch::Tensor& row_offsets, const torch::Tensor& column_indices,
                                const torch::Tensor& lhs_in, const torch::Tensor& rhs_in, bool reduce) {
  TORCH_CHECK(lhs_in.dim() == 2 || lhs_in.dim() == 3, "torch_sputnik.sddmm: lhs must be 2-D or 3-D");
  TORCH_CHECK(lhs_in.dim() == rhs_in.dim(), "torch_sputnik.sddmm: lhs and rhs must have the same rank");
  torch::Tensor lhs = value_operand(lhs_in, "lhs_matrix");
  torch::Tensor rhs = value_operand(rhs_in, "rhs_matrix");
  TORCH_CHECK(lhs.scalar_type() == rhs.scalar_type(), "torch_sputnik.sddmm: lhs (", lhs.scalar_type(), ") and rhs (",
              rhs.scalar_type(), ") must have the same dtype");
  const auto dtype = lhs.scalar_type();
  TORCH_CHECK(!reduce || dtype == torch::kFloat32, "torch_sputnik.sddmm_reduce: float32 only");
  c10::cuda::CUDAGuard guard(lhs.device());
  torch::Tensor row_indices = as_i32_indices(row_indices_in, "row_indices");
  check_topology(m, row_indices, row_offsets, column_indices);
  TORCH_CHECK(lhs.size(-1) == rhs.size(-1), "torch_sputnik.sddmm: last dims of lhs and rhs must match");

  const int64_t nnz = column_indices.size(0);
  const int dim_offset = lhs.dim() - 2;
  const int64_t replicas = dim_offset == 1 ? lhs.size(0) : 1;
  const int64_t k = lhs.size(dim_offset + 1);
  TORCH_CHECK(lhs.size(dim_offset) == m, "torch_sputnik.sddmm: lhs must have m rows");
  TORCH_CHECK(rhs.size(dim_offset) == n, "torch_sputnik.sddmm: rhs must have n rows");
  TORCH_CHECK(dim_offset == 0 || rhs.size(0) == replicas,
              "torch_sputnik.sddmm: first dims of lhs and rhs must match");

  auto options = torch::TensorOptions().dtype(dtype).device(lhs.device());
  torch::Tensor out = (replicas == 1 || reduce) ? torch::empty({nnz}, options)
                                                : torch::empty({replicas, nnz}, options);
  const char* op = reduce ? "sddmm_reduce" : "sddmm";
  if (dtype != torch::kFloat32) {
    if (auto e = get_plan(m, n, row_offsets, column_indices)) {
      check_rc(tsp_plan_sddmm_b16(e->plan, dtype == torch::kBFloat16 ? 1 : 0, (int)k, (int)replicas,
                                  row_indices.data_ptr<int>(), lhs.data_ptr(), rhs.data_ptr(), out.data_ptr(),
                                  current_stream()),
               op);
      return out;
    }
    auto fn16 = dtype == torch::kFloat16 ? tsp_sddmm_f16 : tsp_sddmm_bf16;
    check_rc(fn16(m, (int)k, n, (int)nnz, (int)replicas, row_indices.data_ptr<int>(), row_offsets.data_ptr<int>(),
                  column_indices.data_ptr<int>(), lhs.data_ptr(), rhs.data_ptr(), out.data_ptr(), current_stream()),
             op);
    return out;
  }
  if (auto e = get_plan(m, n, row_offsets, column_indices)) {
    check_rc(tsp_plan_sddmm_f32(e->plan, (int)k, (int)replicas, reduce ? 1 : 0, row_indices.data_ptr<int>(),
                                lhs.data_ptr<float>(), rhs.data_ptr<float>(), out.data_ptr<float>(),
                                current_stream()),
             op);
    return out;
  }
  auto fn = (reduce && replicas > 1) ? tsp_sddmm_reduce_f32 : tsp_sddmm_f32;
  check_rc(fn(m, (int)k, n, (int)nnz, (int)replicas, row_indices.data_ptr<int>(),
              row_offsets.data_ptr<int>(), column_indices.data_ptr<int>(), lhs.data_ptr<float>(),
              rhs.data_ptr<float>(), out.data_ptr<float>(), current_stream()),
           op);
  return out;
}

// src/sputnik.cpp:18-23 -> src/sddmm_cuda.cu:7-57
torch::Tensor sddmm(int m, int n, torch::Tensor row_indices, torch::Tensor row_offsets,
                    torch::Tensor column_indices, torch::Tensor lhs, torch::Tensor rhs) {
  return sddmm_impl(m, n, row_indices, row_offsets, column_indices, lhs, rhs, false);
}

// sddmm summed over the replica dimension -> [nnz]: what modules/sparse_linear.py:44 + autograd's
// broadcast-sum compute for the weight gradient, without the [R, nnz] intermediate.
torch::Tensor sddmm_reduce(int m, int n, torch::Tensor row_indices, torch::Tensor row_offsets,
                           torch::Tensor column_indices, torch::Tensor lhs, torch::Tensor rhs) {
  return sddmm_impl(m, n, row_indices, row_offsets, column_indices, lhs, rhs, true);
}

static torch::Tensor softmax_impl(const torch::Tensor& values_in, const torch::Tensor& row_indices_in,
                                  const torch::Tensor& row_offsets,
                                  const torch::Tensor& column_indices, double scale) {
  TORCH_CHECK(values_in.dim() == 1 || values_in.dim() == 2,
              "torch_sputnik.sparse_softmax: values must have 1 or 2 dimensions");
  torch::Tensor values = value_operand(values_in, "values");
  c10::cuda::CUDAGuard guard(values.device());
  torch::Tensor row_indices = as_i32_indices(row_indices_in, "row_indices");
  const int64_t m = row_indices.size(0);
  check_topology(m, row_indices, row_offsets, column_indices);
  const int64_t nnz = column_indices.size(0);
  TORCH_CHECK(values.size(-1) == nnz,
              "torch_sputnik.sparse_softmax: number of values must equal number of column indices");
  const int64_t replicas = values.dim() == 2 ? values.size(0) : 1;
  torch::Tensor out = torch::empty_like(values);
  if (values.scalar_type() != torch::kFloat32) {
    auto fn16 = values.scalar_type() == torch::kFloat16 ? tsp_csr_softmax_f16 : tsp_csr_softmax_bf16;
    check_rc(fn16((int)m, (int)nnz, (int)replicas, (float)scale, values.data_ptr(), row_indices.data_ptr<int>(),
                  row_offsets.data_ptr<int>(), out.data_ptr(), current_stream()),
             "sparse_softmax");
    return out;
  }
  check_rc(tsp_csr_softmax_f32((int)m, (int)nnz, (int)replicas, (float)scale,
                               values.data_ptr<float>(), row_indices.data_ptr<int>(),
                               row_offsets.data_ptr<int>(), out.data_ptr<float>(), current_stream()),
           "sparse_softmax");
  return out;
}

// src/sputnik.cpp:25-28 -> src/softmax_cuda.cu:7-46
torch::Tensor sparse_softmax(torch::Tensor values, torch::Tensor row_indices,
                             torch::Tensor row_offsets, torch::Tensor column_indices) {
  return softmax_impl(values, row_indices, row_offsets, column_indices, 1.0);
}

// Extension: softmax(scale * x) in one pass (folds modules/sparse_attention.py:72).
torch::Tensor sparse_softmax_scaled(torch::Tensor values, torch::Tensor row_indices,
                                    torch::Tensor row_offsets, torch::Tensor column_indices,
                                    double scale) {
  return softmax_impl(values, row_indices, row_offsets, column_indices, scale);
}

// Extension: dx = scale * y * (dy - rowsum(dy * y)).
torch::Tensor sparse_softmax_backward(torch::Tensor y, torch::Tensor dy, torch::Tensor row_indices_in,
                                      torch::Tensor row_offsets, torch::Tensor column_indices,
                                      double scale) {
  TSP_CHECK_CUDA_F32(y, "y");
  TSP_CHECK_CUDA_F32(dy, "dy");
  TORCH_CHECK(y.sizes() == dy.sizes(), "torch_sputnik.sparse_softmax_backward: y and dy must have the same shape");
  c10::cuda::CUDAGuard guard(y.device());
  torch::Tensor row_indices = as_i32_indices(row_indices_in, "row_indices");
  const int64_t m = row_indices.size(0);
  check_topology(m, row_indices, row_offsets, column_indices);
  const int64_t nnz = column_indices.size(0);
  TORCH_CHECK(y.size(-1) == nnz, "torch_sputnik.sparse_softmax_backward: bad number of values");
  const int64_t replicas = y.dim() == 2 ? y.size(0) : 1;
  torch::Tensor dx = torch::empty_like(y);
  check_rc(tsp_csr_softmax_backward_f32((int)m, (int)nnz, (int)replicas, (float)scale,
                                        y.data_ptr<float>(), dy.data_ptr<float>(),
                                        row_indices.data_ptr<int>(), row_offsets.data_ptr<int>(),
                                        dx.data_ptr<float>(), current_stream()),
           "sparse_softmax_backward");
  return dx;
}

// values[..., perm] with the permutation of a planned transpose (tiled view for large inputs).
static torch::Tensor plan_permute_values_impl(tsp_plan* plan, const torch::Tensor& values) {
  torch::Tensor out = torch::empty_like(values);
  const int replicas = values.dim() == 2 ? (int)values.size(0) : 1;
  if (values.scalar_type() != torch::kFloat32)
    check_rc(tsp_plan_permute_values_b16(plan, replicas, values.data_ptr(), out.data_ptr(), current_stream()),
             "permute_values");
  else
    check_rc(tsp_plan_permute_values_f32(plan, replicas, values.data_ptr<float>(), out.data_ptr<float>(),
                                         current_stream()),
             "permute_values");
  return out;
}

static torch::Tensor permute_values_plain(const torch::Tensor& values, const torch::Tensor& perm);

static torch::Tensor permute_values_impl(const torch::Tensor& values, const torch::Tensor& perm) {
  {
    // the perm of a cached transpose (ops.transposed_topology / csr_transpose_with_perm): take its plan
    PlanCache& c = plan_cache();
    std::shared_ptr<PlanHolder> owner;
    {
      std::lock_guard<std::mutex> lock(c.mu);
      auto it = c.by_perm.find(perm.data_ptr());
      if (it != c.by_perm.end()) {
        owner = it->second.lock();
        if (!owner) c.by_perm.erase(it);
      }
    }
    if (owner && perm._version() == 0 && !current_stream_capturing())
      return plan_permute_values_impl(owner->plan, values);
  }
  return permute_values_plain(values, perm);
}

static torch::Tensor permute_values_plain(const torch::Tensor& values, const torch::Tensor& perm) {
  torch::Tensor out = torch::empty_like(values);
  const int replicas = values.dim() == 2 ? (int)values.size(0) : 1;
  if (values.scalar_type() != torch::kFloat32)
    check_rc(tsp_permute_values_b16((int)perm.size(0), replicas, values.data_ptr(), perm.data_ptr<int>(), out.data_ptr(),
                                    current_stream()),
             "permute_values");
  else
    check_rc(tsp_permute_values_f32((int)perm.size(0), replicas, values.data_ptr<float>(), perm.data_ptr<int>(),
                                    out.data_ptr<float>(), current_stream()),
             "permute_values");
  return out;
}

static std::vector<torch::Tensor> transpose_impl(int m, int n, const torch::Tensor& values_in,
                                                 const torch::Tensor& row_offsets,
                                                 const torch::Tensor& column_indices,
                                                 bool want_perm) {
  // The reference handles 1-D values only (src/transpose_cuda.cu:58); its own 3-D backward
  // passes hand it [R, nnz] (modules/spmm.py:59) and get garbage.  Here [R, nnz] is permuted
  // replica by replica with the shared topology.
  TORCH_CHECK(values_in.dim() == 1 || values_in.dim() == 2,
              "torch_sputnik.csr_transpose: values must have 1 or 2 dimensions");
  torch::Tensor values = value_operand(values_in, "values");
  TSP_CHECK_CUDA_I32(row_offsets, "row_offsets");
  TSP_CHECK_CUDA_I32(column_indices, "column_indices");
  TORCH_CHECK(row_offsets.dim() == 1 && row_offsets.size(0) == m + 1,
              "torch_sputnik.csr_transpose: expected m+1 row offsets");
  TORCH_CHECK(column_indices.dim() == 1, "torch_sputnik.csr_transpose: column_indices must be 1-D");
  const int64_t nnz = column_indices.size(0);
  TORCH_CHECK(values.size(-1) == nnz,
              "torch_sputnik.csr_transpose: expected same number of values and indices");
  c10::cuda::CUDAGuard guard(values.device());
  const int64_t replicas = values.dim() == 2 ? values.size(0) : 1;

  // Planned topology: the sort ran once; every later call (one per backward pass in
  // modules/spmm.py:59, sddmm.py:60, sparse_linear.py:52) is a gather through the cached
  // permutation, and returns the plan's own index arrays -- shared between calls, do not edit.
  if (nnz > 0) {
    if (auto e = get_plan(m, n, row_offsets, column_indices)) {
      ensure_transposed(e, m, n, values.device());
      // (under CUDA-graph capture: the plain gather, which asks the plan nothing)
      std::vector<torch::Tensor> out{current_stream_capturing() ? permute_values_plain(values, e->perm)
                                                                : plan_permute_values_impl(e->plan, values),
                                     e->ro_t, e->ci_t};
      if (want_perm) out.push_back(e->perm);
      return out;
    }
  }

  auto iopt = torch::TensorOptions().dtype(torch::kInt32).device(values.device());
  if (values.scalar_type() != torch::kFloat32) {
    // 16-bit values, unplanned: sort the topology alone, then move the values through the permutation
    torch::Tensor row_offsets_t = torch::empty({n + 1}, iopt);
    torch::Tensor column_indices_t = torch::empty({nnz}, iopt);
    torch::Tensor perm = torch::empty({nnz}, iopt);
    const size_t wsb = tsp_csr_transpose_workspace(m, n, (int)nnz);
    torch::Tensor ws = torch::empty({(int64_t)wsb + 256}, torch::TensorOptions().dtype(torch::kUInt8).device(values.device()));
    check_rc(tsp_csr_transpose_f32(m, n, (int)nnz, 0, nullptr, row_offsets.data_ptr<int>(), column_indices.data_ptr<int>(),
                                   nullptr, row_offsets_t.data_ptr<int>(), column_indices_t.data_ptr<int>(),
                                   perm.data_ptr<int>(), ws.data_ptr(), wsb + 256, current_stream()),
             "csr_transpose");
    std::vector<torch::Tensor> out{nnz > 0 ? permute_values_impl(values, perm) : torch::empty_like(values), row_offsets_t,
                                   column_indices_t};
    if (want_perm) out.push_back(perm);
    return out;
  }
  torch::Tensor values_t = torch::empty_like(values);
  torch::Tensor row_offsets_t = torch::empty({n + 1}, iopt);
  torch::Tensor column_indices_t = torch::empty({nnz}, iopt);
  torch::Tensor perm;
  if (want_perm) perm = torch::empty({nnz}, iopt);
  const size_t ws_bytes = tsp_csr_transpose_workspace(m, n, (int)nnz);
  torch::Tensor ws = torch::empty({(int64_t)ws_bytes},
                                  torch::TensorOptions().dtype(torch::kUInt8).device(values.device()));
  check_rc(tsp_csr_transpose_f32(m, n, (int)nnz, (int)replicas, values.data_ptr<float>(),
                                 row_offsets.data_ptr<int>(), column_indices.data_ptr<int>(),
                                 values_t.data_ptr<float>(), row_offsets_t.data_ptr<int>(),
                                 column_indices_t.data_ptr<int>(),
                                 want_perm ? perm.data_ptr<int>() : nullptr, ws.data_ptr(), ws_bytes,
                                 current_stream()),
           "csr_transpose");
  std::vector<torch::Tensor> out{values_t, row_offsets_t, column_indices_t};
  if (want_perm) out.push_back(perm);
  return out;
}

// Extension: [row_offsets_t, column_indices_t, perm, row_indices_t] of a topology from its plan
// (row_indices_t = heaviest-first swizzle of the transpose); what a backward pass needs to multiply
// by A^T with spmm_perm, without touching any values.
std::vector<torch::Tensor> transposed_topology(int m, int n, torch::Tensor row_offsets, torch::Tensor column_indices) {
  TSP_CHECK_CUDA_I32(row_offsets, "row_offsets");
  TSP_CHECK_CUDA_I32(column_indices, "column_indices");
  TORCH_CHECK(row_offsets.dim() == 1 && row_offsets.size(0) == m + 1,
              "torch_sputnik.transposed_topology: expected m+1 row offsets");
  c10::cuda::CUDAGuard guard(row_offsets.device());
  std::shared_ptr<PlanEntry> e;
  if (column_indices.size(0) > 0) e = get_plan(m, n, row_offsets, column_indices);
  if (e) {
    ensure_transposed(e, m, n, row_offsets.device());
    return {e->ro_t, e->ci_t, e->perm, e->ri_t};
  }
  torch::Tensor probe = torch::empty({column_indices.size(0)}, torch::TensorOptions().dtype(torch::kFloat32).device(row_offsets.device()));
  std::vector<torch::Tensor> t = transpose_impl(m, n, probe, row_offsets, column_indices, true);
  return {t[1], t[2], t[3], heaviest_first(t[1])};
}

// src/sputnik.cpp:30-34 -> src/transpose_cuda.cu:45-102
std::vector<torch::Tensor> csr_transpose(int m, int n, torch::Tensor values,
                                         torch::Tensor row_offsets, torch::Tensor column_indices) {
  return transpose_impl(m, n, values, row_offsets, column_indices, false);
}

// Extension: also returns perm (values_t = values[..., perm]) so a static topology is sorted once.
std::vector<torch::Tensor> csr_transpose_with_perm(int m, int n, torch::Tensor values,
                                                   torch::Tensor row_offsets,
                                                   torch::Tensor column_indices) {
  return transpose_impl(m, n, values, row_offsets, column_indices, true);
}

torch::Tensor permute_values(torch::Tensor values_in, torch::Tensor perm) {
  torch::Tensor values = value_operand(values_in, "values");
  TSP_CHECK_CUDA_I32(perm, "perm");
  TORCH_CHECK(values.dim() == 1 || values.dim() == 2, "torch_sputnik.permute_values: values must be 1-D or 2-D");
  TORCH_CHECK(perm.dim() == 1 && perm.size(0) == values.size(-1),
              "torch_sputnik.permute_values: perm must have nnz elements");
  c10::cuda::CUDAGuard guard(values.device());
  return permute_values_impl(values, perm);
}

// Extension: spmm with the transposed topology of a cached csr_transpose and the ORIGINAL values
// (nonzero p reads values[..., perm[p]]): A^T @ dense without the permute_values pass.
torch::Tensor spmm_perm(int m, int k, torch::Tensor values, torch::Tensor perm, torch::Tensor row_indices,
                        torch::Tensor row_offsets, torch::Tensor column_indices, torch::Tensor dense) {
  return spmm_impl(m, k, values, row_indices, row_offsets, column_indices, dense, c10::nullopt, false,
                   false, "spmm_perm", perm);
}

// Extension: modules/sparse_attention.py:66-82 in one kernel.  q [R, m, d],
// k and v [R, n, d] (or 2-D for R = 1) -> [out [R, m, d], lse [R, m]].  Head dimensions the fused
// kernel does not take go through the three-op chain (lse is then left empty).
std::vector<torch::Tensor> sparse_attention(int m, int n, torch::Tensor row_indices_in, torch::Tensor row_offsets,
                                            torch::Tensor column_indices, torch::Tensor q_in, torch::Tensor k_in,
                                            torch::Tensor v_in, double scale) {
  torch::Tensor q = value_operand(q_in, "q"), k = value_operand(k_in, "k"), v = value_operand(v_in, "v");
  TORCH_CHECK(q.scalar_type() == k.scalar_type() && q.scalar_type() == v.scalar_type(),
              "torch_sputnik.sparse_attention: q, k, v must have the same dtype");
  const auto dtype = q.scalar_type();
  TORCH_CHECK(q.dim() == k.dim() && q.dim() == v.dim() && (q.dim() == 2 || q.dim() == 3),
              "torch_sputnik.sparse_attention: q, k, v must all be 2-D or all 3-D");
  c10::cuda::CUDAGuard guard(q.device());
  torch::Tensor row_indices = as_i32_indices(row_indices_in, "row_indices");
  check_topology(m, row_indices, row_offsets, column_indices);
  const int off = q.dim() - 2;
  const int64_t replicas = off ? q.size(0) : 1, d = q.size(off + 1);
  TORCH_CHECK(q.size(off) == m && k.size(off) == n && v.size(off) == n && k.size(off + 1) == d &&
                  v.size(off + 1) == d && (!off || (k.size(0) == replicas && v.size(0) == replicas)),
              "torch_sputnik.sparse_attention: expected q [R, m, d], k [R, n, d], v [R, n, d]");
  const int64_t nnz = column_indices.size(0);
  const bool aligned = ((reinterpret_cast<uintptr_t>(q.data_ptr()) | reinterpret_cast<uintptr_t>(k.data_ptr()) |
                         reinterpret_cast<uintptr_t>(v.data_ptr())) & 15) == 0;
  const bool fused = aligned && (dtype == torch::kFloat32 ? tsp_sparse_attention_supported(n, (int)d)
                                                          : tsp_sparse_attention_supported_b16(n, (int)d));
  if (!fused) {
    torch::Tensor scores = sddmm_impl(m, n, row_indices, row_offsets, column_indices, q, k, false);
    torch::Tensor probs = softmax_impl(scores, row_indices, row_offsets, column_indices, scale);
    torch::Tensor out = spmm_impl(m, n, probs, row_indices, row_offsets, column_indices, v, c10::nullopt, false,
                                  false, "sparse_attention");
    return {out.reshape(q.sizes()), torch::empty({0}, q.options().dtype(torch::kFloat32))};
  }
  torch::Tensor out = torch::empty_like(q);
  torch::Tensor lse = torch::empty(off ? std::vector<int64_t>{replicas, m} : std::vector<int64_t>{m},
                                   q.options().dtype(torch::kFloat32));
  if (dtype == torch::kFloat32)
    check_rc(tsp_sparse_attention_f32(m, n, (int)d, (int)nnz, (int)replicas, (float)scale, row_indices.data_ptr<int>(),
                                      row_offsets.data_ptr<int>(), column_indices.data_ptr<int>(),
                                      q.data_ptr<float>(), k.data_ptr<float>(), v.data_ptr<float>(),
                                      out.data_ptr<float>(), lse.data_ptr<float>(), current_stream()),
             "sparse_attention");
  else
    check_rc(tsp_sparse_attention_b16(dtype == torch::kBFloat16 ? 1 : 0, m, n, (int)d, (int)nnz, (int)replicas,
                                      (float)scale, row_indices.data_ptr<int>(), row_offsets.data_ptr<int>(),
                                      column_indices.data_ptr<int>(), q.data_ptr(), k.data_ptr(), v.data_ptr(),
                                      out.data_ptr(), lse.data_ptr<float>(), current_stream()),
             "sparse_attention");
  return {out, lse};
}

#include "many_mask.inc"

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.doc() = "B200-native CSR sparse kernels behind the torch_sputnik operator API";
  m.def("spmm", &spmm, "Sparse Matrix Matrix Multiplication: AxB");
  m.def("left_spmm", &left_spmm, "Sparse Matrix Matrix Multiplication: A(2d)xB(3d)");
  m.def("sddmm", &sddmm, "Sampled Dense Dense Matrix Multiplication: (AxB).C = D");
  m.def("sparse_softmax", &sparse_softmax,
        "Computes softmax function across the last dim of a sparse matrix");
  m.def("csr_transpose", &csr_transpose, "Transpose sparse matrix");
  m.def("spmm_bias_relu", &spmm_bias_relu, "relu(AxB + bias[:, None])");
  m.def("spmm_bias", &spmm_bias_relu, "alias of spmm_bias_relu (tests/test_spmm_bias_relu.py:37)");
  m.def("sparse_softmax_scaled", &sparse_softmax_scaled, "softmax(scale * x) over CSR rows");
  m.def("sparse_softmax_backward", &sparse_softmax_backward, "gradient of sparse_softmax_scaled");
  m.def("csr_transpose_with_perm", &csr_transpose_with_perm, "csr_transpose + permutation");
  m.def("permute_values", &permute_values, "values[..., perm]");
  m.def("sddmm_reduce", &sddmm_reduce, "sum over replicas of sddmm -> [nnz] (sparse-linear weight gradient)");
  m.def("spmm_perm", &spmm_perm, "spmm reading values through a permutation (transposed topology, original values)");
  m.def("sparse_attention", &sparse_attention, "fused sddmm -> softmax(scale) -> spmm; returns [out, lse]");
  m.def("spmm_many_mask", &spmm_many_mask, "spmm with one CSR mask per batch element");
  m.def("sddmm_many_mask", &sddmm_many_mask, "sddmm with one CSR mask per batch element");
  m.def("sparse_softmax_many_mask", &sparse_softmax_many_mask, "sparse_softmax with one CSR mask per batch element");
  m.def("sparse_softmax_many_mask_scaled", &sparse_softmax_many_mask_scaled,
        "sparse_softmax_many_mask of scale * values");
  m.def("csr_transpose_many_mask", &csr_transpose_many_mask, "csr_transpose with one CSR mask per batch element");
  m.def("set_variant", [](const std::string& op, int v) { return tsp_set_variant(op.c_str(), v); },
        "force a kernel variant (benchmarking only)");
  m.def("launch_count", []() { return tsp_launch_count(); }, "kernels launched so far");
  m.def("abi_version", []() { return tsp_version(); });
  m.def("transposed_topology", &transposed_topology,
        "[row_offsets_t, column_indices_t, perm, row_indices_t] of a topology, from its cached plan");
  m.def("plan_policy", [](int policy) {
          PlanCache& c = plan_cache();
          std::lock_guard<std::mutex> lock(c.mu);
          const int old = c.policy;
          if (policy >= 0) c.policy = policy;
          return old;
        },
        "0: never plan (every call takes the unplanned C-ABI entry), 1: plan every topology on first sight "
        "(default); negative: query.  Returns the previous policy");
  m.def("plan_cache_clear", []() {
          PlanCache& c = plan_cache();
          std::lock_guard<std::mutex> lock(c.mu);
          c.entries.clear();
        },
        "drop every cached plan");
  m.def("plan_cache_stats", []() {
          PlanCache& c = plan_cache();
          std::lock_guard<std::mutex> lock(c.mu);
          return std::vector<int64_t>{(int64_t)c.entries.size(), (int64_t)c.hits, (int64_t)c.misses,
                                      (int64_t)c.evictions};
        },
        "[entries, hits, misses, evictions]");
}
