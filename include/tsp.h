/*
 * tsp.h -- torch-free C-ABI of the B200-native CSR sparse-kernel library (libtsp.so).
 *
 * This is the drop-in boundary for the hot path of mabdullahsoyturk/Torch-Sputnik:
 * every entry point below replaces one call the reference's PyTorch binding makes
 * into libsputnik / cuSPARSE.  All pointers are DEVICE pointers (fp32 values, int32
 * indices, row-major dense operands), all launches are asynchronous on `stream`
 * (a cudaStream_t passed as void*) and no entry point synchronises the stream.
 * State and allocation: the plain entry points (tsp_spmm_f32, tsp_sddmm_f32, ...) keep no state;
 * those that need a view of the topology (the staged SpMM / SDDMM kernels: a sortedness word,
 * block starts, occupancy bitmaps) or split-K partial sums take it from STREAM-ORDERED scratch
 * (cudaMallocAsync / cudaFreeAsync on `stream`, no synchronisation).  The tsp_plan_* entry points
 * keep that topology state in a caller-owned handle instead: tsp_plan_create and the first use of
 * a view allocate with cudaMalloc (once per topology), after which a planned call allocates
 * nothing and launches only the operator's own kernel.
 * Operands are assumed FINITE where noted (tsp_spmm_f32).
 * Return value: 0 on success, non-zero on error; tsp_last_error() returns a thread-local message
 * for the last non-zero return.
 *
 * "replicas": the reference loops on the host over the leading dimension of 3-D
 * operands and launches one kernel per replica (src/spmm_cuda.cu:48-57).  Here the
 * replica count is an argument and one launch covers all replicas (grid.z).
 *
 * Reference interface each entry point replaces (paths relative to /root/reference):
 *   tsp_spmm_f32            sputnik::CudaSpmm / CudaSpmmBiasRelu
 *                           third_party/sputnik/sputnik/spmm/cuda_spmm.h:51-72,
 *                           called from src/spmm_cuda.cu:48-57 (spmm) and
 *                           src/left_replicated_spmm.cu:32-41 (left_spmm, values shared)
 *   tsp_sddmm_f32           sputnik::CudaSddmm
 *                           third_party/sputnik/sputnik/sddmm/cuda_sddmm.h:25-32,
 *                           called from src/sddmm_cuda.cu:45-54
 *   tsp_csr_softmax_f32     sputnik::SparseSoftmax
 *                           third_party/sputnik/sputnik/softmax/sparse_softmax.h:26-32,
 *                           called from src/softmax_cuda.cu:35-43
 *   tsp_csr_transpose_f32   cusparseCsr2cscEx2_bufferSize + cusparseCsr2cscEx2 (ALG1)
 *                           called from src/transpose_cuda.cu:22-31 and :90-99
 */
#ifndef TSP_H_
#define TSP_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TSP_VERSION 2

/* Error codes (also the return values). */
#define TSP_OK 0
#define TSP_ERR_INVALID_ARGUMENT 1
#define TSP_ERR_CUDA 2
#define TSP_ERR_WORKSPACE 3

/* Library version (TSP_VERSION this library was built with). */
int tsp_version(void);

/* Message for the last error returned on the calling thread ("" if none). */
const char* tsp_last_error(void);

/*
 * SpMM:  out[r] = A_r @ dense[r]  (+ bias[:, None], then max(., 0) when relu != 0)
 *   A_r is the m x k CSR matrix (row_offsets[m+1], column_indices[nnz]) with values
 *   values + r*nnz  (values_shared == 0: values is [replicas, nnz], spmm 3-D)  or
 *   values          (values_shared != 0: one [nnz] vector for every replica, left_spmm).
 *   dense is [replicas, k, n], out is [replicas, m, n]; every element of out is written
 *   (rows with no nonzeros produce bias/zero), so out may be uninitialised memory.
 *   row_indices[m] is the row swizzle (any permutation of 0..m-1; work item b handles
 *   row row_indices[b]).  It may be NULL (identity).  Results do not depend on it.
 *   bias_or_null: [m] fp32 indexed by OUTPUT ROW (cuda_spmm.cu.cc:267-277), or NULL.
 *   relu is honoured with or without bias.
 *   Non-finite operands: a row's result only ever involves the dense rows the row's own nonzeros
 *   name (as in the reference, cuda_spmm.cu.cc:224-238) -- an Inf/NaN in dense row c reaches
 *   exactly the output rows that hold a nonzero in column c, whichever kernel variant runs.
 */
int tsp_spmm_f32(int m, int k, int n, int nnz, int replicas, int values_shared,
                 const int* row_indices, const float* values, const int* row_offsets,
                 const int* column_indices, const float* dense,
                 const float* bias_or_null, int relu, float* out, void* stream);

/*
 * SpMM through a value permutation (extension): as tsp_spmm_f32, but nonzero p of the CSR arrays
 * takes its value from values[r, value_perm[p]].  With (row_offsets, column_indices, perm) the
 * output of tsp_csr_transpose_f32 for a static topology, this multiplies by A_r^T without first
 * writing the permuted [replicas, nnz] values (the backward passes of modules/spmm.py:59-67 and
 * modules/sddmm.py:60-68 do a csr_transpose per call).  value_perm_or_null == NULL: identical to
 * tsp_spmm_f32.
 */
int tsp_spmm_perm_f32(int m, int k, int n, int nnz, int replicas, int values_shared,
                      const int* row_indices, const float* values, const int* value_perm_or_null,
                      const int* row_offsets, const int* column_indices, const float* dense,
                      const float* bias_or_null, int relu, float* out, void* stream);

/*
 * SDDMM: out[r, j] = dot(lhs[r, row_of(j), :], rhs[r, column_indices[j], :])
 *   lhs is [replicas, m, k], rhs is [replicas, n, k] (i.e. the right operand already
 *   transposed, as in the reference), out is [replicas, nnz].  Exactly nnz values per
 *   replica are written.
 */
int tsp_sddmm_f32(int m, int k, int n, int nnz, int replicas,
                  const int* row_indices, const int* row_offsets,
                  const int* column_indices, const float* lhs, const float* rhs,
                  float* out, void* stream);

/*
 * SDDMM summed over replicas:  out[j] = sum_r dot(lhs[r, row_of(j), :], rhs[r, column_indices[j], :])
 *   out is ONE [nnz] vector.  This is the weight gradient of a sparse linear layer: the reference
 *   computes sddmm into [replicas, nnz] (modules/sparse_linear.py:44) and lets autograd sum over
 *   the batch dimension; here the replica loop is folded into the kernel's k loop, so the
 *   [replicas, nnz] intermediate never exists.  Arguments as tsp_sddmm_f32.
 */
int tsp_sddmm_reduce_f32(int m, int k, int n, int nnz, int replicas,
                         const int* row_indices, const int* row_offsets,
                         const int* column_indices, const float* lhs, const float* rhs,
                         float* out, void* stream);

/*
 * CSR softmax: for every row i and replica r, over seg = [row_offsets[i], row_offsets[i+1]):
 *   out[r, seg] = exp(scale*x - max(scale*x)) / sum(exp(scale*x - max(scale*x))),  x = values[r, seg]
 *   The reference has no scale (scale = 1.0f); scale folds the caller-side 1/sqrt(d)
 *   of modules/sparse_attention.py:72 into the same pass.
 */
int tsp_csr_softmax_f32(int m, int nnz, int replicas, float scale,
                        const float* values, const int* row_indices,
                        const int* row_offsets, float* out, void* stream);

/*
 * CSR softmax backward: dx[r, seg] = scale * y * (dy - sum(dy * y)),  y = softmax output.
 * (The reference has no usable gradient for its softmax -- SURVEY.md section 8(f) rank 1.)
 */
int tsp_csr_softmax_backward_f32(int m, int nnz, int replicas, float scale,
                                 const float* y, const float* dy,
                                 const int* row_indices, const int* row_offsets,
                                 float* dx, void* stream);

/*
 * Fused sparse attention forward (extension): the chain
 * sddmm -> sparse_softmax(scale) -> spmm of modules/sparse_attention.py:66-82 in one kernel,
 *   out[r, i, :] = sum_j softmax_j(scale * <q[r, i, :], k[r, col_j, :]>) * v[r, col_j, :]
 * over the nonzeros j of mask row i.  q [replicas, m, d], k and v [replicas, n, d],
 * out [replicas, m, d]; lse_or_null, if non-NULL, receives [replicas, m] log-sum-exp of the scaled
 * scores (what a backward pass needs to rebuild the probabilities).  Rows without nonzeros give
 * zeros (lse = -inf).  tsp_sparse_attention_supported(n, d) != 0 iff the head dimension is handled
 * (d = 32, 64, 96, 128 and n * d < 2^29); operands must be 16-byte aligned.
 */
int tsp_sparse_attention_supported(int n, int d);
/* 16-bit storage (bf16 == 0: fp16, != 0: bfloat16): head dimension 64, 128, 192 or 256; scores,
 * probabilities, the running maximum and all sums fp32; out rounded once; lse stays fp32. */
int tsp_sparse_attention_supported_b16(int n, int d);
int tsp_sparse_attention_b16(int bf16, int m, int n, int d, int nnz, int replicas, float scale,
                             const int* row_indices, const int* row_offsets,
                             const int* column_indices, const void* q, const void* k,
                             const void* v, void* out, float* lse_or_null, void* stream);
int tsp_sparse_attention_f32(int m, int n, int d, int nnz, int replicas, float scale,
                             const int* row_indices, const int* row_offsets,
                             const int* column_indices, const float* q, const float* k,
                             const float* v, float* out, float* lse_or_null, void* stream);

/*
 * CSR transpose (m x n  ->  n x m), stable in the original nonzero order, which is what
 * cusparseCsr2cscEx2(CUSPARSE_CSR2CSC_ALG1) returns: inside each output row the entries
 * appear in ascending original row.  values is [replicas, nnz] (replicas >= 1; the
 * topology is shared, each replica's values are permuted the same way).
 *   values_t[replicas, nnz], row_offsets_t[n+1], column_indices_t[nnz] are fully written.
 *   perm_or_null: if non-NULL receives perm[nnz] with values_t[p] = values[perm[p]]
 *   (lets a caller cache the permutation of a static topology).
 *   replicas == 0: topology only (values / values_t may be NULL).
 *   workspace: device scratch of at least tsp_csr_transpose_workspace(m, n, nnz) bytes,
 *   256-byte aligned; contents on entry are irrelevant.
 */
size_t tsp_csr_transpose_workspace(int m, int n, int nnz);
int tsp_csr_transpose_f32(int m, int n, int nnz, int replicas, const float* values,
                          const int* row_offsets, const int* column_indices,
                          float* values_t, int* row_offsets_t, int* column_indices_t,
                          int* perm_or_null, void* workspace, size_t workspace_bytes,
                          void* stream);

/*
 * tsp_csr_transpose_noalloc_f32: tsp_csr_transpose_f32 for callers that cannot allocate.  It has no counterpart in
 * the reference (whose binding cudaMallocs a cuSPARSE buffer on every call, src/transpose_cuda.cu:22-43): it was
 * added for code under CUDA-graph capture, tight serving loops and harnesses
 * that forbid cudaMalloc.  Neither the caller nor the library allocates: the sort's scratch is a static device
 * array of TSP_NOALLOC_SCRATCH_BYTES inside the library, held by every GPU that loads it until the process exits.
 *   Same arguments, results and stable order as tsp_csr_transpose_f32, minus the workspace.
 *   Limit: tsp_csr_transpose_workspace(m, n, nnz) <= TSP_NOALLOC_SCRATCH_BYTES (e.g. any 4096 x 4096 topology of up
 *   to ~2 M nonzeros); larger problems return TSP_ERR_WORKSPACE with a message naming the need and the limit.
 *   Calls no cudaMalloc / cudaFree / stream-ordered allocator.  All calls share the one array, which keeps the last
 *   call's contents (the kernels initialise everything they read): calls whose kernels may overlap on the GPU --
 *   different streams without ordering between them -- race on it; order them.
 */
#define TSP_NOALLOC_SCRATCH_BYTES (8u << 20)
int tsp_csr_transpose_noalloc_f32(int m, int n, int nnz, int replicas, const float* values,
                                  const int* row_offsets, const int* column_indices,
                                  float* values_t, int* row_offsets_t, int* column_indices_t,
                                  int* perm_or_null, void* stream);

/* values_t[r, p] = values[r, perm[p]] for a cached permutation (static topology). */
int tsp_permute_values_f32(int nnz, int replicas, const float* values, const int* perm,
                           float* values_t, void* stream);

/*
 * Many masks in ONE launch (the torch_sputnik.*_many_mask ops that tests/test_attention_many_masks.py:120-150
 * and tests/transformer/functions.py:20-135 of the reference call and src/sputnik.cpp never binds).
 *   b CSR masks of one shape m x n, `heads` replicas per mask: replica r (0 <= r < b*heads) uses mask r / heads.
 *   row_offsets[b, m+1] with LOCAL offsets per mask, row_indices[b, m] (or NULL), column_indices concatenated,
 *   nnz_prefix[b+1] a DEVICE array with the exclusive prefix of the masks' nonzero counts, max_nnz the largest count.
 *   Dense operands are regular [b*heads, rows, cols]; values / sampled outputs are RAGGED by replica: replica r's
 *   nnz(r/heads) values start at heads * nnz_prefix[r/heads] + (r % heads) * nnz(r/heads).
 *   tsp_permute_values_many_f32: perm is the concatenation of the masks' permutations (indices local to a mask).
 *   One launch of the L1-gather kernels per call, whatever b is (the reference's host loop per replica,
 *   src/spmm_cuda.cu:48-57, would be a loop per replica AND mask).
 */
int tsp_sddmm_many_f32(int b, int heads, int m, int k, int n, int max_nnz, const int* nnz_prefix,
                       const int* row_offsets, const int* column_indices, const float* lhs,
                       const float* rhs, float* out, void* stream);
int tsp_spmm_many_f32(int b, int heads, int m, int k, int n, int max_nnz, const int* nnz_prefix,
                      const int* row_indices, const float* values, const int* row_offsets,
                      const int* column_indices, const float* dense, float* out, void* stream);
int tsp_csr_softmax_many_f32(int b, int heads, int m, int max_nnz, const int* nnz_prefix, float scale,
                             const float* values, const int* row_offsets, float* out, void* stream);
int tsp_permute_values_many_f32(int b, int heads, int max_nnz, const int* nnz_prefix,
                                const float* values, const int* perm, float* values_t, void* stream);

/*
 * 16-bit storage, fp32 arithmetic.  Replaces
 * sputnik::CudaSpmm / CudaSpmmBiasRelu (half2 values, short2 indices, half2 dense; float bias)
 * third_party/sputnik/sputnik/spmm/cuda_spmm.cu.cc:511-570, contract cuda_spmm.h:75-127 -- which the
 * reference binding never reaches (src/spmm_cuda.cu:41-55 hard-codes float) -- and extends the same
 * storage contract to SDDMM and the CSR softmax, which Sputnik has in fp32 only.  values / dense /
 * lhs / rhs / out point at IEEE fp16 (..._f16) or bfloat16 (..._bf16) elements; indices stay int32
 * (no short2 packing, so k is not limited to 32767), any nnz per row, any n (16-byte-aligned rows
 * take the 8-element vector path).  Every product and sum is fp32, each result is rounded to the
 * storage type once.  Large dense-ish problems (where the fp32 path would stage dense tiles in shared
 * memory) widen the operands once into stream-ordered fp32 scratch, run the fp32 kernel and round the
 * result on the way back -- same arithmetic.  bias stays fp32.  Argument meaning otherwise as tsp_spmm_perm_f32 /
 * tsp_sddmm_f32 / tsp_csr_softmax_f32.  tsp_permute_values_b16 moves any 16-bit element type.
 */
int tsp_spmm_f16(int m, int k, int n, int nnz, int replicas, int values_shared,
                 const int* row_indices, const void* values, const int* value_perm_or_null,
                 const int* row_offsets, const int* column_indices, const void* dense,
                 const float* bias_or_null, int relu, void* out, void* stream);
int tsp_spmm_bf16(int m, int k, int n, int nnz, int replicas, int values_shared,
                  const int* row_indices, const void* values, const int* value_perm_or_null,
                  const int* row_offsets, const int* column_indices, const void* dense,
                  const float* bias_or_null, int relu, void* out, void* stream);
int tsp_sddmm_f16(int m, int k, int n, int nnz, int replicas, const int* row_indices,
                  const int* row_offsets, const int* column_indices, const void* lhs,
                  const void* rhs, void* out, void* stream);
int tsp_sddmm_bf16(int m, int k, int n, int nnz, int replicas, const int* row_indices,
                   const int* row_offsets, const int* column_indices, const void* lhs,
                   const void* rhs, void* out, void* stream);
int tsp_csr_softmax_f16(int m, int nnz, int replicas, float scale, const void* values,
                        const int* row_indices, const int* row_offsets, void* out, void* stream);
int tsp_csr_softmax_bf16(int m, int nnz, int replicas, float scale, const void* values,
                         const int* row_indices, const int* row_offsets, void* out, void* stream);
int tsp_permute_values_b16(int nnz, int replicas, const void* values, const int* perm,
                           void* values_t, void* stream);

/*
 * Per-topology plans ("optional transpose-plan cache"; the reference's
 * topology is fixed at module construction, modules/sparse_attention.py:49-52,
 * modules/sparse_linear.py:78-83, and its binding redoes the topology work on every call,
 * src/transpose_cuda.cu:10-43).
 *
 * tsp_plan_create(m, n, nnz, row_offsets, column_indices, stream, &plan): a handle for the m x n
 *   CSR topology.  row_offsets / column_indices are BORROWED device arrays that must stay valid and
 *   unchanged for the plan's lifetime.  Launches the one sortedness check on `stream`; the verdict
 *   reaches the host asynchronously (no synchronisation here), after which planned calls launch
 *   exactly one kernel.  The plan also owns, built on first use: the block-start / occupancy-bitmap
 *   views of the staged SDDMM kernels and the transposed topology + permutation.
 *   Thread-safe: concurrent planned calls on one plan are allowed (internal mutex around view
 *   creation); tsp_plan_destroy must not race with them.  The first use of a view allocates
 *   (cudaMalloc): do it once before capturing a CUDA graph (a missing view during capture makes the
 *   call take the view-free gather kernel instead).
 * tsp_plan_sorted: 1 columns ascending in every row, 0 not, -1 verdict not on the host yet.
 * tsp_plan_transpose: the plan of the TRANSPOSED topology (owned by `plan`, destroyed with it) and its
 *   arrays: row_offsets_t[n+1], column_indices_t[nnz], perm[nnz] with values_t[p] = values[perm[p]],
 *   bit-identical to tsp_csr_transpose_f32 / cusparseCsr2cscEx2(ALG1).  Any out pointer may be NULL.
 *   A^T @ B without materialising values_t: tsp_plan_spmm_f32(transposed, ..., values, perm, ...).
 * tsp_plan_spmm_f32 / tsp_plan_sddmm_f32: tsp_spmm_perm_f32 / tsp_sddmm_f32 (reduce == 0) /
 *   tsp_sddmm_reduce_f32 (reduce != 0) on the plan's topology; m, k (resp. m, n) and nnz come from the
 *   plan, `n` is the dense operand's column count (resp. `k` the inner dimension).
 * tsp_plan_permute_values_f32 / _b16: values_t[r, p] = values[r, perm[p]] with the permutation of the
 *   plan's transpose (built on first use, like tsp_plan_transpose) -- the value half of csr_transpose on a
 *   static topology (src/transpose_cuda.cu:90-99 redoes the whole csr2csc per call).  Same result as
 *   tsp_permute_values_f32 with that perm; large inputs go through a per-topology tiled view of the
 *   permutation (sources of every output tile in ascending order) instead of a sector-per-element gather.
 */
typedef struct tsp_plan tsp_plan;
int tsp_plan_create(int m, int n, int nnz, const int* row_offsets, const int* column_indices,
                    void* stream, tsp_plan** plan_out);
void tsp_plan_destroy(tsp_plan* plan);
int tsp_plan_sorted(tsp_plan* plan);
int tsp_plan_transpose(tsp_plan* plan, void* stream, tsp_plan** transposed_out,
                       const int** row_offsets_t, const int** column_indices_t, const int** perm);
int tsp_plan_spmm_f32(tsp_plan* plan, int n, int replicas, int values_shared,
                      const int* row_indices, const float* values, const int* value_perm_or_null,
                      const float* dense, const float* bias_or_null, int relu, float* out,
                      void* stream);
int tsp_plan_sddmm_f32(tsp_plan* plan, int k, int replicas, int reduce, const int* row_indices,
                       const float* lhs, const float* rhs, float* out, void* stream);
/* 16-bit storage on a plan: bf16 == 0 -> fp16 elements, != 0 -> bfloat16 (see tsp_spmm_f16). */
int tsp_plan_spmm_b16(tsp_plan* plan, int bf16, int n, int replicas, int values_shared,
                      const int* row_indices, const void* values, const int* value_perm_or_null,
                      const void* dense, const float* bias_or_null, int relu, void* out,
                      void* stream);
int tsp_plan_sddmm_b16(tsp_plan* plan, int bf16, int k, int replicas, const int* row_indices,
                       const void* lhs, const void* rhs, void* out, void* stream);
int tsp_plan_permute_values_f32(tsp_plan* plan, int replicas, const float* values, float* values_t,
                                void* stream);
int tsp_plan_permute_values_b16(tsp_plan* plan, int replicas, const void* values, void* values_t,
                                void* stream);

/*
 * Knobs for A/B benchmarking only (not part of the reference surface):
 *   tsp_set_variant("spmm", v) / ("sddmm", v): 0 = heuristic (default), other values force
 *   a kernel variant.  Returns the previous value, or -1 for unknown op.
 */
int tsp_set_variant(const char* op, int variant);

/* Number of kernel launches issued by this library on the calling process so far. */
unsigned long long tsp_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* TSP_H_ */
