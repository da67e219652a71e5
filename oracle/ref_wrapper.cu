// ref_wrapper.cu -- extern "C" doorway into the UNMODIFIED reference kernels.
//
// TEST / MEASUREMENT INFRASTRUCTURE ONLY (see oracle/README.md).  oracle/Makefile compiles the
// reference's own Sputnik sources where they lie under /root/reference/third_party/sputnik for
// sm_100a, together with this file, into oracle/_ref/libsputnik_ref.so.  Nothing here is the
// reference's code: it only forwards to sputnik::CudaSpmm / CudaSpmmBiasRelu / CudaSddmm /
// SparseSoftmax (the symbols src/*.cu of the reference binds) and to cuSPARSE's
// cusparseCsr2cscEx2(ALG1) exactly as src/transpose_cuda.cu:22-31,90-99 calls it, reproducing
// the binding's per-replica host loop (src/spmm_cuda.cu:48-57).
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cusparse.h>

#include "sputnik/sputnik.h"

extern "C" {

int ref_spmm(int m, int k, int n, int nnz, int replicas, int values_shared, const int* row_indices,
             const float* values, const int* row_offsets, const int* column_indices,
             const float* dense, const float* bias_or_null, float* out, void* stream) {
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  for (int r = 0; r < replicas; ++r) {
    const float* v = values + (values_shared ? 0 : (size_t)nnz * r);
    cudaError_t e;
    if (bias_or_null)
      e = sputnik::CudaSpmmBiasRelu(m, k, n, nnz, row_indices, v, row_offsets, column_indices,
                                    dense + (size_t)k * n * r, bias_or_null,
                                    out + (size_t)m * n * r, s);
    else
      e = sputnik::CudaSpmm(m, k, n, nnz, row_indices, v, row_offsets, column_indices,
                            dense + (size_t)k * n * r, out + (size_t)m * n * r, s);
    if (e != cudaSuccess) return (int)e;
  }
  return 0;
}

// The reference's half-precision SpMM (cuda_spmm.cu.cc:511-570; contract cuda_spmm.h:75-127: every
// row holds an even number of nonzeros, n is even, column indices are 16-bit).  The Torch-Sputnik
// binding never reaches it (src/spmm_cuda.cu:41-55 hard-codes float); it pins the 16-bit entry points.
int ref_spmm_half(int m, int k, int n, int nnz, const int* row_indices, const void* values,
                  const int* row_offsets, const void* column_indices_i16, const void* dense,
                  const float* bias_or_null, void* out, void* stream) {
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  cudaError_t e;
  if (bias_or_null)
    e = sputnik::CudaSpmmBiasRelu(m, k, n, nnz, row_indices, static_cast<const half2*>(values), row_offsets,
                                  static_cast<const short2*>(column_indices_i16), static_cast<const half2*>(dense),
                                  bias_or_null, static_cast<half2*>(out), s);
  else
    e = sputnik::CudaSpmm(m, k, n, nnz, row_indices, static_cast<const half2*>(values), row_offsets,
                          static_cast<const short2*>(column_indices_i16), static_cast<const half2*>(dense),
                          static_cast<half2*>(out), s);
  return (int)e;
}

int ref_sddmm(int m, int k, int n, int nnz, int replicas, const int* row_indices,
              const int* row_offsets, const int* column_indices, const float* lhs,
              const float* rhs, float* out, void* stream) {
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  for (int r = 0; r < replicas; ++r) {
    cudaError_t e = sputnik::CudaSddmm(m, k, n, nnz, row_indices, row_offsets, column_indices,
                                       lhs + (size_t)m * k * r, rhs + (size_t)k * n * r,
                                       out + (size_t)nnz * r, s);
    if (e != cudaSuccess) return (int)e;
  }
  return 0;
}

int ref_sparse_softmax(int m, int nnz, int replicas, const float* values, const int* row_indices,
                       const int* row_offsets, const int* column_indices, float* out,
                       void* stream) {
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  for (int r = 0; r < replicas; ++r) {
    cudaError_t e = sputnik::SparseSoftmax(m, -1, nnz, values + (size_t)nnz * r, row_indices,
                                           row_offsets, column_indices, out + (size_t)nnz * r, s);
    if (e != cudaSuccess) return (int)e;
  }
  return 0;
}

// cuSPARSE csr2csc, ALG1, fp32, base 0.  Allocates and frees its own workspace (synchronous).
int ref_csr_transpose(int m, int n, int nnz, const float* values, const int* row_offsets,
                      const int* column_indices, float* values_t, int* row_offsets_t,
                      int* column_indices_t, void* stream) {
  static cusparseHandle_t handle = nullptr;
  if (!handle && cusparseCreate(&handle) != CUSPARSE_STATUS_SUCCESS) return -1;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (cusparseSetStream(handle, s) != CUSPARSE_STATUS_SUCCESS) return -2;
  size_t bytes = 0;
  if (cusparseCsr2cscEx2_bufferSize(handle, m, n, nnz, values, row_offsets, column_indices,
                                    values_t, row_offsets_t, column_indices_t, CUDA_R_32F,
                                    CUSPARSE_ACTION_NUMERIC, CUSPARSE_INDEX_BASE_ZERO,
                                    CUSPARSE_CSR2CSC_ALG1, &bytes) != CUSPARSE_STATUS_SUCCESS)
    return -3;
  void* ws = nullptr;
  if (cudaMalloc(&ws, bytes ? bytes : 4) != cudaSuccess) return -4;
  cudaMemsetAsync(ws, 0, bytes ? bytes : 4, s);
  cusparseStatus_t st = cusparseCsr2cscEx2(
      handle, m, n, nnz, values, row_offsets, column_indices, values_t, row_offsets_t,
      column_indices_t, CUDA_R_32F, CUSPARSE_ACTION_NUMERIC, CUSPARSE_INDEX_BASE_ZERO,
      CUSPARSE_CSR2CSC_ALG1, ws);
  cudaStreamSynchronize(s);
  cudaFree(ws);
  return st == CUSPARSE_STATUS_SUCCESS ? 0 : -5;
}

// Same, with a caller-provided workspace, for timing (no allocation in the timed region).
size_t ref_csr_transpose_workspace(int m, int n, int nnz, const float* values,
                                   const int* row_offsets, const int* column_indices,
                                   float* values_t, int* row_offsets_t, int* column_indices_t) {
  static cusparseHandle_t handle = nullptr;
  if (!handle && cusparseCreate(&handle) != CUSPARSE_STATUS_SUCCESS) return 0;
  size_t bytes = 0;
  cusparseCsr2cscEx2_bufferSize(handle, m, n, nnz, values, row_offsets, column_indices, values_t,
                                row_offsets_t, column_indices_t, CUDA_R_32F,
                                CUSPARSE_ACTION_NUMERIC, CUSPARSE_INDEX_BASE_ZERO,
                                CUSPARSE_CSR2CSC_ALG1, &bytes);
  return bytes;
}

int ref_csr_transpose_ws(int m, int n, int nnz, const float* values, const int* row_offsets,
                         const int* column_indices, float* values_t, int* row_offsets_t,
                         int* column_indices_t, void* ws, void* stream) {
  static cusparseHandle_t handle = nullptr;
  if (!handle && cusparseCreate(&handle) != CUSPARSE_STATUS_SUCCESS) return -1;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (cusparseSetStream(handle, s) != CUSPARSE_STATUS_SUCCESS) return -2;
  cusparseStatus_t st = cusparseCsr2cscEx2(
      handle, m, n, nnz, values, row_offsets, column_indices, values_t, row_offsets_t,
      column_indices_t, CUDA_R_32F, CUSPARSE_ACTION_NUMERIC, CUSPARSE_INDEX_BASE_ZERO,
      CUSPARSE_CSR2CSC_ALG1, ws);
  return st == CUSPARSE_STATUS_SUCCESS ? 0 : -5;
}

}  // extern "C"
