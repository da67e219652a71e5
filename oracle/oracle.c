/*
 * oracle.c -- CPU restatement of the reference's arithmetic for the CSR hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product (torch_sputnik_b200/, libtsp.so) links,
 * imports or calls this file; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline
 * leg may.  Built into oracle/liboracle.so by oracle/Makefile (gcc -O2 -fopenmp).
 *
 * Every function restates a loop the reference uses as ITS ground truth (paths relative to
 * /root/reference):
 *   oracle_spmm      third_party/sputnik/sputnik/spmm/spmm_test.cu.cc:58-73  (fp64 accumulate
 *                    over a row's nonzeros in CSR order, cast to fp32) + the fused epilogue
 *                    of spmm/cuda_spmm.cu.cc:267-277 (bias[row] then max(.,0)).
 *   oracle_sddmm     third_party/sputnik/sputnik/sddmm/sddmm_test.cu.cc:59-73 (fp64 dot).
 *   oracle_csr_softmax   softmax/sparse_softmax.cu.cc:40-68 (max, sum exp, one reciprocal),
 *                    evaluated in fp64 and cast to fp32; and oracle_csr_softmax_f32, the
 *                    same three passes in the kernel's own fp32 lane order (32 strided
 *                    partials, xor butterfly 1,2,4,8,16), which the reference kernel matches
 *                    to the last bit when expf agrees.
 *   oracle_csr_transpose   the result cusparseCsr2cscEx2(ALG1) is documented to return and
 *                    src/transpose_cuda.cu:90-99 relies on: stable counting sort by column.
 *                    cuSPARSE is closed source and not under /root/reference; parity for it
 *                    is pinned by golden vectors produced by cuSPARSE itself on the GPU box
 *                    (tests/golden/make_golden_gpu.py).
 *
 * Parity pin: see tests/golden/README.md -- golden vectors are outputs of the reference's own
 * Sputnik kernels (compiled unmodified for sm_100a, oracle/_ref) and of cuSPARSE on a B200.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

void oracle_spmm(int m, int k, int n, const float* values, const int* row_offsets,
                 const int* column_indices, const float* dense, const float* bias_or_null,
                 int relu, float* out) {
  (void)k;
#pragma omp parallel for schedule(dynamic, 8)
  for (int i = 0; i < m; ++i) {
    for (int j = 0; j < n; ++j) {
      double accum = 0.0;
      for (int l = row_offsets[i]; l < row_offsets[i + 1]; ++l) {
        accum += (double)values[l] * (double)dense[(size_t)column_indices[l] * n + j];
      }
      float r = (float)accum;
      if (bias_or_null) r = r + bias_or_null[i];
      if (relu) r = r > 0.f ? r : 0.f;
      out[(size_t)i * n + j] = r;
    }
  }
}

/* Sum of |a_l * b_lj| in fp64: the forward-error scale the tolerance tests use. */
void oracle_spmm_abs(int m, int k, int n, const float* values, const int* row_offsets,
                     const int* column_indices, const float* dense, float* out) {
  (void)k;
#pragma omp parallel for schedule(dynamic, 8)
  for (int i = 0; i < m; ++i) {
    for (int j = 0; j < n; ++j) {
      double accum = 0.0;
      for (int l = row_offsets[i]; l < row_offsets[i + 1]; ++l) {
        accum += fabs((double)values[l] * (double)dense[(size_t)column_indices[l] * n + j]);
      }
      out[(size_t)i * n + j] = (float)accum;
    }
  }
}

void oracle_sddmm(int m, int k, int n, const int* row_offsets, const int* column_indices,
                  const float* lhs, const float* rhs, float* out) {
  (void)n;
#pragma omp parallel for schedule(dynamic, 8)
  for (int i = 0; i < m; ++i) {
    for (int j = row_offsets[i]; j < row_offsets[i + 1]; ++j) {
      const int idx_n = column_indices[j];
      double acc = 0.0;
      for (int l = 0; l < k; ++l) {
        acc += (double)lhs[(size_t)i * k + l] * (double)rhs[(size_t)idx_n * k + l];
      }
      out[j] = (float)acc;
    }
  }
}

void oracle_sddmm_abs(int m, int k, int n, const int* row_offsets, const int* column_indices,
                      const float* lhs, const float* rhs, float* out) {
  (void)n;
#pragma omp parallel for schedule(dynamic, 8)
  for (int i = 0; i < m; ++i) {
    for (int j = row_offsets[i]; j < row_offsets[i + 1]; ++j) {
      const int idx_n = column_indices[j];
      double acc = 0.0;
      for (int l = 0; l < k; ++l) {
        acc += fabs((double)lhs[(size_t)i * k + l] * (double)rhs[(size_t)idx_n * k + l]);
      }
      out[j] = (float)acc;
    }
  }
}

void oracle_csr_softmax(int m, const float* values, const int* row_offsets, float scale,
                        float* out) {
#pragma omp parallel for schedule(dynamic, 64)
  for (int i = 0; i < m; ++i) {
    const int b = row_offsets[i], e = row_offsets[i + 1];
    if (e <= b) continue;
    double mx = -INFINITY;
    for (int l = b; l < e; ++l) {
      const double x = (double)(values[l] * scale);
      mx = x > mx ? x : mx;
    }
    double norm = 0.0;
    for (int l = b; l < e; ++l) norm += exp((double)(values[l] * scale) - mx);
    norm = 1.0 / norm;
    for (int l = b; l < e; ++l) out[l] = (float)(exp((double)(values[l] * scale) - mx) * norm);
  }
}

/* fp32 in the kernel's lane order (sparse_softmax.cu.cc:40-68): lane t owns elements
 * t, t+32, ...; partial results meet in an xor butterfly with strides 1,2,4,8,16.  expf here
 * is the host libm's, which is correctly rounded almost everywhere but is NOT the device
 * expf, so this variant is used with a 2-ulp tolerance, not bit-exactly. */
void oracle_csr_softmax_f32(int m, const float* values, const int* row_offsets, float* out) {
  for (int i = 0; i < m; ++i) {
    const int b = row_offsets[i], n = row_offsets[i + 1] - b;
    if (n <= 0) continue;
    const float* in = values + b;
    float lane[32];
    for (int t = 0; t < 32; ++t) {
      float mx = -INFINITY;
      for (int idx = t; idx < n; idx += 32) mx = in[idx] > mx ? in[idx] : mx;
      lane[t] = mx;
    }
    for (int s = 1; s < 32; s *= 2) {
      float nxt[32];
      for (int t = 0; t < 32; ++t) nxt[t] = lane[t ^ s] > lane[t] ? lane[t ^ s] : lane[t];
      memcpy(lane, nxt, sizeof(lane));
    }
    const float mx = lane[0];
    for (int t = 0; t < 32; ++t) {
      float norm = 0.f;
      for (int idx = t; idx < n; idx += 32) norm += expf(in[idx] - mx);
      lane[t] = norm;
    }
    for (int s = 1; s < 32; s *= 2) {
      float nxt[32];
      for (int t = 0; t < 32; ++t) nxt[t] = lane[t] + lane[t ^ s];
      memcpy(lane, nxt, sizeof(lane));
    }
    const float inv = 1.0f / lane[0];
    for (int idx = 0; idx < n; ++idx) out[b + idx] = expf(in[idx] - mx) * inv;
  }
}

/* dx = scale * y * (dy - sum(dy*y)) per row, fp64. */
void oracle_csr_softmax_backward(int m, const float* y, const float* dy, const int* row_offsets,
                                 float scale, float* dx) {
#pragma omp parallel for schedule(dynamic, 64)
  for (int i = 0; i < m; ++i) {
    const int b = row_offsets[i], e = row_offsets[i + 1];
    double dot = 0.0;
    for (int l = b; l < e; ++l) dot += (double)y[l] * (double)dy[l];
    for (int l = b; l < e; ++l) dx[l] = (float)((double)scale * (double)y[l] * ((double)dy[l] - dot));
  }
}

/* Stable CSR (m x n) -> CSR of the transpose (n x m).  perm_or_null[p] = source nonzero. */
void oracle_csr_transpose(int m, int n, int nnz, const float* values, const int* row_offsets,
                          const int* column_indices, float* values_t, int* row_offsets_t,
                          int* column_indices_t, int* perm_or_null) {
  int* cursor = (int*)calloc((size_t)n + 1, sizeof(int));
  for (int j = 0; j < nnz; ++j) cursor[column_indices[j] + 1]++;
  for (int c = 0; c < n; ++c) cursor[c + 1] += cursor[c];
  memcpy(row_offsets_t, cursor, ((size_t)n + 1) * sizeof(int));
  for (int i = 0; i < m; ++i) {
    for (int j = row_offsets[i]; j < row_offsets[i + 1]; ++j) {
      const int p = cursor[column_indices[j]]++;
      values_t[p] = values[j];
      column_indices_t[p] = i;
      if (perm_or_null) perm_or_null[p] = j;
    }
  }
  free(cursor);
}
