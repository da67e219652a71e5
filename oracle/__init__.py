"""oracle -- CPU restatement of the reference's arithmetic for the CSR hot path.

TEST INFRASTRUCTURE ONLY.  Only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may import this package; the
product (``torch_sputnik_b200``) never does and fails loudly without its CUDA library.

Layers (each function cites the reference lines it restates; paths relative to /root/reference):

* ``oracle.c`` via ctypes (``liboracle.so``): fp64-accumulate loops of Sputnik's own gtests
  (spmm_test.cu.cc:58-73, sddmm_test.cu.cc:59-73), the softmax of sparse_softmax.cu.cc:40-68,
  stable CSR transpose (= cusparseCsr2cscEx2 ALG1 as used at src/transpose_cuda.cu:90-99).
* ``dense_*`` functions: the dense torch masked-matmul references of the Torch-Sputnik test
  scripts (tests/test_spmm.py:9-10,32-34; tests/test_sddmm.py:8-13; tests/test_softmax.py:9-22).
* ``csr.py``: seeded synthetic CSR generators with the semantics of tests/connectors.py:47-59
  and tests/sparse_matrix.py:9-41.

Parity pin: tests/golden/*.npz are outputs of the reference's own kernels and cuSPARSE on a
B200 (tests/golden/make_golden_gpu.py); tests/test_oracle_golden.py checks this oracle
against them.
"""
from .cpu import (  # noqa: F401
    available,
    build,
    csr_softmax,
    csr_softmax_backward,
    csr_softmax_f32,
    csr_transpose,
    dense_sddmm,
    dense_softmax,
    dense_spmm,
    sddmm,
    sddmm_abs,
    spmm,
    spmm_abs,
)
from .csr import Csr, diffsort_ascending, random_csr, row_of, sorted_row_swizzle  # noqa: F401
