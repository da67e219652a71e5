from .functional import (CsrSoftmax, Sddmm, SparseLinearFunction, Spmm, TransposedTopology,  # noqa: F401
                         dense_to_sparse, diffsort)
from .layers import SparseAttention, SparseLinear, generate_mask  # noqa: F401
