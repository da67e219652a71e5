from .functional import (CsrSoftmax, ManyMaskTopology, Sddmm, SparseLinearFunction, Spmm,  # noqa: F401
                         TransposedTopology, dense_to_sparse, diffsort, many_mask_attention)
from .layers import SparseAttention, SparseLinear, generate_mask  # noqa: F401
