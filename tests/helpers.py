"""Shared helpers for the parity tests."""
import numpy as np
import torch

import oracle
from oracle.csr import dense_operand, random_csr  # noqa: F401

# Tolerances (north_star / SURVEY.md 8c).
REL_FWD = 1e-5      # |x - ref64| <= REL_FWD * sum_i |a_i b_i|   (forward-error scale of the dot)
ABS_SPUTNIK = 1e-4  # Sputnik gtests, inputs in (-1, 1)
ABS_TORCH = 1e-2    # Torch-Sputnik test scripts


def to_dev(csr, dev):
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    return t(csr.values), t(csr.row_indices), t(csr.row_offsets), t(csr.column_indices)


def assert_close_fwd(got, ref, scale, rel=REL_FWD, what=""):
    """got, ref, scale: numpy arrays; scale = sum |a_i b_i| per element."""
    got = np.asarray(got, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    bound = rel * np.asarray(scale, dtype=np.float64) + 1e-30
    err = np.abs(got - ref)
    worst = np.max(err / bound) if err.size else 0.0
    assert worst <= 1.0, f"{what}: max err/bound = {worst:.3f} (max abs err {err.max():.3e})"


def spmm_case(m, k, n, density=None, nnz=None, seed=0, values="pm1", swizzle="heaviest", round_to=1):
    a = random_csr(m, k, density, nnz=nnz, seed=seed, values=values, swizzle=swizzle, round_to=round_to)
    b = dense_operand((k, n), seed + 1000, values)
    return a, b
