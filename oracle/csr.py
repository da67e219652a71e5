"""Seeded synthetic CSR inputs (TEST INFRASTRUCTURE -- see oracle/__init__.py).

Semantics follow the reference's test helpers (paths relative to /root/reference):

* mask: exactly ``nnz`` positions chosen uniformly without replacement, ``nnz`` rounded UP to a
  multiple of ``round_to`` -- tests/connectors.py:47-59 (``Uniform.__call__``); the same
  distribution as Sputnik's RANDOM_UNIFORM generator (sputnik/matrix_utils.cu.cc:200-218).
* CSR extraction: row-major scan, column indices ascending inside a row, int32 everywhere --
  tests/sparse_matrix.py:9-41 (``_dense_to_sparse``).
* row swizzle: ``heaviest`` = argsort(-row_len) (tests/sparse_matrix.py:22, Sputnik
  SortedRowSwizzle matrix_utils.cu.cc:302-317); ``ascending`` = what ``diffsort`` of
  modules/spmm.py:4-6 really returns (lightest row first).
* values: ``u01`` = U[0,1) (tests/initializers.py:29-31), ``pm1`` = U(-1,1)
  (matrix_utils.cu.cc:82,197).
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np


@dataclass
class Csr:
    m: int
    k: int
    values: np.ndarray          # float32 [nnz]
    row_indices: np.ndarray     # int32 [m]
    row_offsets: np.ndarray     # int32 [m+1]
    column_indices: np.ndarray  # int32 [nnz]

    @property
    def nnz(self) -> int:
        return int(self.column_indices.shape[0])

    def to_dense(self, values: np.ndarray | None = None) -> np.ndarray:
        v = self.values if values is None else values
        out = np.zeros((self.m, self.k), dtype=v.dtype)
        out[row_of(self.row_offsets), self.column_indices] = v
        return out


def row_of(row_offsets: np.ndarray) -> np.ndarray:
    """Row index of every nonzero (length nnz)."""
    lens = np.diff(row_offsets.astype(np.int64))
    return np.repeat(np.arange(lens.shape[0], dtype=np.int32), lens)


def sorted_row_swizzle(row_offsets: np.ndarray) -> np.ndarray:
    """Heaviest row first -- tests/sparse_matrix.py:22."""
    return np.argsort(-1 * np.diff(row_offsets), kind="stable").astype(np.int32)


def diffsort_ascending(row_offsets: np.ndarray) -> np.ndarray:
    """What modules/spmm.py:4-6 ``diffsort`` yields: argsort(descending) of MINUS the row
    length, i.e. lightest row first (ties in unspecified order; we use a stable sort)."""
    return np.argsort(np.diff(row_offsets), kind="stable").astype(np.int32)


def random_csr(m: int, k: int, density: float | None = None, *, nnz: int | None = None,
               seed: int = 0, round_to: int = 1, values: str = "u01",
               swizzle: str = "heaviest") -> Csr:
    rng = np.random.default_rng(seed)
    size = m * k
    if nnz is None:
        assert density is not None
        dormant = int(round((1.0 - density) * size))
        nnz = size - dormant
    if round_to > 1:
        nnz = (nnz + round_to - 1) // round_to * round_to
    nnz = max(0, min(size, nnz))
    if nnz * 2 > size:
        # choose the zeros instead (cheaper for dense masks)
        mask = np.ones(size, dtype=bool)
        mask[rng.choice(size, size - nnz, replace=False)] = False
        flat = np.flatnonzero(mask)
    else:
        flat = np.sort(rng.choice(size, nnz, replace=False))
    rows = (flat // k).astype(np.int64)
    cols = (flat % k).astype(np.int32)
    row_offsets = np.zeros(m + 1, dtype=np.int64)
    np.add.at(row_offsets, rows + 1, 1)
    row_offsets = np.cumsum(row_offsets).astype(np.int32)
    if values == "u01":
        vals = rng.random(nnz, dtype=np.float32)
        vals[vals == 0.0] = 0.5  # keep the mask and the value pattern identical
    elif values == "pm1":
        vals = (rng.random(nnz, dtype=np.float32) * 2.0 - 1.0).astype(np.float32)
        vals[vals == 0.0] = 0.5
    else:
        raise ValueError(values)
    if swizzle == "heaviest":
        ri = sorted_row_swizzle(row_offsets)
    elif swizzle == "ascending":
        ri = diffsort_ascending(row_offsets)
    elif swizzle == "identity":
        ri = np.arange(m, dtype=np.int32)
    else:
        raise ValueError(swizzle)
    return Csr(m, k, vals, ri, row_offsets, cols)


def even_rows(a: Csr) -> Csr:
    """Drop the last nonzero of every odd-length row: the layout constraint of the reference's half
    SpMM (third_party/sputnik/sputnik/spmm/cuda_spmm.h:75-113: half2 values, short2 indices)."""
    lens = np.diff(a.row_offsets)
    keep = np.ones(a.values.shape[0], dtype=bool)
    odd = np.flatnonzero(lens % 2 == 1)
    keep[a.row_offsets[odd + 1] - 1] = False
    ro = np.concatenate([[0], np.cumsum(lens - (lens % 2))]).astype(np.int32)
    return Csr(a.m, a.k, a.values[keep], sorted_row_swizzle(ro), ro, a.column_indices[keep])


def dense_operand(shape, seed: int, dist: str = "u01") -> np.ndarray:
    rng = np.random.default_rng(seed)
    x = rng.random(shape, dtype=np.float32)
    if dist == "pm1":
        x = (x * 2.0 - 1.0).astype(np.float32)
    return x


def from_dense_mask(mask: np.ndarray, seed: int = 0, values: str = "pm1") -> Csr:
    """CSR of an arbitrary boolean mask (skewed-pattern test cases), columns ascending per row."""
    m, k = mask.shape
    rows, cols = np.nonzero(mask)
    row_offsets = np.zeros(m + 1, dtype=np.int64)
    np.add.at(row_offsets, rows + 1, 1)
    row_offsets = np.cumsum(row_offsets).astype(np.int32)
    rng = np.random.default_rng(seed)
    vals = rng.random(rows.shape[0], dtype=np.float32)
    if values == "pm1":
        vals = (vals * 2.0 - 1.0).astype(np.float32)
    vals[vals == 0.0] = 0.5
    return Csr(m, k, vals, sorted_row_swizzle(row_offsets), row_offsets, cols.astype(np.int32))
