"""CPU: the C-ABI library loads and exports every symbol include/tsp.h declares."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "tsp.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(tsp_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_reference_surface():
    names = _declared()
    for must in ("tsp_spmm_f32", "tsp_sddmm_f32", "tsp_csr_softmax_f32", "tsp_csr_transpose_f32",
                 "tsp_csr_transpose_workspace", "tsp_last_error"):
        assert must in names


def test_library_exports_every_declared_symbol():
    from torch_sputnik_b200 import cabi
    lib = ctypes.CDLL(cabi.LIB_PATH)
    missing = [n for n in _declared() if not hasattr(lib, n)]
    assert not missing, f"libtsp.so lacks {missing}"
    assert set(cabi.SYMBOLS) == set(_declared())
    lib.tsp_version.restype = ctypes.c_int
    assert lib.tsp_version() == 2


def test_argument_errors_do_not_need_a_gpu():
    from torch_sputnik_b200 import cabi
    lib = cabi.load()
    rc = lib.tsp_spmm_f32(-1, 4, 4, 0, 1, 0, None, None, None, None, None, None, 0, None, None)
    assert rc == 1 and "negative" in cabi.last_error()
    rc = lib.tsp_sddmm_f32(4, 4, 4, 4, 70000, None, None, None, None, None, None, None)
    assert rc == 1 and "65535" in cabi.last_error()
    # counting path: a [chunks][n] table of at most 4 ints per nonzero; radix path (wide matrix): 16 B per nonzero
    assert 4 * 4096 <= lib.tsp_csr_transpose_workspace(4096, 4096, 1 << 20) <= 17 * (1 << 20)
    assert lib.tsp_csr_transpose_workspace(512, 100000, 1 << 20) >= 4 * 4 * (1 << 20)
    plan = ctypes.c_void_p()
    rc = lib.tsp_plan_create(-1, 4, 0, None, None, None, ctypes.byref(plan))
    assert rc == 1 and plan.value is None and "negative" in cabi.last_error()
    assert lib.tsp_plan_spmm_f32(None, 4, 1, 0, None, None, None, None, None, 0, None, None) == 1
    assert lib.tsp_plan_sorted(None) == -1
    lib.tsp_plan_destroy(None)


def test_torch_module_surface(ops):
    """Same op names as src/sputnik.cpp:36-42 of the reference (+ the unbound spmm_bias)."""
    for name in ("spmm", "left_spmm", "sddmm", "sparse_softmax", "csr_transpose", "spmm_bias",
                 "spmm_bias_relu"):
        assert callable(getattr(ops, name))
    import sys
    assert sys.modules["torch_sputnik"] is ops


def test_product_does_not_import_the_oracle():
    """The product path must never route through oracle/ (tier rule 3)."""
    pkg = os.path.join(ROOT, "torch_sputnik_b200")
    for base, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                text = open(os.path.join(base, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M), f
                assert "liboracle" not in text and "libsputnik_ref" not in text, f
