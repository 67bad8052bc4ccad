"""CPU test: the product library loads without a GPU and exports every symbol that
include/czb200.h declares (no compute call is made here)."""
import ctypes as C
import os
import re

import pytest

import cubez_b200
from cubez_b200 import ffi, lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "czb200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    names = re.findall(r"\b([a-z][a-z0-9_]*)\s*\(", src)
    return sorted({n for n in names if n.endswith("_") or n.startswith("czb_")})


@pytest.mark.parametrize("real", ["f32", "f64"])
def test_library_exports_every_declared_symbol(real):
    dll = lib.load(real)
    declared = _declared_symbols()
    assert len(declared) >= 35 + 20
    for name in declared:
        assert hasattr(dll, name), "libczb200_%s.so does not export %s" % (real, name)
    assert dll.czb_real_bytes() == (4 if real == "f32" else 8)


def test_ffi_table_covers_the_reference_boundary():
    # the 34 symbols the unmodified reference driver needs + blas_copy_in_
    needed = """bc_k_ blas_bicg_1_ blas_bicg_2_ blas_calc_ax_ blas_calc_rk_ blas_clear_ blas_copy_
      blas_dot1_ blas_dot2_ blas_triad_ calc_ax_maf_ calc_rk_maf_ err_t_ exact_t_ fileout_t_
      imask_k_ jacobi_ jacobi_maf_ pcr_ pcr_eda_ pcr_eda_maf_ pcr_esa_ pcr_esa_maf_ pcr_j_esa_
      pcr_maf_ pcr_rb_ pcr_rb_esa_ pcr_rb_esa_maf_ pcr_rb_maf_ psor2sma_core_
      psor2sma_core_maf_ psor_ psor_maf_ search_pivot_""".split()
    assert len(needed) == 34
    for n in needed:
        assert n in ffi.REFERENCE_SYMBOLS
    for n in ffi.CZB_SYMBOLS:
        assert n in _declared_symbols()


def test_missing_library_fails_loudly(monkeypatch):
    monkeypatch.setattr(lib, "_LIBS", {})
    monkeypatch.setattr(lib, "lib_path", lambda real="f32": "/nonexistent/libczb200.so")
    with pytest.raises(RuntimeError, match="no CPU/PyTorch fallback"):
        lib.load("f32")


def test_product_does_not_reference_the_oracle():
    """Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline may touch oracle/."""
    pkg = os.path.join(ROOT, "cubez_b200")
    for dp, _, fns in os.walk(pkg):
        for fn in fns:
            if fn.endswith((".py", ".cu", ".cuh", ".cpp", ".h")) or fn == "Makefile":
                txt = open(os.path.join(dp, fn), errors="ignore").read()
                assert "liboracle" not in txt and "oracle/" not in txt.replace("the CPU oracle", ""), os.path.join(dp, fn)
