"""CPU run of the DEVICE code of cubez_b200/csrc/maf.cu (tests/tools/emu: CUDA built-ins emulated, one thread per
block, launches executed sequentially) against the oracle, bit for bit, on a stretched mesh with seeded random fields.
Covers what a GPU-less container can: index arithmetic, launch geometry (make_tile), colour / wavefront line selection,
shared-memory ping-pong, the folded 2x2 last stage and the reduction finishers.  The -m gpu tests
(test_gpu_zz_maf.py) remain the parity tests proper."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

import oracle_util as ou

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "tools", "emu"))
import emu_build  # noqa: E402

_LIBS = {}


def _emu(real):
    if not _LIBS:
        _LIBS.update({k: C.CDLL(v) for k, v in emu_build.build().items()})
    lib = _LIBS[real]
    assert lib.emu_real_bytes() == (4 if real == "f32" else 8)
    return lib


def _case(real, sz):
    o = ou.OracleCZ(sz, real)
    X, B = ou.random_fields(o.shape, o.dt)
    mesh = o.mesh(stretch=0.3)
    return o, X, B, mesh, [ou.ptr(m) for m in mesh]


@pytest.mark.parametrize("real", ["f32", "f64"])
@pytest.mark.parametrize("sz", [[12, 7, 23], [9, 14, 37], [40, 13, 70]])
def test_emulated_maf_point_kernels_equal_oracle(real, sz):
    lib = _emu(real)
    o, X, B, mesh, mp = _case(real, sz)
    R = o.creal
    # jacobi_maf
    o.ac1 = o.dt(0.8)
    Xo = X.copy(); ro = C.c_double(0); o.sweep_maf("jacobi_maf", Xo, B, ro, mesh)
    Xn = X.copy(); res = C.c_double(-1)
    lib.emu_jacobi_maf(ou.ptr(X), ou.ptr(Xn), ou.ptr(B), o._sz, o._idx, *mp, R(0.8), C.byref(res))
    assert np.array_equal(Xn, Xo)
    assert res.value == pytest.approx(ro.value, rel=1e-12)
    # red-black, colour 0 then 1 accumulated
    o.ac1 = o.dt(1.5)
    Xo = X.copy(); ro = C.c_double(0); o.sweep_maf("sor2sma_maf", Xo, B, ro, mesh)
    Xe = X.copy()
    for color in (0, 1):
        lib.emu_rb_color_maf(ou.ptr(Xe), ou.ptr(B), o._sz, o._idx, *mp, R(1.5), C.c_int(color), C.c_int(color), C.byref(res))
    assert np.array_equal(Xe, Xo)
    assert res.value == pytest.approx(ro.value, rel=1e-12)
    # lexicographic SOR through hyperplane wavefronts
    Xo = X.copy(); ro = C.c_double(0); o.sweep_maf("psor_maf", Xo, B, ro, mesh)
    Xe = X.copy()
    lib.emu_psor_maf(ou.ptr(Xe), ou.ptr(B), o._sz, o._idx, *mp, R(1.5), C.byref(res))
    assert np.array_equal(Xe, Xo)
    assert res.value == pytest.approx(ro.value, rel=1e-12)
    # pivot, operator (+ fused dot), residual vector
    Po = o.pivot(mesh)
    Pe = np.zeros_like(X); lib.emu_search_pivot(ou.ptr(Pe), o._sz, o._idx, *mp)
    assert np.array_equal(Pe, Po)
    flop = C.c_double(0)
    Ao = np.zeros_like(X); o.dll.calc_ax_maf_(ou.ptr(Ao), ou.ptr(X), o._sz, o._idx, o.g(), *mp, ou.ptr(Po), C.byref(flop))
    Ae = np.zeros_like(X); dot = C.c_double(0)
    lib.emu_calc_ax_maf(ou.ptr(Ae), ou.ptr(X), o._sz, o._idx, *mp, ou.ptr(Pe), ou.ptr(B), C.byref(dot))
    assert np.array_equal(Ae, Ao)
    assert dot.value == pytest.approx(float(o.dot2(Ao, B)), rel=1e-5 if real == "f32" else 1e-12)
    Ro = np.zeros_like(X); o.dll.calc_rk_maf_(ou.ptr(Ro), ou.ptr(X), ou.ptr(B), o._sz, o._idx, o.g(), *mp, ou.ptr(Po), C.byref(flop))
    Re = np.zeros_like(X); lib.emu_calc_rk_maf(ou.ptr(Re), ou.ptr(X), ou.ptr(B), o._sz, o._idx, *mp, ou.ptr(Pe))
    assert np.array_equal(Re, Ro)


@pytest.mark.parametrize("real", ["f32", "f64"])
@pytest.mark.parametrize("sz", [[9, 9, 9], [12, 7, 23], [10, 8, 34], [8, 6, 66], [6, 6, 131], [7, 5, 300]])
def test_emulated_maf_line_kernel_equals_oracle(real, sz):
    """k_pcr_lines_maf: red-black (both colours) and the i+j wavefronts of the lexicographic routines; line lengths of
    2^m and 2^m +- 1 cells."""
    lib = _emu(real)
    o, X, B, mesh, mp = _case(real, sz)
    R = o.creal
    M = o.mask()
    o.ac1 = o.dt(1.3)
    pn = ou.get_num_stage(o.idx[5] - o.idx[4] + 1)
    res = C.c_double(-1)
    Xo = X.copy(); ro = C.c_double(0); o.sweep_maf("pcr_rb_maf", Xo, B, ro, mesh)
    Xe = X.copy()
    for color in (0, 1):
        lib.emu_pcr_maf(ou.ptr(Xe), ou.ptr(M), ou.ptr(B), o._sz, o._idx, C.c_int(pn), C.c_int(color), *mp, R(1.3),
                        C.c_int(color), C.byref(res))
    assert np.array_equal(Xe, Xo), np.abs(Xe - Xo).max()
    assert res.value == pytest.approx(ro.value, rel=1e-12)
    Xo = X.copy(); ro = C.c_double(0); o.sweep_maf("pcr_maf", Xo, B, ro, mesh)
    Xe = X.copy()
    lib.emu_pcr_maf(ou.ptr(Xe), ou.ptr(M), ou.ptr(B), o._sz, o._idx, C.c_int(pn), C.c_int(-1), *mp, R(1.3), C.c_int(0),
                    C.byref(res))
    assert np.array_equal(Xe, Xo), np.abs(Xe - Xo).max()
    assert res.value == pytest.approx(ro.value, rel=1e-12)
