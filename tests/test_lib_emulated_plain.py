"""CPU run of the library's uniform-grid per-call ABI (abi.cu over kernels.cu) in the CPU-emulated build of
tests/tools/emu, against the oracle: the same call with the same arguments on both libraries must leave the same
arrays (REAL fields bit for bit, double sums to 1e-12).  The GPU parity tests (tests/test_gpu_kernels.py) have shown
this on hardware; here it guards the host side (argument order, staging, flop formulas) and the plain-CUDA kernels on
every CPU test run.  pcr*_ (pcr.cu) and the fused kernels are device-only (TMA / tensor memory / inline PTX) and are
not part of the emulated build."""
import ctypes as C

import numpy as np
import pytest

import oracle_util as ou
from test_maf_kernels_emulated import _emu


def _both(real, sz=(11, 9, 14)):
    o = ou.OracleCZ(list(sz), real)
    return _emu(real), o


@pytest.mark.parametrize("real", ["f32", "f64"])
def test_emulated_sweeps_and_operator(real):
    lib, o = _both(real)
    X, B = ou.random_fields(o.shape, o.dt)
    cf = np.array([1, 1, 1, 1, 1, 1, 6], dtype=o.dt)
    for omg in (0.8, 1.5):
        om = o.creal(omg)
        out = []
        for dll in (o.dll, lib):
            x = X.copy(); w = np.zeros_like(X); res = C.c_double(0.5); flop = C.c_double(0)
            dll.jacobi_(ou.ptr(x), o._sz, o._idx, o.g(), ou.ptr(cf), C.byref(om), ou.ptr(B), C.byref(res), ou.ptr(w), C.byref(flop))
            out.append((x, res.value, flop.value))
        assert np.array_equal(out[0][0], out[1][0]) and out[0][2] == out[1][2]
        assert out[0][1] == pytest.approx(out[1][1], rel=1e-12)
        out = []
        for dll in (o.dll, lib):
            x = X.copy(); res = C.c_double(0); flop = C.c_double(0); ip = C.c_int(0)
            for color in (0, 1):
                cc = C.c_int(color)
                dll.psor2sma_core_(ou.ptr(x), o._sz, o._idx, o.g(), ou.ptr(cf), C.byref(ip), C.byref(cc), C.byref(om), ou.ptr(B),
                                   C.byref(res), C.byref(flop))
            x2 = X.copy(); r2 = C.c_double(0)
            dll.psor_(ou.ptr(x2), o._sz, o._idx, o.g(), ou.ptr(cf), C.byref(om), ou.ptr(B), C.byref(r2), C.byref(flop))
            out.append((x, x2, res.value, r2.value, flop.value))
        assert np.array_equal(out[0][0], out[1][0]) and np.array_equal(out[0][1], out[1][1]) and out[0][4] == out[1][4]
        assert out[0][2] == pytest.approx(out[1][2], rel=1e-12) and out[0][3] == pytest.approx(out[1][3], rel=1e-12)
    out = []
    for dll in (o.dll, lib):
        a = np.zeros_like(X); r = np.zeros_like(X); flop = C.c_double(0)
        dll.blas_calc_ax_(ou.ptr(a), ou.ptr(X), o._sz, o._idx, o.g(), ou.ptr(cf), C.byref(flop))
        dll.blas_calc_rk_(ou.ptr(r), ou.ptr(X), ou.ptr(B), o._sz, o._idx, o.g(), ou.ptr(cf), C.byref(flop))
        out.append((a, r, flop.value))
    assert np.array_equal(out[0][0], out[1][0]) and np.array_equal(out[0][1], out[1][1]) and out[0][2] == out[1][2]


@pytest.mark.parametrize("real", ["f32", "f64"])
def test_emulated_blas1_mask_boundary_and_error(real):
    lib, o = _both(real)
    X, Y = ou.random_fields(o.shape, o.dt)
    Z, = ou.random_fields(o.shape, o.dt, n=1, seed=11)
    a, b = o.creal(0.37), o.creal(-1.21)
    E0 = np.zeros_like(X); o.dll.exact_t_(o._sz, o.g(), ou.ptr(E0), o._dh, o._org)   # same err_t input for both
    out = []
    for dll in (o.dll, lib):
        flop = C.c_double(0)
        z1 = Z.copy(); dll.blas_triad_(ou.ptr(z1), ou.ptr(X), ou.ptr(Y), C.byref(a), o._sz, o._idx, o.g(), C.byref(flop))
        p1 = X.copy(); dll.blas_bicg_1_(ou.ptr(p1), ou.ptr(Y), ou.ptr(Z), C.byref(a), C.byref(b), o._sz, o._idx, o.g(), C.byref(flop))
        z2 = Z.copy(); dll.blas_bicg_2_(ou.ptr(z2), ou.ptr(X), ou.ptr(Y), C.byref(a), C.byref(b), o._sz, o._idx, o.g(), C.byref(flop))
        d1, d2 = o.creal(0), o.creal(0)
        dll.blas_dot1_(C.byref(d1), ou.ptr(X), o._sz, o._idx, o.g(), C.byref(flop))
        dll.blas_dot2_(C.byref(d2), ou.ptr(X), ou.ptr(Y), o._sz, o._idx, o.g(), C.byref(flop))
        m = Z.copy(); dll.imask_k_(ou.ptr(m), o._sz, o._idx, o.g())
        c1 = Z.copy(); dll.blas_copy_(ou.ptr(c1), ou.ptr(X), o._sz, o.g())
        c2 = Z.copy(); dll.blas_copy_in_(ou.ptr(c2), ou.ptr(X), o._sz, o.g())
        c3 = Z.copy(); dll.blas_clear_(ou.ptr(c3), o._sz, o.g())
        p = Z.copy(); dll.bc_k_(o._sz, o.g(), ou.ptr(p), o._dh, o._org, o._nID)
        e = np.zeros_like(X); dll.exact_t_(o._sz, o.g(), ou.ptr(e), o._dh, o._org)
        err = C.c_double(0); loc = (C.c_int * 3)()
        e2 = E0.copy(); dll.err_t_(o._sz, o._idx, o.g(), C.byref(err), ou.ptr(X), ou.ptr(e2), loc)
        out.append(dict(z1=z1, p1=p1, z2=z2, m=m, c1=c1, c2=c2, c3=c3, p=p, e=e, e2=e2, d1=d1.value, d2=d2.value,
                        err=err.value, loc=tuple(loc), flop=flop.value))
    ref, emu = out
    for k in ("z1", "p1", "z2", "m", "c1", "c2", "c3", "e2"):
        assert np.array_equal(ref[k], emu[k]), k
    # bc_k / exact_t evaluate sin / sinh: host libm table (bc_k) resp. device functions (exact_t) -> a few ulp
    assert np.array_equal(ref["p"], emu["p"])
    np.testing.assert_allclose(emu["e"], ref["e"], rtol=0, atol=4 * np.finfo(o.dt).eps * max(1.0, float(np.abs(ref["e"]).max())))
    rel = 1e-6 if real == "f32" else 1e-13
    assert emu["d1"] == pytest.approx(ref["d1"], rel=rel) and emu["d2"] == pytest.approx(ref["d2"], rel=rel, abs=rel)
    assert emu["err"] == ref["err"] and emu["loc"] == ref["loc"] and emu["flop"] == ref["flop"]


@pytest.mark.parametrize("real", ["f32", "f64"])
@pytest.mark.parametrize("sz", [[9, 9, 9], [10, 8, 34], [6, 6, 131]])
def test_emulated_line_sor_per_call_abi(real, sz):
    """pcr_rb_, pcr_rb_esa_, pcr_, pcr_eda_, pcr_esa_, pcr_j_esa_ (abi.cu -> k_pcr_tables + k_pcr_lines, the
    block-per-line path) against the oracle, bit for bit."""
    lib = _emu(real)
    o = ou.OracleCZ(sz, real)
    X, B = ou.random_fields(o.shape, o.dt)
    M = o.mask()
    n = o.idx[5] - o.idx[4] + 1
    pn = C.c_int(ou.get_num_stage(n)); s = C.c_int(2 ** (pn.value - 2)); ip = C.c_int(0)
    om = o.creal(1.3)
    w = [ou.ptr(a) for a in o.work1d]
    for kind in ("pcr_rb", "pcr_rb_esa", "pcr", "pcr_eda", "pcr_esa", "pcr_j_esa"):
        if pn.value < 3 and kind in ("pcr_rb_esa", "pcr", "pcr_esa"):
            continue
        out = []
        for dll in (o.dll, lib):
            x = X.copy(); res = C.c_double(0.5); flop = C.c_double(0)
            head = (o._sz, o._idx, o.g(), C.byref(pn))
            tail = (C.byref(om), C.byref(res), C.byref(flop))
            f3 = (ou.ptr(x), ou.ptr(M), ou.ptr(B))
            if kind == "pcr_rb":
                for color in (0, 1):
                    cc = C.c_int(color)
                    dll.pcr_rb_(*head, C.byref(ip), C.byref(cc), *f3, *w, *tail)
            elif kind == "pcr_rb_esa":
                for color in (0, 1):
                    cc = C.c_int(color)
                    dll.pcr_rb_esa_(*head, C.byref(ip), C.byref(cc), C.byref(s), *f3, *w, *tail)
            elif kind == "pcr":
                dll.pcr_(*head, *f3, *w, *tail)
            elif kind == "pcr_eda":
                dll.pcr_eda_(*head, *f3, *w[:3], *tail)
            elif kind == "pcr_esa":
                dll.pcr_esa_(*head, C.byref(s), *f3, *w, *tail)
            else:
                src = np.zeros_like(X); wrk = np.zeros_like(X)
                dll.pcr_j_esa_(*head, C.byref(s), *f3, *w, ou.ptr(src), ou.ptr(wrk), *tail)
            out.append((x, res.value, flop.value))
        assert np.array_equal(out[0][0], out[1][0]), (kind, np.abs(out[0][0] - out[1][0]).max())
        assert out[0][1] == pytest.approx(out[1][1], rel=1e-12), kind
        assert out[0][2] == out[1][2], kind


def test_emulated_line_sor_solves_match_golden():
    """czb_solve with the uniform-grid line-SOR names (LSOR_PCR_RB / LSOR_PCR / LSOR_PCR_J_ESA loops of solver.cu) and
    PBiCGstab with a line-SOR preconditioner against the reference driver's golden histories."""
    from test_maf_kernels_emulated import GOLDEN, _solve
    lib = _emu("f32")
    for solver in ("pcr_rb", "pcr_rb_esa", "pcr_j_esa", "pcr", "pcr_eda", "pcr_esa"):
        g = [c for c in GOLDEN if c["real"] == "f32" and c["n"] == 32 and c["solver"] == solver][0]
        o, itr, res, hist, P, flop = _solve(lib, "f32", 32, solver, None, g["itr_max"], float(g["coef"]))
        assert itr == g["itr"], solver
        np.testing.assert_allclose(hist, [float(h) for h in g["history"]], rtol=2e-6, err_msg=solver)
        Po, RHSo = o.setup_laplace()
        itr_o, res_o, hist_o = o.stationary(solver, Po, RHSo, g["itr_max"], float(g["coef"]))
        assert np.array_equal(P, Po), solver
    g = [c for c in GOLDEN if c["real"] == "f32" and c["n"] == 32 and c["solver"] == "pbicgstab" and c["precond"] == "pcr_rb"][0]
    o, itr, res, hist, P, flop = _solve(lib, "f32", 32, "pbicgstab", "pcr_rb", g["itr_max"], float(g["coef"]))
    assert abs(itr - g["itr"]) <= 1
    m = min(len(hist), len(g["history"]))
    np.testing.assert_allclose(hist[:m], [float(h) for h in g["history"]][:m], rtol=1e-4, atol=1e-12)


@pytest.mark.parametrize("real", ["f32", "f64"])
def test_emulated_library_reproduces_the_reference_fortran_vectors(real):
    """The library's own per-call ABI (CPU-emulated build) against the committed outputs of the reference's Fortran
    executed through oracle/f90_to_c.py: every array argument bit for bit on non-unit coefficients and a stretched mesh
    (the -m gpu twin is tests/test_gpu_f90c_vectors.py)."""
    import json
    import os
    import sys
    import f90c_harness as H
    here = os.path.dirname(os.path.abspath(__file__))
    sys.path.insert(0, os.path.join(here, "golden"))
    import make_f90c_golden as MG
    digests = json.load(open(os.path.join(here, "golden", "f90c_digests.json")))
    sigs = {s["name"]: s for s in json.load(open(H.SIG_JSON))}
    n = H.check_library_against_fortran_vectors(_emu(real), real, MG.SIZES[:4], digests, sigs, MG.run_routine, MG.record)
    assert n >= 130
