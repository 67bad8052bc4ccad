"""CPU tests of the ORACLE (test infrastructure): pins the C++ restatement of the reference's
Fortran kernels against every known answer the reference offers and against
the committed golden histories produced by the reference's unmodified C++ driver."""
import ctypes as C
import json
import os

import numpy as np
import pytest

import oracle_util as ou

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = json.load(open(os.path.join(HERE, "golden", "histories.json")))


def _golden(real, n, solver, precond=None, coef=None):
    for c in GOLDEN:
        if (c["real"], c["n"], c["solver"], c["precond"]) == (real, n, solver, precond) and (coef is None or c["coef"] == coef):
            return c
    raise KeyError((real, n, solver, precond))


def test_thomas_3x3_known_answer():
    # reference example/tri/test1.cpp:25-31 : b=[2,4,-1] c=[3,-3] a=[4,3] d=[8,3,3] -> (1,2,3)
    d = ou.load("f64")
    n = C.c_int(3)
    a = np.array([0.0, 4.0, 3.0]); b = np.array([2.0, 4.0, -1.0]); c = np.array([3.0, -3.0, 0.0])
    rhs = np.array([8.0, 3.0, 3.0]); w = np.zeros(3)
    d.tdma_0_(C.byref(n), ou.ptr(rhs), ou.ptr(a), ou.ptr(b), ou.ptr(c), ou.ptr(w))
    np.testing.assert_allclose(rhs, [1.0, 2.0, 3.0], rtol=0, atol=1e-14)


@pytest.mark.parametrize("real,tol", [("f32", 2e-5), ("f64", 1e-12)])
@pytest.mark.parametrize("mode", [0, 1])
def test_pcr_line_linear_known_answer(real, tol, mode):
    # reference example/tri/test3.cpp:37-50, test5.cpp:146-177: 1-D Laplace, N=23, Dirichlet
    # -3 / 9 -> x_i = -3 + 0.5 i ; solved with the PCR line solver (2x2 and 4x4 last stage)
    # and cross-checked with Thomas tdma_p (obsolete.f90:65-84).
    d = ou.load(real); dt = np.float32 if real == "f32" else np.float64
    n = 23
    a = np.full(n, -0.5, dt); c = np.full(n, -0.5, dt); a[0] = 0; c[-1] = 0
    rhs = np.zeros(n, dt); rhs[0] = 0.5 * -3.0; rhs[-1] = 0.5 * 9.0
    x = np.zeros(n, dt)
    nn, mm = C.c_int(n), C.c_int(mode)
    d.oracle_pcr_line_(C.byref(nn), C.byref(mm), ou.ptr(a), ou.ptr(c), ou.ptr(rhs), ou.ptr(x))
    exact = -3.0 + 0.5 * np.arange(1, n + 1)
    np.testing.assert_allclose(x, exact, rtol=0, atol=tol * 10)
    t = rhs.copy(); w = np.zeros(n, dt)
    d.tdma_p_(C.byref(nn), ou.ptr(t), ou.ptr(a), ou.ptr(c), ou.ptr(w))
    np.testing.assert_allclose(x, t, rtol=0, atol=tol * 10)


def test_analytic_solution_error_is_discretisation_level():
    # cz_Evaluate.cpp:550-563 "Error max" of converged runs: O(h^2), far below the field's O(1)
    for c in GOLDEN:
        if c["solver"] in ("sor2sma", "pbicgstab", "pcr_rb") and c["n"] == 32:
            assert float(c["errmax"]) < 1.0e-3, c


def test_memo_pcr_eda_same_iteration_count_as_pcr():
    # reference doc/Memo.md:131-134
    assert _golden("f32", 32, "pcr")["itr"] == _golden("f32", 32, "pcr_eda")["itr"]
    assert _golden("f32", 32, "pcr")["history"] == _golden("f32", 32, "pcr_esa")["history"]


CASES = [("f32", 32, "jacobi", None), ("f32", 32, "sor2sma", None), ("f32", 32, "psor", None),
         ("f32", 32, "pcr_rb", None), ("f32", 32, "pcr_rb_esa", None), ("f32", 32, "pcr_j_esa", None),
         ("f32", 32, "pcr", None), ("f32", 32, "pcr_eda", None),
         ("f32", 32, "pbicgstab", "sor2sma"), ("f32", 32, "pbicgstab", "jacobi"), ("f32", 32, "pbicgstab", "none"),
         ("f32", 32, "pbicgstab", "pcr_rb"),
         ("f64", 32, "jacobi", None), ("f64", 32, "sor2sma", None), ("f64", 32, "pcr_rb", None),
         ("f64", 32, "pbicgstab", "sor2sma")]


@pytest.mark.parametrize("real,n,solver,precond", CASES)
def test_python_loops_reproduce_reference_driver_history(real, n, solver, precond):
    """tests/oracle_util.OracleCZ (Python restatement of cz_Poisson.cpp over the oracle kernels)
    must print the same history as the reference's own driver did (tests/golden)."""
    g = _golden(real, n, solver, precond)
    o = ou.OracleCZ([n] * 3, real)
    P, RHS = o.setup_laplace()
    if solver == "pbicgstab":
        itr, res, hist = o.pbicgstab(P, RHS, g["itr_max"], float(g["coef"]), precond)
    else:
        itr, res, hist = o.stationary(solver, P, RHS, g["itr_max"], float(g["coef"]))
    assert itr == g["itr"]
    assert ["%13.6e" % h for h in hist] == ["%13s" % h for h in g["history"]]


def test_double_and_real_sum_modes_agree_to_float_rounding():
    # ORACLE_SUM=real bounds the reference's own drift; here only the default
    # (double accumulation) is exercised and must be independent of the thread count.
    o = ou.OracleCZ([32] * 3, "f32")
    X, B = ou.random_fields(o.shape, o.dt)
    o.ac1 = o.dt(1.1)
    r1 = C.c_double(0); Xa = X.copy(); o.sweep_rb(Xa, B, r1)
    os.environ["OMP_NUM_THREADS"] = "3"
    r2 = C.c_double(0); Xb = X.copy(); o.sweep_rb(Xb, B, r2)
    assert np.array_equal(Xa, Xb) and r1.value == r2.value


# ---- the C++ port against an independent numpy restatement of the same Fortran (oracle/np_restatement.py) ----
import sys
sys.path.insert(0, os.path.join(os.path.dirname(HERE), "oracle"))
import np_restatement as npr  # noqa: E402


def _np_case(real, sz):
    o = ou.OracleCZ(sz, real)
    X, B = ou.random_fields(o.shape, o.dt)
    return o, X, B


@pytest.mark.parametrize("real", ["f32", "f64"])
@pytest.mark.parametrize("sz", [[9, 9, 9], [12, 7, 23], [16, 16, 37]])
def test_port_equals_numpy_restatement_sweeps(real, sz):
    o, X, B = _np_case(real, sz)
    cf = [1, 1, 1, 1, 1, 1, 6]
    # Jacobi (cz_solver.f90:284-387)
    o.ac1 = o.dt(0.8)
    Xo = X.copy(); ro = C.c_double(0); o.sweep_jacobi(Xo, B, ro)
    Xn = X.copy(); rn = npr.jacobi(Xn, B, o.idx, cf, 0.8)
    assert np.array_equal(Xo, Xn)
    assert ro.value == pytest.approx(rn, rel=1e-12)
    # red-black SOR, colour 0 then colour 1 (cz_solver.f90:404-493, cz_Poisson.cpp:197-203)
    o.ac1 = o.dt(1.5)
    Xo = X.copy(); ro = C.c_double(0); o.sweep_rb(Xo, B, ro)
    Xn = X.copy(); rn = npr.psor2sma_core(Xn, B, o.idx, cf, 0, 0, 1.5) + npr.psor2sma_core(Xn, B, o.idx, cf, 0, 1, 1.5)
    assert np.array_equal(Xo, Xn)
    assert ro.value == pytest.approx(rn, rel=1e-12)
    # operator and residual vector (cz_blas.f90:579-644, 658-723)
    cfa = np.array(cf, dtype=o.dt); flop = C.c_double(0)
    Ao = np.zeros_like(X); o.dll.blas_calc_ax_(ou.ptr(Ao), ou.ptr(X), o._sz, o._idx, o.g(), ou.ptr(cfa), C.byref(flop))
    An = np.zeros_like(X); npr.blas_calc_ax(An, X, o.idx, cf)
    assert np.array_equal(Ao, An)
    Ro = np.zeros_like(X); o.dll.blas_calc_rk_(ou.ptr(Ro), ou.ptr(X), ou.ptr(B), o._sz, o._idx, o.g(), ou.ptr(cfa), C.byref(flop))
    Rn = np.zeros_like(X); npr.blas_calc_rk(Rn, X, B, o.idx, cf)
    assert np.array_equal(Ro, Rn)


@pytest.mark.parametrize("real", ["f32", "f64"])
@pytest.mark.parametrize("sz", [[9, 9, 9], [12, 7, 23], [10, 8, 34], [8, 6, 66], [6, 6, 131]])
def test_port_equals_numpy_restatement_line_sor(real, sz):
    """pcr_rb (cz_solver.f90:497-662) incl. line lengths of 2^m and 2^m +- 1 cells; the numpy version carries the
    coefficient recurrences a, c, e along with d exactly as the Fortran does (the CUDA kernels tabulate them)."""
    o, X, B = _np_case(real, sz)
    o.ac1 = o.dt(1.3)
    M = o.mask()
    n = o.idx[5] - o.idx[4] + 1
    pn = ou.get_num_stage(n)
    Xo = X.copy(); ro = C.c_double(0); o.sweep_pcr("pcr_rb", Xo, B, ro)
    Xn = X.copy(); rn = npr.pcr_rb(Xn, M, B, o.idx, pn, 0, 1.3) + npr.pcr_rb(Xn, M, B, o.idx, pn, 1, 1.3)
    assert np.array_equal(Xo, Xn), np.abs(Xo - Xn).max()
    assert ro.value == pytest.approx(rn, rel=1e-12)
    # pcr_rb_esa (cz_solver.f90:1261-1469): 4x4 systems by Cramer's rule in the last stage
    if pn >= 3:
        Xo = X.copy(); ro = C.c_double(0); o.sweep_pcr("pcr_rb_esa", Xo, B, ro)
        Xn = X.copy(); rn = npr.pcr_rb_esa(Xn, M, B, o.idx, pn, 0, 1.3) + npr.pcr_rb_esa(Xn, M, B, o.idx, pn, 1, 1.3)
        assert np.array_equal(Xo, Xn), np.abs(Xo - Xn).max()
        assert ro.value == pytest.approx(rn, rel=1e-12)


@pytest.mark.parametrize("real", ["f32", "f64"])
@pytest.mark.parametrize("sz", [[9, 9, 9], [7, 6, 23], [6, 5, 34]])
@pytest.mark.parametrize("kind", ["pcr_j_esa", "pcr", "pcr_eda", "pcr_esa"])
def test_port_equals_numpy_restatement_other_line_sor(real, sz, kind):
    """Line Jacobi and the lexicographic (Gauss-Seidel between lines) variants, cz_solver.f90:666-1257, 1473-1676."""
    o, X, B = _np_case(real, sz)
    o.ac1 = o.dt(0.9 if kind == "pcr_j_esa" else 1.3)
    M = o.mask()
    pn = ou.get_num_stage(o.idx[5] - o.idx[4] + 1)
    Xo = X.copy(); ro = C.c_double(0); o.sweep_pcr(kind, Xo, B, ro)
    Xn = X.copy(); rn = getattr(npr, kind)(Xn, M, B, o.idx, pn, float(o.ac1))
    assert np.array_equal(Xo, Xn), np.abs(Xo - Xn).max()
    assert ro.value == pytest.approx(rn, rel=1e-12)


@pytest.mark.parametrize("real", ["f32", "f64"])
def test_port_equals_numpy_restatement_psor_and_blas1(real):
    o, X, B = _np_case(real, [9, 8, 11])
    cf = [1, 1, 1, 1, 1, 1, 6]
    o.ac1 = o.dt(1.5)
    Xo = X.copy(); ro = C.c_double(0); o.sweep_psor(Xo, B, ro)
    Xn = X.copy(); rn = npr.psor(Xn, B, o.idx, cf, 1.5)
    assert np.array_equal(Xo, Xn)
    assert ro.value == pytest.approx(rn, rel=1e-12)
    Y, Z = ou.random_fields(o.shape, o.dt, seed=7)
    flop = C.c_double(0)
    a, b = o.creal(0.37), o.creal(-1.21)
    Zo = Z.copy(); o.dll.blas_triad_(ou.ptr(Zo), ou.ptr(X), ou.ptr(Y), C.byref(a), o._sz, o._idx, o.g(), C.byref(flop))
    Zn = Z.copy(); npr.blas_triad(Zn, X, Y, a.value, o.idx)
    assert np.array_equal(Zo, Zn)
    Po = X.copy(); o.dll.blas_bicg_1_(ou.ptr(Po), ou.ptr(Y), ou.ptr(Z), C.byref(a), C.byref(b), o._sz, o._idx, o.g(), C.byref(flop))
    Pn = X.copy(); npr.blas_bicg_1(Pn, Y, Z, a.value, b.value, o.idx)
    assert np.array_equal(Po, Pn)
    Zo = Z.copy(); o.dll.blas_bicg_2_(ou.ptr(Zo), ou.ptr(X), ou.ptr(Y), C.byref(a), C.byref(b), o._sz, o._idx, o.g(), C.byref(flop))
    Zn = Z.copy(); npr.blas_bicg_2(Zn, X, Y, a.value, b.value, o.idx)
    assert np.array_equal(Zo, Zn)


# ---- stretched-grid (MAF) twins, SURVEY.md §8f-2: cz_maf.f90, cz_blas.f90:738-1039 ----------------------------
MAF_CASES = [("f32", 32, "jacobi_maf", None), ("f32", 32, "sor2sma_maf", None), ("f32", 32, "psor_maf", None),
             ("f32", 32, "pcr_rb_maf", None), ("f32", 32, "pcr_rb_esa_maf", None), ("f32", 32, "pcr_maf", None),
             ("f32", 32, "pcr_eda_maf", None), ("f32", 32, "pcr_esa_maf", None),
             ("f32", 32, "pbicgstab_maf", "sor2sma_maf"), ("f32", 32, "pbicgstab_maf", "pcr_rb_maf"),
             ("f32", 32, "pbicgstab_maf", "none"),
             ("f64", 32, "sor2sma_maf", None), ("f64", 32, "pcr_rb_maf", None), ("f64", 32, "pbicgstab_maf", "sor2sma_maf")]


@pytest.mark.parametrize("real,n,solver,precond", MAF_CASES)
def test_python_loops_reproduce_reference_driver_history_maf(real, n, solver, precond):
    g = _golden(real, n, solver, precond)
    o = ou.OracleCZ([n] * 3, real)
    P, RHS = o.setup_laplace()
    if solver == "pbicgstab_maf":
        itr, res, hist = o.pbicgstab(P, RHS, g["itr_max"], float(g["coef"]), precond, maf=True)
    else:
        itr, res, hist = o.stationary(solver, P, RHS, g["itr_max"], float(g["coef"]))
    assert itr == g["itr"]
    assert ["%13.6e" % h for h in hist] == ["%13s" % h for h in g["history"]]


def test_maf_twins_on_the_uniform_mesh_reproduce_the_plain_solvers():
    """What pins the MAF port: on the driver's uniform mesh the metric coefficients are 1/h^2 up to rounding, so the
    reference's own runs of X_maf must track X (golden histories written by the unmodified reference driver)."""
    for real, solver, rel in (("f32", "jacobi", 1e-3), ("f32", "sor2sma", 1e-3), ("f32", "psor", 1e-3), ("f64", "sor2sma", 1e-9)):
        a, b = _golden(real, 32, solver), _golden(real, 32, solver + "_maf")
        assert a["itr"] == b["itr"]
        np.testing.assert_allclose([float(h) for h in b["history"]], [float(h) for h in a["history"]], rtol=rel)
    # the line routines too (round 1 had their residual sqrt(2) too large: `tmp(k) = tmp(k) + dp*dp` and
    # `res1 = res1 + dp*dp` are the two branches of #ifdef _SVR, cz_maf.f90:642-646, not two statements)
    a, b = _golden("f64", 32, "pcr_rb"), _golden("f64", 32, "pcr_rb_maf")
    assert a["itr"] == b["itr"]
    np.testing.assert_allclose([float(h) for h in b["history"]], [float(h) for h in a["history"]], rtol=2e-4)
    assert _golden("f32", 32, "pcr_maf")["history"] == _golden("f32", 32, "pcr_eda_maf")["history"] == _golden("f32", 32, "pcr_esa_maf")["history"]
    assert _golden("f32", 32, "pcr_rb_maf")["history"] == _golden("f32", 32, "pcr_rb_esa_maf")["history"]
    for c in GOLDEN:
        if c["solver"].endswith("_maf") and c["solver"] != "jacobi_maf" and c["precond"] != "none":
            assert float(c["errmax"]) < 1.0e-3, c


@pytest.mark.parametrize("real", ["f32", "f64"])
@pytest.mark.parametrize("sz", [[9, 9, 9], [12, 7, 23], [10, 8, 34]])
def test_port_equals_numpy_restatement_maf(real, sz):
    """The C++ port of the MAF routines against the independent numpy restatement, bit for bit, on a STRETCHED mesh
    (so that the first- and second-derivative metric terms are all non-trivial) and random fields."""
    o, X, B = _np_case(real, sz)
    mesh = o.mesh(stretch=0.3)
    xc, yc, zc = mesh
    M = o.mask()
    # point sweeps
    o.ac1 = o.dt(0.8)
    Xo = X.copy(); ro = C.c_double(0); o.sweep_maf("jacobi_maf", Xo, B, ro, mesh)
    Xn = X.copy(); rn = npr.jacobi_maf(Xn, B, o.idx, xc, yc, zc, 0.8)
    assert np.array_equal(Xo, Xn)
    assert ro.value == pytest.approx(rn, rel=1e-12)
    o.ac1 = o.dt(1.5)
    Xo = X.copy(); ro = C.c_double(0); o.sweep_maf("sor2sma_maf", Xo, B, ro, mesh)
    Xn = X.copy(); rn = npr.psor2sma_core_maf(Xn, B, o.idx, xc, yc, zc, 0, 0, 1.5) + npr.psor2sma_core_maf(Xn, B, o.idx, xc, yc, zc, 0, 1, 1.5)
    assert np.array_equal(Xo, Xn)
    assert ro.value == pytest.approx(rn, rel=1e-12)
    # pivot, operator, residual vector
    Po = o.pivot(mesh)
    Pn = np.zeros_like(X); npr.search_pivot(Pn, o.idx, xc, yc, zc)
    assert np.array_equal(Po, Pn)
    flop = C.c_double(0); mp = [ou.ptr(m) for m in mesh]
    Ao = np.zeros_like(X); o.dll.calc_ax_maf_(ou.ptr(Ao), ou.ptr(X), o._sz, o._idx, o.g(), *mp, ou.ptr(Po), C.byref(flop))
    An = np.zeros_like(X); npr.calc_ax_maf(An, X, o.idx, xc, yc, zc, Pn)
    assert np.array_equal(Ao, An)
    Ro = np.zeros_like(X); o.dll.calc_rk_maf_(ou.ptr(Ro), ou.ptr(X), ou.ptr(B), o._sz, o._idx, o.g(), *mp, ou.ptr(Po), C.byref(flop))
    Rn = np.zeros_like(X); npr.calc_rk_maf(Rn, X, B, o.idx, xc, yc, zc, Pn)
    assert np.array_equal(Ro, Rn)
    assert flop.value == 2 * 63.0 * (sz[0] - 2) * (sz[1] - 2) * (sz[2] - 2)
    # line SOR: red-black (both argument lists) and lexicographic (three argument lists)
    pn = ou.get_num_stage(o.idx[5] - o.idx[4] + 1)
    o.ac1 = o.dt(1.3)
    Xn = X.copy(); rn = npr.pcr_rb_maf(Xn, M, B, o.idx, pn, 0, 1.3, xc, yc, zc) + npr.pcr_rb_maf(Xn, M, B, o.idx, pn, 1, 1.3, xc, yc, zc)
    for kind in ("pcr_rb_maf", "pcr_rb_esa_maf"):
        Xo = X.copy(); ro = C.c_double(0); o.sweep_maf(kind, Xo, B, ro, mesh)
        assert np.array_equal(Xo, Xn), (kind, np.abs(Xo - Xn).max())
        assert ro.value == pytest.approx(rn, rel=1e-12)
    if sz[0] * sz[1] <= 100:
        Xn = X.copy(); rn = npr.pcr_maf(Xn, M, B, o.idx, pn, 1.3, xc, yc, zc)
        for kind in ("pcr_maf", "pcr_eda_maf", "pcr_esa_maf"):
            Xo = X.copy(); ro = C.c_double(0); o.sweep_maf(kind, Xo, B, ro, mesh)
            assert np.array_equal(Xo, Xn), (kind, np.abs(Xo - Xn).max())
            assert ro.value == pytest.approx(rn, rel=1e-12)


@pytest.mark.parametrize("real", ["f32", "f64"])
def test_port_equals_numpy_restatement_psor_maf(real):
    o, X, B = _np_case(real, [8, 7, 10])
    mesh = o.mesh(stretch=0.3)
    o.ac1 = o.dt(1.5)
    Xo = X.copy(); ro = C.c_double(0); o.sweep_maf("psor_maf", Xo, B, ro, mesh)
    Xn = X.copy(); rn = npr.psor_maf(Xn, B, o.idx, *mesh, 1.5)
    assert np.array_equal(Xo, Xn)
    assert ro.value == pytest.approx(rn, rel=1e-12)


# ---- the oracle's PCR stage recurrences against OUTPUTS OF THE REFERENCE ITSELF ------------------------------------------
# src/cz_cpp/cz_pcr.cpp (CZ::pcr / CZ::pcr_kernel, :42-96) is the reference's own C++ statement of the recurrences the
# Fortran line solvers run (cz_solver.f90:584-608).  oracle/Makefile compiles it unmodified behind an exact-I/O harness
# (oracle/ref_pcr_harness.cpp -> oracle/_ref/ref_pcr_*); tests/golden/ref_pcr_stages.json holds its outputs
# (tests/golden/make_ref_pcr.py).  Double: the port must agree bit for bit.  Float: the C++ evaluates
# 1.0 / (1.0 - ap*c - cp*a) with double literals where the Fortran stays in REAL -> a few ulp.
def _port_stages(real, n, stages, a, c, d):
    dll = ou.load(real)
    oa, oc, od = a.copy(), c.copy(), d.copy()
    dll.oracle_pcr_stages_(C.byref(C.c_int(n)), C.byref(C.c_int(stages)), ou.ptr(oa), ou.ptr(oc), ou.ptr(od))
    return oa, oc, od


def _check_stages(real, got, ref):
    for g, r in zip(got, ref):
        if real == "f64":
            assert np.array_equal(g, r), np.abs(g - r).max()
        else:
            np.testing.assert_allclose(g.astype(np.float64), r, rtol=3e-6, atol=1e-30)   # <= 1 ulp of the denominator per stage, up to 7 stages


def test_pcr_stages_equal_the_reference_cpp_pcr_golden():
    gold = json.load(open(os.path.join(HERE, "golden", "ref_pcr_stages.json")))
    assert len(gold["cases"]) >= 40
    for cs in gold["cases"]:
        dt = np.float64 if cs["real"] == "f64" else np.float32
        a, c, d = (np.array([float.fromhex(v) for v in cs[k]], dtype=dt) for k in ("a", "c", "d"))
        ref = [np.array([float.fromhex(v) for v in cs[k]], dtype=np.float64) for k in ("ref_a", "ref_c", "ref_d")]
        _check_stages(cs["real"], _port_stages(cs["real"], cs["n"], cs["stages"], a, c, d), ref)


@pytest.mark.parametrize("real", ["f64", "f32"])
def test_pcr_stages_equal_the_reference_cpp_pcr_live(real):
    """The same against the harness run here (skipped where /root/reference was absent at build time), on the line lengths
    of the BASELINE configs (62, 254, 510, 1022 cells) with the solver's own stage counts pn-1 and pn-2."""
    exe = ou.ref_binary("ref_pcr_%s" % real)
    if exe is None:
        pytest.skip("oracle/_ref/ref_pcr_%s not built (needs /root/reference at build time)" % real)
    import subprocess
    dt = np.float64 if real == "f64" else np.float32
    rng = np.random.default_rng(11)
    for n in (62, 254, 510, 1022):
        pn = int(np.floor(np.log2(n))) + 1
        r = dt(1.0) / dt(6.0)
        a = np.full(n, -r, dt); c = np.full(n, -r, dt); a[0] = 0; c[-1] = 0
        d = rng.uniform(-1, 1, n).astype(dt)
        for stages in (pn - 1, pn - 2):
            inp = "%d %d\n" % (n, stages) + "\n".join(" ".join(float(v).hex() for v in arr) for arr in (a, c, d)) + "\n"
            p = subprocess.run([exe], input=inp, capture_output=True, text=True, timeout=60)
            assert p.returncode == 0, p.stderr
            ref = [np.array([float.fromhex(t) for t in line.split()]) for line in p.stdout.strip().split("\n")]
            _check_stages(real, _port_stages(real, n, stages, a, c, d), ref)
