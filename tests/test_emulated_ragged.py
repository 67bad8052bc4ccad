"""Ragged and tiny grids through the library's own solve loops (czb_solve of the CPU-emulated build, tests/tools/emu)
against the oracle: seeded random non-cubic sizes from 3 nodes per axis (ONE inner cell) upwards, every stationary solver
name incl. the stretched-grid twins, random relaxation factors and iteration counts, non-zero right-hand side and a
random interior start field -- same iteration count, bit-identical field, history to summation order; PBiCGstab with
every preconditioner on ragged grids to the tolerance of tests/test_gpu_solve.py.  The reference has no test of this kind
(SURVEY section 4); the sizes are the edge cases its driver admits (main.cpp:19-29 takes any gsz_x gsz_y gsz_z)."""
import ctypes as C

import numpy as np
import pytest

import oracle_util as ou
from test_maf_kernels_emulated import _emu

STATIONARY = ["jacobi", "psor", "sor2sma", "pcr_rb", "pcr_rb_esa", "pcr_j_esa", "pcr", "pcr_esa",
              "jacobi_maf", "psor_maf", "sor2sma_maf", "pcr_rb_maf", "pcr_rb_esa_maf", "pcr_maf", "pcr_esa_maf"]
PRECOND = ["none", "jacobi", "psor", "sor2sma", "pcr", "pcr_rb", "pcr_rb_esa"]


def _solve(lib, o, name, pre, X, B, iters, coef):
    ctx = lib.czb_ctx_create(o._sz, o._idx, o._nID, o.creal(o.pitch), o._org)
    hist = (C.c_double * (iters + 2))(); res = C.c_double(0); flop = C.c_double(0)
    itr = lib.czb_solve(ctx, name.encode(), pre.encode() if pre else None, ou.ptr(X), ou.ptr(B), iters, o.creal(coef),
                        C.byref(res), hist, iters + 2, C.byref(flop))
    lib.czb_ctx_destroy(ctx)
    return itr, np.array(hist[:iters + 2])


@pytest.mark.parametrize("lo,hi,klo,khi,seed,cases", [(3, 7, 4, 8, 7, 40), (4, 22, 4, 40, 1, 40)],
                         ids=["tiny", "ragged"])
def test_stationary_solvers_on_ragged_grids(lo, hi, klo, khi, seed, cases):
    # k >= 4 nodes: a line of one cell has no 4x4 last stage (the library refuses it with a message; the reference
    # reads beyond its arrays there)
    rng = np.random.default_rng(seed)
    for _ in range(cases):
        real = "f32" if rng.random() < 0.7 else "f64"
        sz = [int(rng.integers(lo, hi)), int(rng.integers(lo, hi)), int(rng.integers(klo, khi))]
        solver = STATIONARY[int(rng.integers(len(STATIONARY)))]
        coef = float(rng.choice([0.8, 1.0, 1.3])) if "jacobi" in solver else float(rng.choice([1.0, 1.5, 1.9]))
        iters = int(rng.integers(1, 9))
        lib, o = _emu(real), ou.OracleCZ(sz, real)
        P, RHS = o.setup_laplace()
        X, B = ou.random_fields(o.shape, o.dt, seed=int(rng.integers(1 << 30)))
        P[3:-3, 3:-3, 3:-3] = X[3:-3, 3:-3, 3:-3]
        RHS[3:-3, 3:-3, 3:-3] = 0.05 * B[3:-3, 3:-3, 3:-3]
        Pg, Rg = P.copy(), RHS.copy()
        itr, hist = _solve(lib, o, solver, None, Pg, Rg, iters, coef)
        itr_o, _, hist_o = o.stationary(solver, P, RHS, iters, coef)
        case = (real, sz, solver, coef, iters)
        assert itr == itr_o, case
        assert np.array_equal(Pg, P), (case, np.abs(Pg - P).max())
        np.testing.assert_allclose(hist[:len(hist_o)], hist_o, rtol=1e-11, err_msg=str(case))


def test_pbicgstab_with_every_preconditioner_on_ragged_grids():
    rng = np.random.default_rng(3)
    for _ in range(30):
        real = "f32" if rng.random() < 0.6 else "f64"
        sz = [int(rng.integers(4, 18)), int(rng.integers(4, 18)), int(rng.integers(5, 36))]
        maf = rng.random() < 0.4
        pc = PRECOND[int(rng.integers(len(PRECOND)))]
        pre = pc + ("_maf" if maf and pc != "none" else "")
        coef = 0.8 if "jacobi" in pc else float(rng.choice([1.0, 1.3]))
        iters = int(rng.integers(2, 6))
        lib, o = _emu(real), ou.OracleCZ(sz, real)
        P, RHS = o.setup_laplace()
        Pg, Rg = P.copy(), RHS.copy()
        itr, hist = _solve(lib, o, "pbicgstab_maf" if maf else "pbicgstab", pre, Pg, Rg, iters, coef)
        itr_o, _, hist_o = o.pbicgstab(P, RHS, iters, coef, pre, maf=maf)
        case = (real, sz, maf, pre, coef, iters)
        tol = 2e-3 if real == "f32" else 1e-9          # REAL-rounded Krylov scalars: last-bit differences grow per iteration
        assert itr == itr_o, case
        np.testing.assert_allclose(hist[:len(hist_o)], hist_o, rtol=tol, err_msg=str(case))
        np.testing.assert_allclose(Pg, P, rtol=0, atol=tol * np.abs(P).max(), err_msg=str(case))


def test_per_call_abi_on_index_windows_and_general_coefficients():
    """The reference ABI takes the inner-index window `idx` verbatim (a rank with a neighbour starts at 1 / ends at sz,
    cz_miscel.cpp:20-52) and any coefficient vector `cf`: seeded random windows on ragged grids from 3 nodes per axis,
    random `cf`, both values of the colour offset `ip` -- every array the calls write equals the oracle's bit for bit."""
    rng = np.random.default_rng(21)
    for _ in range(60):
        real = "f32" if rng.random() < 0.6 else "f64"
        sz = [int(rng.integers(3, 16)), int(rng.integers(3, 16)), int(rng.integers(3, 30))]
        idx = []
        for a in range(3):
            lo = int(rng.integers(1, min(3, sz[a]) + 1))
            idx += [lo, int(rng.integers(max(lo, sz[a] - 2), sz[a] + 1))]
        lib, o = _emu(real), ou.OracleCZ(sz, real, idx=idx)
        X, B = ou.random_fields(o.shape, o.dt, seed=int(rng.integers(1 << 30)))
        cf = np.array([1, 1, 1, 1, 1, 1, 6], dtype=o.dt)
        if rng.random() < 0.4:
            cf = (cf * rng.uniform(0.5, 1.5, 7)).astype(o.dt)
        om = o.creal(float(rng.choice([0.8, 1.0, 1.5]))); ip = C.c_int(int(rng.integers(0, 2)))
        outs = []
        for dll in (o.dll, lib):
            r = []
            x = X.copy(); w = np.zeros_like(X); res = C.c_double(0.5); flop = C.c_double(0)
            dll.jacobi_(ou.ptr(x), o._sz, o._idx, o.g(), ou.ptr(cf), C.byref(om), ou.ptr(B), C.byref(res), ou.ptr(w), C.byref(flop))
            r += [x, res.value, flop.value]
            x = X.copy(); res = C.c_double(0); flop = C.c_double(0)
            for color in (0, 1):
                cc = C.c_int(color)
                dll.psor2sma_core_(ou.ptr(x), o._sz, o._idx, o.g(), ou.ptr(cf), C.byref(ip), C.byref(cc), C.byref(om), ou.ptr(B),
                                   C.byref(res), C.byref(flop))
            r += [x, res.value, flop.value]
            x = X.copy(); res = C.c_double(0)
            dll.psor_(ou.ptr(x), o._sz, o._idx, o.g(), ou.ptr(cf), C.byref(om), ou.ptr(B), C.byref(res), C.byref(flop))
            r += [x, res.value]
            a = np.zeros_like(X); rk = np.zeros_like(X); flop = C.c_double(0)
            dll.blas_calc_ax_(ou.ptr(a), ou.ptr(X), o._sz, o._idx, o.g(), ou.ptr(cf), C.byref(flop))
            dll.blas_calc_rk_(ou.ptr(rk), ou.ptr(X), ou.ptr(B), o._sz, o._idx, o.g(), ou.ptr(cf), C.byref(flop))
            r += [a, rk, flop.value]
            d1 = o.creal(0); d2 = o.creal(0)
            dll.blas_dot1_(C.byref(d1), ou.ptr(X), o._sz, o._idx, o.g(), C.byref(flop))
            dll.blas_dot2_(C.byref(d2), ou.ptr(X), ou.ptr(B), o._sz, o._idx, o.g(), C.byref(flop))
            r += [d1.value, d2.value]
            m = np.zeros_like(X); dll.imask_k_(ou.ptr(m), o._sz, o._idx, o.g())
            outs.append(r + [m])
        for n, (u, v) in enumerate(zip(*outs)):
            if isinstance(u, np.ndarray):
                assert np.array_equal(u, v), (real, sz, idx, n)
            else:
                assert v == pytest.approx(u, rel=2e-6 if real == "f32" else 1e-12), (real, sz, idx, n)
