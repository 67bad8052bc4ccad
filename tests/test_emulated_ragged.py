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
