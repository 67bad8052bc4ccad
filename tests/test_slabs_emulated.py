"""world_size-2, -3 (j-slabs) and 2..8 (3-D bricks) runs of the library's OWN multi-rank code on the CPU: solver.cu / comm.cpp / kernels.cu / maf.cu of
the CPU-emulated build (tests/tools/emu), one OS process per rank, NCCL replaced by a file-based stand-in loaded through
comm.cpp's CZB_NCCL_LIB switch.  Same property as the multi-GPU test (tests/mgpu_worker.py): the gathered slabs
reproduce the single-domain run -- bit-exact fields and equal iteration counts for the stationary solvers (halo
exchange after each colour, global colour parity, global Dirichlet / mesh indices), histories to summation-order
tolerance.  Includes the stretched-grid (`_maf`) solvers, whose slab path has no multi-GPU test yet."""
import os
import subprocess
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
WORKER = os.path.join(HERE, "tools", "emu", "slab_worker.py")


def _run(tmp, nranks, n, solver, pre, omega, iters, eps):
    work = os.path.join(str(tmp), "%s_%s_%d" % (solver, pre, nranks))
    os.makedirs(work)
    procs = [subprocess.Popen([sys.executable, WORKER, str(r), str(nranks), work, str(n), solver, pre or "-", str(omega),
                               str(iters), repr(eps)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
             for r in range(nranks)]
    outs = [p.communicate(timeout=900)[0] for p in procs]
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o[-3000:]
    parts = [np.load(os.path.join(work, "out_%d.npz" % r)) for r in range(nranks)]
    return np.concatenate([q["own"] for q in parts], 0), [int(q["itr"]) for q in parts], [q["hist"] for q in parts]


CASES = [("sor2sma", None, 1.5, 25), ("jacobi", None, 0.8, 25), ("sor2sma_maf", None, 1.5, 25), ("jacobi_maf", None, 0.8, 20),
         ("pcr_rb_maf", None, 1.5, 15), ("pcr_rb_esa_maf", None, 1.5, 6), ("pcr_rb", None, 1.5, 12), ("pcr_j_esa", None, 0.9, 12)]


@pytest.mark.parametrize("solver,pre,omega,iters", CASES)
@pytest.mark.parametrize("nranks", [2, 3])
def test_slabs_reproduce_the_single_domain_bit_for_bit(tmp_path, solver, pre, omega, iters, nranks):
    n = 20
    ref, itr1, h1 = _run(tmp_path, 1, n, solver, pre, omega, iters, 0.0)
    full, itrs, hs = _run(tmp_path, nranks, n, solver, pre, omega, iters, 0.0)
    assert full.shape == ref.shape
    assert np.array_equal(full, ref), np.abs(full - ref).max()
    assert all(i == itr1[0] for i in itrs)
    for h in hs:                                   # every rank derives the same history from the all-reduced sums
        np.testing.assert_allclose(h, h1[0], rtol=1e-12)


@pytest.mark.parametrize("solver,pre", [("pbicgstab", "sor2sma"), ("pbicgstab_maf", "sor2sma_maf"), ("pbicgstab_maf", "pcr_rb_maf")])
def test_slab_bicgstab_matches_the_single_domain(tmp_path, solver, pre):
    n = 20
    ref, itr1, h1 = _run(tmp_path, 1, n, solver, pre, 1.1, 40, 1e-5)
    full, itrs, hs = _run(tmp_path, 2, n, solver, pre, 1.1, 40, 1e-5)
    assert all(abs(i - itr1[0]) <= 1 for i in itrs) and len(set(itrs)) == 1
    np.testing.assert_allclose(full, ref, rtol=0, atol=1e-5 * np.abs(ref).max())
    m = min(len(hs[0]), len(h1[0])) - 1
    np.testing.assert_allclose(hs[0][:m], h1[0][:m], rtol=1e-3)


# ---- 3-D bricks (reference: CBrick division, cz_Evaluate.cpp:103-159, `gdv_x gdv_y gdv_z`) ----------------
def _run_bricks(tmp, gdv, n, solver, pre, omega, iters, eps):
    nranks = gdv[0] * gdv[1] * gdv[2]
    work = os.path.join(str(tmp), "%s_%s_%s" % (solver, pre, "x".join(map(str, gdv))))
    os.makedirs(work)
    procs = [subprocess.Popen([sys.executable, WORKER, str(r), str(nranks), work, str(n), solver, pre or "-", str(omega),
                               str(iters), repr(eps), ",".join(map(str, gdv))], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
             for r in range(nranks)]
    outs = [p.communicate(timeout=900)[0] for p in procs]
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o[-3000:]
    parts = [np.load(os.path.join(work, "out_%d.npz" % r)) for r in range(nranks)]
    full = np.full((n - 2, n - 2, n - 2), np.nan, np.float32)          # inner cells of the global grid, (j, i, k)
    for q in parts:
        j0, i0, k0 = [int(v) - 3 for v in q["at"]]                       # global C index 3 = first inner node (Fortran 2)
        o = q["own"]
        assert np.isnan(full[j0:j0 + o.shape[0], i0:i0 + o.shape[1], k0:k0 + o.shape[2]]).all()     # no overlap
        full[j0:j0 + o.shape[0], i0:i0 + o.shape[1], k0:k0 + o.shape[2]] = o
    assert not np.isnan(full).any()                                      # the bricks tile the cube
    return full, [int(q["itr"]) for q in parts], [q["hist"] for q in parts], outs


BRICKS = [([2, 1, 1], "sor2sma", None, 1.5, 20), ([1, 1, 2], "sor2sma", None, 1.5, 20), ([2, 2, 1], "jacobi", None, 0.8, 20),
          ([2, 1, 2], "sor2sma", None, 1.8, 15), ([2, 2, 2], "sor2sma", None, 1.5, 12), ([3, 1, 2], "jacobi", None, 0.8, 12),
          ([2, 2, 1], "pcr_rb", None, 1.5, 8), ([2, 1, 1], "pcr_j_esa", None, 0.9, 8), ([2, 1, 2], "sor2sma_maf", None, 1.5, 12),
          ([1, 2, 2], "jacobi_maf", None, 0.8, 10)]


@pytest.mark.parametrize("gdv,solver,pre,omega,iters", BRICKS, ids=lambda v: "x".join(map(str, v)) if isinstance(v, list) else str(v))
def test_bricks_reproduce_the_single_domain_bit_for_bit(tmp_path, gdv, solver, pre, omega, iters):
    n = 20
    ref, itr1, h1 = _run(tmp_path, 1, n, solver, pre, omega, iters, 0.0)
    full, itrs, hs, _ = _run_bricks(tmp_path, gdv, n, solver, pre, omega, iters, 0.0)
    inner = ref[:, 3:-3, 3:-3]                    # _run returns the owned inner PLANES (all i, k incl. guards)
    assert full.shape == inner.shape
    assert np.array_equal(full, inner), np.abs(full - inner).max()
    assert all(i == itr1[0] for i in itrs)
    for h in hs:
        np.testing.assert_allclose(h, h1[0], rtol=1e-12)


def test_brick_bicgstab_matches_the_single_domain(tmp_path):
    n = 20
    ref, itr1, h1 = _run(tmp_path, 1, n, "pbicgstab", "sor2sma", 1.1, 40, 1e-5)
    full, itrs, hs, _ = _run_bricks(tmp_path, [2, 1, 2], n, "pbicgstab", "sor2sma", 1.1, 40, 1e-5)
    assert all(abs(i - itr1[0]) <= 1 for i in itrs) and len(set(itrs)) == 1
    inner = ref[:, 3:-3, 3:-3]
    np.testing.assert_allclose(full, inner, rtol=0, atol=1e-5 * np.abs(inner).max())
    m = min(len(hs[0]), len(h1[0])) - 1
    np.testing.assert_allclose(hs[0][:m], h1[0][:m], rtol=1e-3)


def test_line_solvers_refuse_a_division_along_k(tmp_path):
    """PCR lines run along k: a division along z would cut them -> czb_solve reports a solver error (returns 0)."""
    full, itrs, hs, outs = None, None, None, None
    nranks, gdv = 2, [1, 1, 2]
    work = os.path.join(str(tmp_path), "refuse")
    os.makedirs(work)
    procs = [subprocess.Popen([sys.executable, WORKER, str(r), str(nranks), work, "20", "pcr_rb", "-", "1.5", "5", "0.0", "1,1,2"],
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True) for r in range(nranks)]
    outs = [p.communicate(timeout=300)[0] for p in procs]
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o[-2000:]
        assert "gdv_z = 1" in o
    assert all(int(np.load(os.path.join(work, "out_%d.npz" % r))["itr"]) == 0 for r in range(nranks))
