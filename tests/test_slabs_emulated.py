"""world_size-2 and -3 runs of the library's OWN multi-rank code on the CPU: solver.cu / comm.cpp / kernels.cu / maf.cu of
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
