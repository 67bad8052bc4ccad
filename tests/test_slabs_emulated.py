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


# ---- the `cz` driver as its own launcher (cubez_b200/csrc/cz_main.cpp: launch_ranks) ------------------------------------------
def _cz_env(work):
    sys.path.insert(0, os.path.join(HERE, "tools", "emu"))
    import emu_build
    return emu_build.build_cz(), dict(os.environ, CZB_NCCL_LIB=emu_build.build_nccl(), CZB_NCCL_EMU_DIR=work, CZB_RANK_GRACE_S="1")


def test_cz_driver_launches_its_ranks_and_writes_the_serial_history(tmp_path):
    """`cz ... --gpus 2` and the reference's own `gdv_x gdv_y gdv_z` form (main.cpp:19-29): the parent re-executes itself once
    per rank; rank 0 writes the history file of the serial run, the parent prints the global Error max."""
    one, two, three = [os.path.join(str(tmp_path), d) for d in ("one", "two", "three")]
    res = {}
    for work, extra in ((one, []), (two, ["--gpus", "2"]), (three, ["1", "2", "1"])):
        os.makedirs(work)
        exe, env = _cz_env(work)
        p = subprocess.run([exe, "32", "32", "32", "sor2sma", "400", "1.5"] + extra, cwd=work, env=env, capture_output=True,
                           text=True, timeout=600)
        assert p.returncode == 0, p.stdout + p.stderr
        res[work] = (open(os.path.join(work, "sor2sma.txt")).read(), p.stdout.split("Error max = ")[1].split()[0],
                     p.stdout.split("Iter = ")[1].split()[0])
    h1 = [float(l.split(",")[1]) for l in res[one][0].split("\n")[1:] if l.strip()]
    for work in (two, three):
        assert res[work][2] == res[one][2]                       # the same iteration count
        assert res[work][1] == res[one][1]                       # bit-identical field -> the same Error max digits
        h = [float(l.split(",")[1]) for l in res[work][0].split("\n")[1:] if l.strip()]
        np.testing.assert_allclose(h, h1, rtol=2e-6)             # residual sums differ by summation order only


def test_cz_launcher_ends_the_job_when_a_rank_dies(tmp_path):
    """A rank that dies would leave its peer waiting in the next exchange for ever (mpirun's job under the reference):
    the launcher notices, gives the rest CZB_RANK_GRACE_S seconds, ends them and reports a solver error."""
    import signal
    import time
    work = str(tmp_path)
    exe, env = _cz_env(work)
    q = subprocess.Popen([exe, "48", "48", "48", "jacobi", "2000000", "0.8", "--gpus", "2"], cwd=work, env=dict(env, CZ_EPS="0"),
                         stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    try:
        kids = []
        for _ in range(100):                                     # the two ranks are the parent's only children
            time.sleep(0.1)
            kids = open("/proc/%d/task/%d/children" % (q.pid, q.pid)).read().split()
            if len(kids) == 2:
                break
        assert len(kids) == 2, kids
        time.sleep(1.0)
        os.kill(int(kids[1]), signal.SIGKILL)
        t0 = time.time()
        out, err = q.communicate(timeout=120)
    finally:
        if q.poll() is None:
            q.kill()
    assert q.returncode != 0 and "Solver error" in out and "rank 1 ended abnormally" in err, (q.returncode, out, err)
    assert time.time() - t0 < 30
    for k in kids:                                               # both ranks were reaped by the launcher
        assert not os.path.exists("/proc/%s" % k), k


def test_cz_driver_on_ragged_grids_and_random_divisions(tmp_path):
    """Seeded random non-cubic grids cut into random bricks through the `cz` launcher (2 .. 12 processes): the history
    file and the iteration count are those of the undivided run.  "Error max" is compared to 1e-5 only: a rank evaluates
    the analytic solution from ITS origin (`origin + (i - 1) * pitch` in REAL, cz_Evaluate.cpp:136-138), which rounds
    differently from the global expression -- the reference does the same."""
    import re
    rng = np.random.default_rng(5)
    solvers = ["jacobi", "sor2sma", "pcr_rb", "pcr_j_esa", "jacobi_maf", "sor2sma_maf", "pcr_rb_maf", "pbicgstab", "pbicgstab_maf"]

    def run(work, args):
        os.makedirs(work)
        exe, env = _cz_env(work)
        p = subprocess.run([exe] + args, cwd=work, env=env, capture_output=True, text=True, timeout=600)
        assert p.returncode == 0, (args, p.stdout[-2000:], p.stderr[-2000:])
        h = [float(l.split(",")[1]) for l in open(os.path.join(work, args[3] + ".txt")).read().split("\n")[1:] if l.strip()]
        return int(re.search(r"Iter = (\d+)", p.stdout).group(1)), float(p.stdout.split("Error max = ")[1].split()[0]), h

    for t in range(12):
        sz = [int(rng.integers(6, 20)) for _ in range(3)]
        solver = solvers[int(rng.integers(len(solvers)))]
        gdv = [int(rng.integers(1, 3)), int(rng.integers(1, 4)), 1 if "pcr" in solver else int(rng.integers(1, 3))]
        if gdv == [1, 1, 1]:
            gdv = [1, 2, 1]
        bicg = solver.startswith("pbicgstab")
        pre = [str(rng.choice(["sor2sma", "jacobi", "none"]))] if bicg else []
        if bicg and solver.endswith("_maf") and pre[0] != "none":
            pre = [pre[0] + "_maf"]
        coef = "0.8" if ("jacobi" in solver or (pre and "jacobi" in pre[0])) else str(rng.choice(["1.0", "1.5", "1.8"]))
        args = [str(s) for s in sz] + [solver, str(int(rng.integers(3, 15))), coef] + pre
        a = run(os.path.join(str(tmp_path), "a%d" % t), args)
        b = run(os.path.join(str(tmp_path), "b%d" % t), args + [str(g) for g in gdv])
        assert a[0] == b[0], (args, gdv)
        assert len(a[2]) == len(b[2]), (args, gdv)
        np.testing.assert_allclose(b[2], a[2], rtol=2e-3 if bicg else 2e-6, err_msg=str((args, gdv)))
        assert b[1] == pytest.approx(a[1], rel=1e-5 if not bicg else 1e-3), (args, gdv)
