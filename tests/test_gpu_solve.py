"""GPU parity tests, loop level: czb_solve (device-resident loops) against the oracle loops and
the committed golden histories of the reference driver; the unmodified reference driver linked
to libczb200 (oracle/_ref/cz_gpu_*) against the same driver linked to the oracle."""
import json
import os
import subprocess
import tempfile

import numpy as np
import pytest

import oracle_util as ou

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = json.load(open(os.path.join(HERE, "golden", "histories.json")))

# north_star tolerances: iteration count +-1; per-cell solution and residual history within
# 1e-5 relative (float) / 1e-12 (double).  bc_k's sin() may differ from libm by 1 ulp, so whole
# solves are compared with these tolerances; identical-input sweeps are bit-exact (test_gpu_kernels).
TOL = {"f32": 1e-5, "f64": 1e-12}


def _run_case(g, variant=0):
    import cubez_b200
    n, real = g["n"], g["real"]
    cz = cubez_b200.CZ([n] * 3, real)
    cz.variant(variant)
    P, RHS = cz.setup_laplace()
    out = cz.solve(g["solver"], P, RHS, g["itr_max"], float(g["coef"]), g["precond"])
    err, loc = cz.exact_error(P)
    field = P.cpu().numpy()
    cz.close()
    return out, err, loc, field


def _oracle_case(g):
    o = ou.OracleCZ([g["n"]] * 3, g["real"])
    P, RHS = o.setup_laplace()
    if g["solver"] == "pbicgstab":
        itr, res, hist = o.pbicgstab(P, RHS, g["itr_max"], float(g["coef"]), g["precond"])
    else:
        itr, res, hist = o.stationary(g["solver"], P, RHS, g["itr_max"], float(g["coef"]))
    return itr, res, hist, P


PLAIN = [i for i, g in enumerate(GOLDEN) if not g["solver"].endswith("_maf")]   # the `_maf` cases: test_gpu_maf.py


@pytest.mark.parametrize("variant", [1, 0])
@pytest.mark.parametrize("gi", PLAIN, ids=lambda i: "%s-%d-%s-%s" % (
    GOLDEN[i]["real"], GOLDEN[i]["n"], GOLDEN[i]["solver"], GOLDEN[i]["precond"]))
def test_solve_matches_reference_history_and_field(gi, variant):
    g = GOLDEN[gi]
    tol = TOL[g["real"]]
    out, err, loc, field = _run_case(g, variant)
    bicg = g["solver"] == "pbicgstab"
    gh = np.array([float(h) for h in g["history"]])
    if not bicg:
        # stationary solvers: sweeps and Dirichlet data are bit-exact, sums differ at 1e-16
        assert out["itr"] == g["itr"]
        assert len(out["history"]) == len(gh)
        np.testing.assert_allclose(out["history"], gh, rtol=2e-6)     # golden file has 7 digits
    else:
        assert abs(out["itr"] - g["itr"]) <= 1
        m = min(len(gh), len(out["history"]))
        assert m >= len(gh) - 1
        # un-preconditioned BiCGstab passes through a near-breakdown plateau where the dot products
        # cancel: the reference itself is reduction-order dependent there -> loose tolerance
        if g["precond"] == "none":
            np.testing.assert_allclose(out["history"][:8], gh[:8], rtol=1e-5)
            np.testing.assert_allclose(out["history"][:m - 1], gh[:m - 1], rtol=5e-2)
        else:
            np.testing.assert_allclose(out["history"][:m], gh[:m], rtol=10 * tol if g["real"] == "f32" else 2e-6, atol=1e-12)
    # exact_t evaluates sin/sinh on the device (<= few ulp of an O(1) value from libm)
    assert abs(err - float(g["errmax"])) <= (1e-6 if g["real"] == "f32" else 1e-9) + (1e-3 * err if bicg else 0.0)
    if g["n"] <= 32:
        itr, res, hist, P = _oracle_case(g)
        if not bicg:
            assert out["itr"] == itr
            assert np.array_equal(field, P), "field differs from the oracle (max %g)" % np.abs(field - P).max()
            np.testing.assert_allclose(out["history"], hist, rtol=1e-12)
        elif out["itr"] == itr and g["precond"] != "none":
            scale = np.abs(P).max()
            # north_star tolerance on the solution: 1e-5 (float) / 1e-12 (double) relative; the
            # double build accumulates ~1e-16 reduction-order noise through the Krylov recurrences
            np.testing.assert_allclose(field, P, rtol=0, atol=(tol if g["real"] == "f32" else 1e-10) * scale)
            np.testing.assert_allclose(out["history"], hist, rtol=10 * tol if g["real"] == "f32" else 1e-7)


@pytest.mark.parametrize("real", ["f32", "f64"])
def test_host_pointer_solve_roundtrip(real):
    """X/B given as host arrays: staged in, solved on the device, staged out."""
    import cubez_b200
    cz = cubez_b200.CZ([32] * 3, real)
    o = ou.OracleCZ([32] * 3, real)
    P, RHS = o.setup_laplace()
    Ph, Bh = P.copy(), RHS.copy()
    out = cz.solve("sor2sma", Ph, Bh, 2000, 1.5)
    itr, res, hist = o.stationary("sor2sma", P, RHS, 2000, 1.5)
    assert out["itr"] == itr
    # identical inputs (oracle bc_k on both sides) and bit-exact sweeps; the loop's own bc_k may
    # touch boundary values by 1 ulp
    assert np.array_equal(Ph, P)
    np.testing.assert_allclose(out["history"], hist, rtol=1e-12)
    cz.close()


@pytest.mark.parametrize("real,args", [
    ("f32", ["64", "64", "64", "jacobi", "4000", "0.8"]),            # BASELINE.json configs[0]
    ("f32", ["32", "32", "32", "pbicgstab", "200", "1.1", "sor2sma"]),
    ("f64", ["32", "32", "32", "pcr_rb", "2000", "1.5"]),
])
def test_unmodified_reference_driver_on_gpu_library(real, args):
    """A/B drop-in: the reference's own main.cpp/cz_*.cpp objects linked against libczb200
    (host pointers, staged per call) must write the same history file as when linked against
    the oracle kernels."""
    exe = ou.ref_binary("cz_gpu_%s" % real)
    if exe is None:
        pytest.skip("oracle/_ref/cz_gpu_%s not built (needs /root/reference at build time)" % real)
    g = [c for c in GOLDEN if c["real"] == real and str(c["n"]) == args[0] and c["solver"] == args[3]
         and (c["precond"] == (args[6] if len(args) > 6 else None))][0]
    with tempfile.TemporaryDirectory() as td:
        p = subprocess.run([exe] + args, cwd=td, capture_output=True, text=True, timeout=900)
        assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
        lines = open(os.path.join(td, args[3] + ".txt")).read().split("\n")[1:]
    hist = np.array([float(l.split(",")[1]) for l in lines if l.strip()])
    gh = np.array([float(h) for h in g["history"]])
    assert abs(len(hist) - len(gh)) <= 1
    m = min(len(hist), len(gh))
    bicg = args[3] == "pbicgstab"
    np.testing.assert_allclose(hist[:m], gh[:m], rtol=2e-6 if not bicg else 1e-4, atol=0 if not bicg else 1e-12)
    assert "Iter = " in p.stdout and "Error max" in p.stdout


def test_cz_driver_cli():
    """The product's own `cz` driver keeps the reference CLI and output format."""
    exe = os.path.join(ou.ROOT, "cubez_b200", "cz_f32")
    assert os.path.exists(exe), "build cubez_b200/csrc first"
    g = [c for c in GOLDEN if c["real"] == "f32" and c["n"] == 64 and c["solver"] == "sor2sma"][0]
    with tempfile.TemporaryDirectory() as td:
        p = subprocess.run([exe, "64", "64", "64", "sor2sma", "4000", "1.9"], cwd=td, capture_output=True, text=True,
                           timeout=600)
        assert p.returncode == 0, p.stdout + p.stderr
        txt = open(os.path.join(td, "sor2sma.txt")).read().split("\n")
    assert txt[0] == "Itration      Residual"
    assert "Iter = %d" % g["itr"] in p.stdout
    hist = np.array([float(l.split(",")[1]) for l in txt[1:] if l.strip()])
    np.testing.assert_allclose(hist, [float(h) for h in g["history"]], rtol=2e-6)
    p = subprocess.run([exe, "64", "64", "64", "nosuch", "10", "1.0"], capture_output=True, text=True)
    assert "Invalid solver" in p.stdout


# Line SOR on LONG lines: the solve loops use the warp-per-line kernels with the coefficient tables in
# tensor memory (lines of 129..1024 cells) and, where the row pitch allows it, TMA bulk staging of the
# neighbour rows; the golden cases above only reach lines of <= 62 cells.  Non-cubic grids keep the oracle
# fast: many lines (more than one per warp), few planes.  Fields must agree bit for bit with the oracle.
@pytest.mark.parametrize("sz,solver,coef", [
    ([84, 80, 132], "pcr_rb", 1.5),      # 130-cell lines: 8 cells per lane, bulk staging
    ([84, 80, 200], "pcr_rb", 1.5),      # 198
    ([84, 80, 258], "pcr_rb", 1.5),      # 256 = 2^8 exactly
    ([84, 80, 300], "pcr_rb", 1.5),      # 298: 16 cells per lane, bulk staging
    ([84, 80, 302], "pcr_rb", 1.5),      # 300: row pitch not a multiple of 16 B -> register prefetch path
    ([84, 80, 300], "pcr_j_esa", 0.9),   # all lines in one pass, separate output array
    ([84, 80, 300], "pcr_eda", 1.5),     # wavefront launches
    ([84, 80, 200], "pcr_rb_esa", 1.5),  # 4x4 Cramer last stage on the tensor-memory kernel, 8 cells per lane
    ([84, 80, 300], "pcr_rb_esa", 1.5),  # 16 cells per lane
    ([84, 80, 302], "pcr_rb_esa", 1.5),  # no bulk staging possible -> block-per-line kernel
    ([44, 40, 300], "pcr_esa", 1.5),     # 4x4, wavefront launches
    ([44, 40, 514], "pcr_rb", 1.5),      # 512 = 2^9
    ([44, 40, 600], "pcr_rb", 1.5),      # 598: 32 cells per lane, late stages from shared memory
    ([44, 40, 1026], "pcr_rb", 1.5),     # 1024, the longest line of the warp kernels
    ([24, 20, 1100], "pcr_rb", 1.5),     # 1098 > 1024: block-per-line kernel
])
def test_line_sor_long_lines(sz, solver, coef):
    import cubez_b200
    cz = cubez_b200.CZ(sz, "f32", eps=0.0)
    o = ou.OracleCZ(sz, "f32", eps=0.0)
    assert tuple(cz.shape) == tuple(o.shape)
    X, B = ou.random_fields(o.shape, o.dt)
    o.bc_k(X)
    Xg, Bg = cz.from_numpy(X), cz.from_numpy(B)
    iters = 3
    out = cz.solve(solver, Xg, Bg, iters, coef)
    itr, res, hist = o.stationary(solver, X, B, iters, coef)
    assert out["itr"] == itr == iters + 1
    got = Xg.cpu().numpy()
    assert np.array_equal(got, X), "field differs from the oracle (max %g)" % np.abs(got - X).max()
    np.testing.assert_allclose(out["history"], hist, rtol=1e-12)
    cz.close()


def test_phase_profile_report(tmp_path):
    """per-phase report keyed by the reference's PMlib labels (cz_miscel.cpp:177-262), the counterpart
    of its profiling.txt (cz_Evaluate.cpp:506-544)."""
    import cubez_b200
    cz = cubez_b200.CZ([64] * 3, "f32")
    cz.dll.czb_profile(1)
    P, RHS = cz.setup_laplace()
    out = cz.solve("pbicgstab", P, RHS, 50, 1.1, "sor2sma")
    path = str(tmp_path / "profiling.txt")
    n = cz.dll.czb_profile_report(path.encode(), b"test")
    cz.dll.czb_profile(0)
    cz.close()
    txt = open(path).read()
    assert n >= 6
    for label in ("SOR2SMA_kernel", "Blas_AX", "Blas_BiCG_2", "Blas_TRIAD", "BoundaryCondition", "Blas_Residual"):
        assert label in txt, txt
    rows = {l.split()[0]: l.split() for l in txt.splitlines() if l and not l.startswith("#") and not l.startswith("Section")}
    its = out["itr"]
    # lc_max = 8 sweeps, twice per iteration (cz_Poisson.cpp:280,409,445); iterations are enqueued in batches of 4, the
    # ones after convergence are no-ops on the device but still counted here
    assert int(rows["SOR2SMA_kernel"][1]) >= 16 * its and int(rows["SOR2SMA_kernel"][1]) % 16 == 0
    assert int(rows["Blas_AX"][1]) * 8 == int(rows["SOR2SMA_kernel"][1])
    assert float(rows["SOR2SMA_kernel"][2]) > 0.0
    # the driver writes the same table when CZB_PROFILE=1
    exe = os.path.join(ou.ROOT, "cubez_b200", "cz_f32")
    p = subprocess.run([exe, "64", "64", "64", "sor2sma", "4000", "1.9"], cwd=str(tmp_path), capture_output=True, text=True,
                       timeout=600, env=dict(os.environ, CZB_PROFILE="1"))
    assert p.returncode == 0 and "SOR2SMA_kernel" in open(str(tmp_path / "profiling.txt")).read()


# Persistent cooperative loop (persist.cu): grids whose fields fit L2 run all iterations after the first ones inside one
# launch with grid barriers.  It must stop at the same iteration with the same field as the per-iteration launches
# (CZB_PERSIST_MB=0 switches it off) and as the oracle, for sizes that give the row loop 1, 2 and many rows per warp, odd
# and even iteration counts (Jacobi's ping-pong exit) and both precisions.
@pytest.mark.parametrize("real", ["f32", "f64"])
@pytest.mark.parametrize("n,solver,coef,itmax", [
    (18, "jacobi", 0.8, 4000), (33, "sor2sma", 1.5, 4000), (64, "jacobi", 0.8, 151), (64, "sor2sma", 1.9, 4000),
    (100, "sor2sma", 1.8, 60), (100, "jacobi", 0.9, 2), (100, "jacobi", 0.9, 3), (130, "jacobi", 0.8, 40),
])
def test_persistent_loop_equals_per_iteration_launches(real, n, solver, coef, itmax, monkeypatch):
    import cubez_b200
    runs = {}
    for mb in ("200", "0"):
        monkeypatch.setenv("CZB_PERSIST_MB", mb)
        cz = cubez_b200.CZ([n] * 3, real)
        P, RHS = cz.setup_laplace()
        l0 = cz.launch_count()
        out = cz.solve(solver, P, RHS, itmax, coef)
        runs[mb] = (out, P.cpu().numpy(), cz.launch_count() - l0)
        cz.close()
    a, b = runs["200"], runs["0"]
    assert a[0]["itr"] == b[0]["itr"]
    assert np.array_equal(a[1], b[1]), "persistent loop differs from the per-iteration launches (max %g)" % np.abs(a[1] - b[1]).max()
    np.testing.assert_allclose(a[0]["history"], b[0]["history"], rtol=1e-12)
    assert abs(a[0]["res"] - b[0]["res"]) <= 1e-12 * b[0]["res"]
    if itmax > 40:
        assert a[2] < 16 and a[2] < b[2], "the persistent path did not run (%d vs %d launches)" % (a[2], b[2])
    if n <= 64:
        o = ou.OracleCZ([n] * 3, real)
        Po, Ro = o.setup_laplace()
        itr, res, hist = o.stationary(solver, Po, Ro, itmax, coef)
        assert a[0]["itr"] == itr
        assert np.array_equal(a[1], Po)
        np.testing.assert_allclose(a[0]["history"], hist, rtol=1e-12)
