"""CPU run of the LIBRARY ITSELF for the stretched-grid (MAF) row of SURVEY.md §8f-2: cubez_b200/csrc (abi.cu, solver.cu,
kernels.cu, maf.cu, runtime.cu) compiled for the CPU by tests/tools/emu (CUDA built-ins emulated with one thread per
block, launches executed sequentially, device memory = malloc) and driven through the same C ABI the GPU tests use.

* kernel level: every `*_maf_` entry point on a stretched mesh with seeded random fields == the oracle, bit for bit;
* loop level: czb_solve with the `*_maf` names == the golden histories written by the unmodified reference driver.

This covers what a GPU-less container can: index arithmetic, launch geometry, colour / wavefront line selection,
shared-memory ping-pong, the folded 2x2 last stage, reduction finishers, host-pointer staging and the loop control.
It cannot show barrier placement or races; the -m gpu tests (test_gpu_maf.py) remain the parity tests proper, and
this emulated library is test infrastructure that never ships (the product aborts without a CUDA device)."""
import ctypes as C
import json
import os
import sys

import numpy as np
import pytest

import oracle_util as ou

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "tools", "emu"))
import emu_build  # noqa: E402

GOLDEN = json.load(open(os.path.join(HERE, "golden", "histories.json")))
_LIBS = {}


def _emu(real):
    import cubez_b200.ffi as ffi
    if not _LIBS:
        for k, path in emu_build.build_lib(sanitize=os.environ.get("EMU_SANITIZE") or None).items():
            dll = C.CDLL(path)
            ffi.bind_reference_symbols(dll)
            ffi.bind_czb_symbols(dll, C.c_float if k == "f32" else C.c_double)
            dll.czb_init(0)
            _LIBS[k] = dll
    lib = _LIBS[real]
    assert lib.czb_real_bytes() == (4 if real == "f32" else 8)
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
    g = o.g()
    cells = float((sz[0] - 2) * (sz[1] - 2) * (sz[2] - 2))
    tmp = np.zeros(sz[2] + 4, dtype=o.dt)
    # jacobi_maf_
    o.ac1 = o.dt(0.8)
    Xo = X.copy(); ro = C.c_double(0.25); o.sweep_maf("jacobi_maf", Xo, B, ro, mesh)
    Xe = X.copy(); W = np.zeros_like(X); res = C.c_double(0.25); flop = C.c_double(0); om = R(0.8)
    lib.jacobi_maf_(ou.ptr(Xe), o._sz, o._idx, g, *mp, C.byref(om), ou.ptr(B), C.byref(res), ou.ptr(W), ou.ptr(tmp), C.byref(flop))
    assert np.array_equal(Xe, Xo)
    assert res.value == pytest.approx(ro.value, rel=1e-12)
    assert flop.value == 66.0 * cells
    # psor2sma_core_maf_, colour 0 then 1
    o.ac1 = o.dt(1.5)
    Xo = X.copy(); ro = C.c_double(0); o.sweep_maf("sor2sma_maf", Xo, B, ro, mesh)
    Xe = X.copy(); res = C.c_double(0); flop = C.c_double(0); om = R(1.5); ip = C.c_int(0)
    for color in (0, 1):
        cc = C.c_int(color)
        lib.psor2sma_core_maf_(ou.ptr(Xe), o._sz, o._idx, g, *mp, C.byref(ip), C.byref(cc), C.byref(om), ou.ptr(B),
                               C.byref(res), ou.ptr(tmp), C.byref(flop))
    assert np.array_equal(Xe, Xo)
    assert res.value == pytest.approx(ro.value, rel=1e-12)
    assert flop.value == 66.0 * cells
    # psor_maf_ through hyperplane wavefronts
    Xo = X.copy(); ro = C.c_double(0); o.sweep_maf("psor_maf", Xo, B, ro, mesh)
    Xe = X.copy(); res = C.c_double(0)
    lib.psor_maf_(ou.ptr(Xe), o._sz, o._idx, g, *mp, C.byref(om), ou.ptr(B), C.byref(res), C.byref(flop))
    assert np.array_equal(Xe, Xo)
    assert res.value == pytest.approx(ro.value, rel=1e-12)
    # search_pivot_, calc_ax_maf_, calc_rk_maf_
    Po = o.pivot(mesh)
    Pe = np.zeros_like(X); lib.search_pivot_(ou.ptr(Pe), o._sz, o._idx, g, *mp)
    assert np.array_equal(Pe, Po)
    flop = C.c_double(0); fo = C.c_double(0)
    Ao = np.zeros_like(X); o.dll.calc_ax_maf_(ou.ptr(Ao), ou.ptr(X), o._sz, o._idx, g, *mp, ou.ptr(Po), C.byref(fo))
    Ae = np.zeros_like(X); lib.calc_ax_maf_(ou.ptr(Ae), ou.ptr(X), o._sz, o._idx, g, *mp, ou.ptr(Pe), C.byref(flop))
    assert np.array_equal(Ae, Ao)
    Ro = np.zeros_like(X); o.dll.calc_rk_maf_(ou.ptr(Ro), ou.ptr(X), ou.ptr(B), o._sz, o._idx, g, *mp, ou.ptr(Po), C.byref(fo))
    Re = np.zeros_like(X); lib.calc_rk_maf_(ou.ptr(Re), ou.ptr(X), ou.ptr(B), o._sz, o._idx, g, *mp, ou.ptr(Pe), C.byref(flop))
    assert np.array_equal(Re, Ro)
    assert flop.value == fo.value == 2 * 63.0 * cells


@pytest.mark.parametrize("real", ["f32", "f64"])
@pytest.mark.parametrize("sz", [[9, 9, 9], [12, 7, 23], [10, 8, 34], [8, 6, 66], [6, 6, 131], [7, 5, 300]])
def test_emulated_maf_line_kernel_equals_oracle(real, sz):
    """The five MAF line routines (k_pcr_lines_maf: colours and i+j wavefronts); line lengths of 2^m and 2^m +- 1."""
    lib = _emu(real)
    o, X, B, mesh, mp = _case(real, sz)
    M = o.mask()
    o.ac1 = o.dt(1.3)
    pn = C.c_int(ou.get_num_stage(o.idx[5] - o.idx[4] + 1)); s = C.c_int(2 ** (pn.value - 2)); ip = C.c_int(0)
    om = o.creal(1.3)
    g = o.g()
    w = [None] * 6
    for kind in ("pcr_rb_maf", "pcr_rb_esa_maf", "pcr_maf", "pcr_eda_maf", "pcr_esa_maf"):
        Xo = X.copy(); ro = C.c_double(0.5); fo = o.sweep_maf(kind, Xo, B, ro, mesh)
        Xe = X.copy(); res = C.c_double(0.5); flop = C.c_double(0)
        head = (o._sz, o._idx, g, C.byref(pn))
        tail = (C.byref(om), C.byref(res), None, C.byref(flop))
        fields = (ou.ptr(Xe), ou.ptr(M), ou.ptr(B))
        if kind == "pcr_rb_maf":
            for color in (0, 1):
                cc = C.c_int(color)
                lib.pcr_rb_maf_(*head, C.byref(ip), C.byref(cc), *fields, *mp, *w, *tail)
        elif kind == "pcr_rb_esa_maf":
            for color in (0, 1):
                cc = C.c_int(color)
                lib.pcr_rb_esa_maf_(*head, C.byref(ip), C.byref(cc), C.byref(s), *fields, *mp, *w, *tail)
        elif kind == "pcr_maf":
            lib.pcr_maf_(*head, *fields, *mp, *w, *tail)
        elif kind == "pcr_eda_maf":
            lib.pcr_eda_maf_(*head, *fields, *mp, *w[:3], *tail)
        else:
            lib.pcr_esa_maf_(*head, C.byref(s), *fields, *mp, *w, *tail)
        assert np.array_equal(Xe, Xo), (kind, np.abs(Xe - Xo).max())
        assert res.value == pytest.approx(ro.value, rel=1e-12), kind
        assert flop.value == fo, kind


def _solve(lib, real, n, solver, precond, itr_max, coef, mesh=None):
    o = ou.OracleCZ([n] * 3, real)
    P, RHS = o.setup_laplace()                         # Dirichlet data from the oracle's bc_k (host libm, as the library's table)
    ctx = lib.czb_ctx_create(o._sz, o._idx, o._nID, o.creal(o.pitch), o._org)
    if mesh is not None:
        lib.czb_ctx_set_mesh(ctx, *[ou.ptr(m) for m in mesh])
    hist = (C.c_double * (itr_max + 2))()
    res = C.c_double(0); flop = C.c_double(0)
    itr = lib.czb_solve(ctx, solver.encode(), precond.encode() if precond else None, ou.ptr(P), ou.ptr(RHS), itr_max,
                        o.creal(coef), C.byref(res), hist, itr_max + 2, C.byref(flop))
    lib.czb_ctx_destroy(ctx)
    nh = min(itr, itr_max - 1 if solver.startswith("pbicgstab") else itr_max)
    return o, itr, res.value, np.array(hist[:nh]), P, flop.value


MAF = [i for i, g in enumerate(GOLDEN) if g["solver"].endswith("_maf") and g["real"] == "f32"] + \
      [i for i, g in enumerate(GOLDEN) if g["solver"] == "sor2sma_maf" and g["real"] == "f64"]


@pytest.mark.parametrize("gi", MAF, ids=lambda i: "%s-%d-%s-%s" % (
    GOLDEN[i]["real"], GOLDEN[i]["n"], GOLDEN[i]["solver"], GOLDEN[i]["precond"]))
def test_emulated_maf_solve_matches_reference_history(gi):
    """czb_solve (solver.cu loops: batching, device-resident convergence flag, ping-pong, preconditioner dispatch,
    BiCGstab scalars) with the MAF kernels against the golden histories of the unmodified reference driver."""
    g = GOLDEN[gi]
    real = g["real"]
    lib = _emu(real)
    o, itr, res, hist, P, flop = _solve(lib, real, g["n"], g["solver"], g["precond"], g["itr_max"], float(g["coef"]))
    gh = np.array([float(h) for h in g["history"]])
    if g["solver"] != "pbicgstab_maf":
        assert itr == g["itr"]
        assert len(hist) == len(gh)
        np.testing.assert_allclose(hist, gh, rtol=2e-6)               # golden file has 7 digits
        Po, RHSo = o.setup_laplace()
        itr_o, res_o, hist_o = o.stationary(g["solver"], Po, RHSo, g["itr_max"], float(g["coef"]))
        assert itr == itr_o
        assert np.array_equal(P, Po), "field differs from the oracle (max %g)" % np.abs(P - Po).max()
        np.testing.assert_allclose(hist, hist_o, rtol=1e-12)
        assert flop > 0
    else:
        assert abs(itr - g["itr"]) <= 1
        m = min(len(gh), len(hist))
        assert m >= len(gh) - 1
        if g["precond"] == "none":
            np.testing.assert_allclose(hist[:6], gh[:6], rtol=1e-5)
        else:
            np.testing.assert_allclose(hist[:m], gh[:m], rtol=1e-4, atol=1e-12)


def test_emulated_plain_solvers_still_match_golden():
    """The same emulated library on the uniform-grid baseline kernels (kernels.cu; fused.cu is stubbed out, so the loops
    take the per-colour path): guards the loop code shared by the plain and the MAF solvers."""
    lib = _emu("f32")
    for solver, coef in (("jacobi", "0.8"), ("sor2sma", "1.5"), ("psor", "1.5")):
        g = [c for c in GOLDEN if c["real"] == "f32" and c["n"] == 32 and c["solver"] == solver and c["coef"] == coef][0]
        o, itr, res, hist, P, flop = _solve(lib, "f32", 32, solver, None, g["itr_max"], float(coef))
        assert itr == g["itr"]
        np.testing.assert_allclose(hist, [float(h) for h in g["history"]], rtol=2e-6)
    for pc in ("sor2sma", "jacobi"):
        g = [c for c in GOLDEN if c["real"] == "f32" and c["n"] == 32 and c["solver"] == "pbicgstab" and c["precond"] == pc][0]
        o, itr, res, hist, P, flop = _solve(lib, "f32", 32, "pbicgstab", pc, g["itr_max"], float(g["coef"]))
        gh = np.array([float(h) for h in g["history"]])
        assert abs(itr - g["itr"]) <= 1
        m = min(len(gh), len(hist))
        np.testing.assert_allclose(hist[:m], gh[:m], rtol=1e-4, atol=1e-12)


def test_emulated_stretched_mesh_solve_equals_oracle():
    lib = _emu("f32")
    o = ou.OracleCZ([24] * 3, "f32")
    mesh = o.mesh(stretch=0.2)
    o2, itr, res, hist, P, flop = _solve(lib, "f32", 24, "sor2sma_maf", None, 500, 1.5, mesh=mesh)
    Po, RHSo = o.setup_laplace()
    itr_o, res_o, hist_o = o.stationary("sor2sma_maf", Po, RHSo, 500, 1.5, mesh=mesh)
    assert itr == itr_o
    assert np.array_equal(P, Po)
    np.testing.assert_allclose(hist, hist_o, rtol=1e-12)


def test_emulated_fileout_sph_records(tmp_path, monkeypatch):
    """fileout_t_ (cz_utility.f90:17-44) with CZB_FILEOUT=1: the .sph record sequence of the reference's _aurora_ build.
    Run in a fresh process because the switch is read once."""
    import subprocess
    code = r'''
import ctypes as C, os, sys, struct
import numpy as np
sys.path.insert(0, %r); sys.path.insert(0, %r)
import emu_build, cubez_b200.ffi as ffi
dll = C.CDLL(emu_build.build_lib()["f32"]); ffi.bind_reference_symbols(dll)
sz = (C.c_int * 3)(5, 4, 3); g = C.c_int(2)
s = np.arange(9 * 8 * 7, dtype=np.float32).reshape(8, 9, 7)          # [j][i][k] with guards
dh = C.c_float(0.25); org = (C.c_float * 3)(1.0, 2.0, 3.0)
dll.fileout_t_(sz, C.byref(g), C.c_void_p(s.ctypes.data), C.byref(dh), org, b"p_00000.sph")
raw = open("p_00000.sph", "rb").read()
recs, o = [], 0
while o < len(raw):
    n = struct.unpack_from("<i", raw, o)[0]; recs.append(raw[o + 4:o + 4 + n]); assert struct.unpack_from("<i", raw, o + 4 + n)[0] == n; o += n + 8
assert [len(r) for r in recs] == [8, 12, 12, 12, 8, 5 * 4 * 3 * 4], [len(r) for r in recs]
assert struct.unpack("<2i", recs[0]) == (1, 1) and struct.unpack("<3i", recs[1]) == (5, 4, 3)
assert struct.unpack("<3f", recs[2]) == (1.0, 2.0, 3.0) and struct.unpack("<3f", recs[3]) == (0.25, 0.25, 0.25)
data = np.frombuffer(recs[5], dtype=np.float32).reshape(3, 4, 5)     # [k][j][i]
assert np.array_equal(data, s[2:6, 2:7, 2:5].transpose(2, 0, 1))
print("ok")
''' % (os.path.join(HERE, "tools", "emu"), os.path.dirname(HERE))
    env = dict(os.environ, CZB_FILEOUT="1")
    p = subprocess.run([sys.executable, "-c", code], cwd=str(tmp_path), env=env, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0 and "ok" in p.stdout, p.stdout + p.stderr


def test_emulated_cz_driver_cli_maf(tmp_path):
    """The product's `cz` driver (reference CLI, console and history-file format) over the emulated library: a `_maf`
    run writes the golden history, reports the analytic flop count and, with CZB_FILEOUT=1, the two .sph dumps."""
    import subprocess
    exe = emu_build.build_cz()
    g = [c for c in GOLDEN if c["real"] == "f32" and c["n"] == 32 and c["solver"] == "sor2sma_maf"][0]
    p = subprocess.run([exe, "32", "32", "32", "sor2sma_maf", str(g["itr_max"]), g["coef"]], cwd=str(tmp_path),
                       env=dict(os.environ, CZB_FILEOUT="1"), capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stdout + p.stderr
    txt = open(os.path.join(str(tmp_path), "sor2sma_maf.txt")).read().split("\n")
    assert txt[0] == "Itration      Residual"
    assert "Iter = %d" % g["itr"] in p.stdout
    hist = np.array([float(l.split(",")[1]) for l in txt[1:] if l.strip()])
    np.testing.assert_allclose(hist, [float(h) for h in g["history"]], rtol=2e-6)
    err = float(p.stdout.split("Error max = ")[1].split()[0])
    assert abs(err - float(g["errmax"])) < 1e-6
    flop = float(p.stdout.split("Flop count  : ")[1].split()[0])
    assert flop == pytest.approx(66.0 * 30 ** 3 * g["itr"], rel=1e-6)        # cz_maf.f90:330-333, both colours
    for name in ("p_00000.sph", "e_00000.sph"):
        assert os.path.getsize(os.path.join(str(tmp_path), name)) == (8 + 12 + 12 + 12 + 8 + 32 ** 3 * 4) + 6 * 8


@pytest.mark.parametrize("real,sz,color", [("f32", [8, 7, 34], 0), ("f32", [8, 7, 34], 1), ("f32", [6, 5, 131], -1),
                                           ("f64", [7, 6, 70], 1), ("f32", [5, 5, 300], 0)])
def test_line_kernel_with_real_threads_under_thread_sanitizer(real, sz, color, tmp_path):
    """k_pcr_lines_maf with one std::thread per CUDA thread and std::barrier for __syncthreads, built with
    -fsanitize=thread (tests/tools/emu/simt_line.cpp): no data race on the shared a / c / d rows (ThreadSanitizer
    would make the program exit non-zero) and the field still equals the oracle bit for bit."""
    import subprocess
    exe = emu_build.build_simt()[real]
    o, X, B, mesh, mp = _case(real, sz)
    M = o.mask()
    o.ac1 = o.dt(1.3)
    pn = ou.get_num_stage(o.idx[5] - o.idx[4] + 1)
    # oracle: one colour, or all lines in (j,i) order
    Xo = X.copy(); ro = C.c_double(0)
    if color >= 0:
        flop = C.c_double(0); om = o.creal(1.3); cpn = C.c_int(pn); ip = C.c_int(0); cc = C.c_int(color)
        tmp = np.zeros(sz[2] + 4, dtype=o.dt); w = [ou.ptr(a) for a in o.work1d]
        o.dll.pcr_rb_maf_(o._sz, o._idx, o.g(), C.byref(cpn), C.byref(ip), C.byref(cc), ou.ptr(Xo), ou.ptr(M), ou.ptr(B), *mp, *w,
                          C.byref(om), C.byref(ro), ou.ptr(tmp), C.byref(flop))
    else:
        o.sweep_maf("pcr_maf", Xo, B, ro, mesh)
    inp, outp = os.path.join(str(tmp_path), "in.bin"), os.path.join(str(tmp_path), "out.bin")
    with open(inp, "wb") as f:
        f.write(np.array(list(sz) + list(o.idx) + [pn, color], dtype=np.int32).tobytes())
        f.write(np.array([1.3], dtype=o.dt).tobytes())
        for a in (X, M, B, *mesh):
            f.write(np.ascontiguousarray(a, dtype=o.dt).tobytes())
    p = subprocess.run([exe, inp, outp], capture_output=True, text=True, timeout=900,
                       env=dict(os.environ, TSAN_OPTIONS="halt_on_error=0 exitcode=66"))
    assert p.returncode == 0, "ThreadSanitizer / driver: rc %d\n%s" % (p.returncode, p.stderr[-3000:])
    assert "WARNING: ThreadSanitizer" not in p.stderr
    raw = open(outp, "rb").read()
    Xs = np.frombuffer(raw[:X.nbytes], dtype=o.dt).reshape(X.shape)
    res = np.frombuffer(raw[X.nbytes:], dtype=np.float64)[0]
    assert np.array_equal(Xs, Xo), np.abs(Xs - Xo).max()
    assert res == pytest.approx(ro.value, rel=1e-12)


def test_emulated_kernels_under_address_sanitizer():
    """The kernel-level cases above (and two solves) once more on an AddressSanitizer build of the emulated library:
    every "device" allocation is a malloc block there, so an out-of-bounds load or store of a kernel -- the kind of bug
    that would be an illegal-address fault on the GPU -- aborts the child process."""
    import subprocess
    if os.environ.get("EMU_SANITIZE"):
        pytest.skip("already inside the sanitizer run")
    asan = subprocess.run(["gcc", "-print-file-name=libasan.so"], capture_output=True, text=True).stdout.strip()
    if not os.path.isabs(asan) or not os.path.exists(asan):
        pytest.skip("libasan not available")
    env = dict(os.environ, EMU_SANITIZE="address", LD_PRELOAD=asan, ASAN_OPTIONS="detect_leaks=0")
    sel = "point_kernels or line_kernel_equals or stretched_mesh or sweeps_and_operator or blas1 or (solve_matches and (pcr_maf or pbicgstab_maf-sor2sma_maf)) or tiny"
    p = subprocess.run([sys.executable, "-m", "pytest", os.path.join(HERE, "test_maf_kernels_emulated.py"),
                        os.path.join(HERE, "test_lib_emulated_plain.py"), os.path.join(HERE, "test_emulated_ragged.py"),   # "tiny": grids from 3 nodes per axis
                        "-x", "-q", "-p", "no:cacheprovider", "-k", sel],
                       env=env, capture_output=True, text=True, timeout=1500)
    assert p.returncode == 0 and " passed" in p.stdout, p.stdout[-3000:] + p.stderr[-3000:]
    assert "AddressSanitizer" not in p.stderr


def test_emulated_kernels_under_undefined_behaviour_sanitizer():
    """The ragged / tiny-grid cases and the plain per-call ABI cases on a -fsanitize=undefined build of the emulated
    library: no signed overflow, misaligned or out-of-range index arithmetic, bad shift or float-to-int conversion in
    the host code and the plain-CUDA kernels (UBSan reports "runtime error" on stderr and carries on)."""
    import subprocess
    if os.environ.get("EMU_SANITIZE"):
        pytest.skip("already inside the sanitizer run")
    ubsan = subprocess.run(["gcc", "-print-file-name=libubsan.so"], capture_output=True, text=True).stdout.strip()
    if not os.path.isabs(ubsan) or not os.path.exists(ubsan):
        pytest.skip("libubsan not available")
    env = dict(os.environ, EMU_SANITIZE="undefined", LD_PRELOAD=ubsan)
    p = subprocess.run([sys.executable, "-m", "pytest", os.path.join(HERE, "test_emulated_ragged.py"),
                        os.path.join(HERE, "test_lib_emulated_plain.py"), "-x", "-q", "-s", "-p", "no:cacheprovider"],
                       env=env, capture_output=True, text=True, timeout=1500)
    assert p.returncode == 0 and " passed" in p.stdout, p.stdout[-3000:] + p.stderr[-3000:]
    assert "runtime error" not in p.stdout + p.stderr, (p.stdout + p.stderr)[-3000:]
