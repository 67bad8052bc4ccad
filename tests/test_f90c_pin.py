"""Pins the oracle port (oracle/cz_f90_port.cpp) to OUTPUTS OF THE REFERENCE'S OWN FORTRAN.

No Fortran compiler exists in this image or on the GPU box, so the reference's kernels (src/cz_f90/*.f90) are executed
through a mechanical Fortran -> C translation of their unmodified text (oracle/f90_to_c.py; `make -C oracle f90c` builds
oracle/_ref/libczf90c_*.so and links the reference's unmodified C++ driver to it -> oracle/_ref/cz_f90c_*).

* everywhere (no /root/reference, no oracle/_ref needed): the port reproduces the committed golden vectors that the
  translated Fortran wrote -- every array argument of every routine bit for bit, residual sums and dot products bit
  for bit in the 1-thread summation order (tests/golden/f90c_digests.json), and the driver's residual histories digit
  for digit (tests/golden/f90c_histories.json);
* where oracle/_ref/libczf90c_*.so exists (this container; it also travels to the GPU box): the same comparison live,
  on more grids, plus the bounds-check build's findings;
* where /root/reference exists: the committed vectors and signatures are regenerated and must be unchanged.
"""
import ctypes as C
import json
import os
import subprocess
import sys

import numpy as np
import pytest

import f90c_harness as H
import oracle_util as ou

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import make_f90c_golden as MG   # noqa: E402

DIGESTS = json.load(open(os.path.join(HERE, "golden", "f90c_digests.json")))
HISTORIES = json.load(open(os.path.join(HERE, "golden", "f90c_histories.json")))
SIGS = {s["name"]: s for s in json.load(open(H.SIG_JSON))}
ROUTINES = [n for n in SIGS if n not in H.SKIP]
HAVE_F90C = all(os.path.exists(H.f90c_path(r, c)) for r in ("f32", "f64") for c in (False, True))
HAVE_REF = os.path.isdir(H.REF_F90)
needs_f90c = pytest.mark.skipif(not HAVE_F90C, reason="oracle/_ref/libczf90c_*.so not built (needs /root/reference once)")


@pytest.fixture(autouse=True)
def _restore_sum_mode():
    yield
    for real in ("f32", "f64"):
        ou.load(real).oracle_set_sum_mode(0)


def _port(real, real_order):
    lib = ou.load(real)
    lib.oracle_set_sum_mode(1 if real_order else 0)
    return lib


def test_signature_table_covers_the_fortran_core():
    """39 subroutines = the five files of the reference's FCORE library (src/cz_f90/CMakeLists.txt); the 35 symbols of
    cz_Ffunc.h that the path uses are all among them, with the argument counts cubez_b200/ffi.py binds."""
    import cubez_b200.ffi as ffi
    assert len(SIGS) == 39
    for sym, nargs in ffi.REFERENCE_SYMBOLS.items():
        assert sym[:-1] in SIGS, sym
        assert len(SIGS[sym[:-1]]["args"]) == nargs, sym
    if HAVE_REF:
        assert json.loads(json.dumps(H.signatures())) == list(SIGS.values()), "regenerate: tests/golden/make_f90c_golden.py"


@pytest.mark.parametrize("real", ["f32", "f64"])
def test_port_reproduces_the_reference_fortran_golden_vectors(real):
    """Runs without the reference: 200 (routine, grid) cases per precision.  Summation in the reference's 1-thread
    order (ORACLE_SUM=real) makes residuals and dot products comparable bit for bit as well."""
    port = _port(real, True)
    n = 0
    for case in DIGESTS:
        if case["real"] != real:
            continue
        sig = SIGS[case["routine"]]
        got = MG.record(sig, MG.run_routine(port, sig, real, case["sz"]))
        assert got == case["out"], (case["routine"], case["sz"], [k for k in got if got[k] != case["out"][k]])
        n += 1
    assert n >= 190


@pytest.mark.parametrize("real", ["f32", "f64"])
def test_port_default_summation_differs_in_the_sums_only(real):
    """ORACLE_SUM=double (what the GPU tests compare with): identical fields, sums within the accumulation error."""
    port = _port(real, False)
    tol = 3e-6 if real == "f32" else 1e-13
    for case in DIGESTS:
        if case["real"] != real or case["sz"] != MG.SIZES[0]:
            continue
        sig = SIGS[case["routine"]]
        got = MG.record(sig, MG.run_routine(port, sig, real, case["sz"]))
        for k, v in got.items():
            if k in ("res", "r"):
                assert float.fromhex(v) == pytest.approx(float.fromhex(case["out"][k]), rel=tol), (case["routine"], k)
            else:
                assert v == case["out"][k], (case["routine"], k)


def _oracle_history(c):
    o = ou.OracleCZ([c["n"]] * 3, c["real"])
    P, RHS = o.setup_laplace()
    if c["solver"].startswith("pbicgstab"):
        itr, res, hist = o.pbicgstab(P, RHS, c["itr_max"], float(c["coef"]), c["precond"], maf=c["solver"].endswith("_maf"))
    else:
        itr, res, hist = o.stationary(c["solver"], P, RHS, c["itr_max"], float(c["coef"]))
    return ["%13.6e" % h for h in hist]


@pytest.mark.parametrize("real", ["f32", "f64"])
def test_solver_loops_over_the_port_reproduce_the_reference_fortran_histories(real):
    """The reference's unmodified driver over its own (translated) Fortran wrote these histories; the Python mirror of
    the driver loops (tests/oracle_util.py) over the port prints the same digits, iteration for iteration."""
    _port(real, True)
    n = 0
    for c in HISTORIES:
        if c["real"] != real or c["n"] > 32:
            continue
        assert _oracle_history(c) == ["%13s" % h for h in c["history"]], (c["solver"], c["precond"])
        n += 1
    assert n >= 8


# ------------------------------------------------------------------------------------------------------------
# live comparison against the translated Fortran
# ------------------------------------------------------------------------------------------------------------

LIVE_SIZES = [[12, 7, 23], [9, 10, 34], [6, 5, 66], [5, 6, 18], [7, 5, 31], [5, 5, 50], [4, 4, 258], [130, 6, 300], [5, 40, 1024]]


@needs_f90c
@pytest.mark.parametrize("real", ["f32", "f64"])
@pytest.mark.parametrize("routine", ROUTINES)
def test_port_equals_translated_fortran_live(real, routine):
    sig = SIGS[routine]
    port = _port(real, True)
    ref = H.load_f90c(real)
    for sz in LIVE_SIZES:
        if not MG.in_reference_domain(routine, sz):
            continue
        a = MG.run_routine(port, sig, real, sz)
        b = MG.run_routine(ref, sig, real, sz)
        for nm in H.compared(sig):
            assert np.array_equal(a[nm].view(np.uint8), b[nm].view(np.uint8)), (routine, sz, nm)


@needs_f90c
def test_bounds_check_build_same_results_and_known_findings():
    """The F2C_BOUNDS build (index check on every array reference, the access itself unchanged) gives the same bits.
    On the driver's own line lengths (2^m - 2: 62 here; 14..510 checked when the golden vectors were made) NO access of
    any routine leaves its declared bounds.  On other lengths (32 here) the zero-padded line routines (pcr_eda, pcr_esa,
    pcr_rb_esa, pcr_j_esa and their _maf twins) read a, d(k + 2^(pn-1)) beyond idx(5)+s and store d1 / dw beyond
    sz(3)+2 -- past the driver's work arrays too (cz_Evaluate.cpp:255-263, 300-311); the port's zero padding is the
    defined reading of those accesses, and pcr_eda's are excluded (MG.in_reference_domain) because it ALLOCATES a, c, d."""
    chk = H.load_f90c("f32", True)
    ref = H.load_f90c("f32", False)
    found = {}
    for routine in ROUTINES:
        sig = SIGS[routine]
        for sz in ([6, 6, 64], [5, 5, 34]):     # 62- and 32-cell lines
            if not MG.in_reference_domain(routine, sz):
                continue
            chk.f2c_bounds_reset()
            a = MG.run_routine(chk, sig, "f32", sz)
            b = MG.run_routine(ref, sig, "f32", sz)
            for nm in H.compared(sig):
                assert np.array_equal(a[nm].view(np.uint8), b[nm].view(np.uint8)), (routine, sz, nm)
            for (loc, arr, d, i, lo, hi, cnt) in H.bounds_report(chk):
                found.setdefault((routine, sz[2] - 2), []).append((loc, arr, i, lo, hi))
    assert not {k: v for k, v in found.items() if k[1] == 62}
    assert sorted({k[0] for k in found}) == ["pcr_esa", "pcr_esa_maf", "pcr_j_esa", "pcr_rb_esa", "pcr_rb_esa_maf"]
    assert {arr for v in found.values() for (_, arr, _, _, _) in v} == {"a", "d", "d1", "dw"}
    point = [r for r in ROUTINES if not r.startswith(("pcr", "tdma"))]
    assert not [k for k in found if k[0] in point], "point kernels / BLAS-1 stay inside their declared bounds"


def _run_driver(exe, c, env_extra):
    return MG.driver_case(exe, c["real"], c["n"], c["solver"], c["itr_max"], c["coef"], c["precond"], env_extra)


@pytest.mark.skipif(not (ou.ref_binary("cz_f90c_f32") and ou.ref_binary("cz_ref_f32")), reason="oracle/_ref drivers not built")
def test_reference_driver_prints_the_same_over_port_and_over_its_own_fortran():
    """Same unmodified driver binary sources, two kernel libraries: the history files and the console lines
    (Iter / Res / Error max and its location) are text-identical.  32^3 cases (the 64^3 ones take a minute)."""
    n = 0
    for c in HISTORIES:
        if c["n"] > 32:
            continue
        a = _run_driver("cz_ref", c, {"ORACLE_SUM": "real"})
        b = _run_driver("cz_f90c", c, None)
        assert a["history"] == b["history"] == c["history"], (c["solver"], c["precond"])
        assert a["console"] == b["console"] == c["console"], (c["solver"], c["precond"])
        n += 1
    assert n >= 30


@pytest.mark.skipif(not HAVE_REF, reason="needs /root/reference")
def test_translator_is_deterministic_and_rejects_what_it_does_not_know(tmp_path):
    sys.path.insert(0, os.path.join(ou.ROOT, "oracle"))
    import f90_to_c
    src = os.path.join(H.REF_F90, "cz_solver.f90")
    a, names = f90_to_c.translate_file(src, "f32")
    b, _ = f90_to_c.translate_file(src, "f32")
    assert a == b and "psor2sma_core" in names
    # Fortran's grammar, not C's: -e*ap*a(kl) negates the product; REAL literals are single precision in f32
    assert "(-((v_e * v_ap) * v_a[" in a and "(1.0f / 6.0f)" in a
    assert "(18.0f * ((double)" in a and "(18.0e0 * 0.5e0)" in a     # 18.0 is REAL, 18.0d0*0.5d0 double
    bad = tmp_path / "bad.f90"
    for text in ("subroutine s(x)\nimplicit none\nreal :: x\nx = y\nend subroutine s\n",                  # undeclared name
                 "subroutine s(x)\nimplicit none\nreal :: x\nx = x**2.0\nend subroutine s\n",             # real power
                 "subroutine s(x)\nimplicit none\nreal :: x\ncall t(x)\nend subroutine s\n",              # call
                 "subroutine s(x)\nimplicit none\nreal :: x\nwhere (x > 0) x = 1\nend subroutine s\n"):   # where
        bad.write_text(text)
        with pytest.raises((SyntaxError, NameError, TypeError)):
            f90_to_c.translate_file(str(bad), "f32")


OMP_SO = {r: os.path.join(ou.ROOT, "oracle", "_ref", "libczf90c_omp_%s.so" % r) for r in ("f32", "f64")}


@pytest.mark.skipif(not (HAVE_F90C and all(os.path.exists(p) for p in OMP_SO.values())), reason="oracle/_ref/libczf90c_omp_*.so not built")
@pytest.mark.parametrize("real", ["f32", "f64"])
def test_openmp_translation_gives_the_serial_fields(real):
    """`f90_to_c.py --openmp` maps the reference's `!$OMP` directives to C pragmas (the CPU arm bench.py times).  With 4
    threads every array of the race-free routines equals the serial translation bit for bit; sums differ by the
    reduction order only.  Not compared at 4 threads, because the REFERENCE races there: `psor`, `psor_maf`, `pcr`,
    `pcr_eda`, `pcr_esa` and their `_maf` twins (in-place Gauss-Seidel between cells / lines inside a parallel loop),
    `err_t` (unsynchronised max), and `jacobi`, `jacobi_maf`, whose `END DO NOWAIT` lets a thread start the copy-back
    `p = wk2` while its neighbours still read `p` (cz_solver.f90:355-372) -- which is why the oracle states the
    1-thread order.  With 1 thread everything equals the serial translation."""
    gomp = C.CDLL("libgomp.so.1")
    ser, omp = H.load_f90c(real), C.CDLL(OMP_SO[real])
    report = dict(l.strip().split(": ", 1) for l in open(os.path.join(ou.ROOT, "oracle", "_ref", "f90c_omp", real, "openmp_report.txt")))
    mapped = [k for k, v in report.items() if v == "openmp"]
    # every subroutine that has directives is mapped (tdma_mp's are malformed: no END PARALLEL DO; not on the path)
    assert sorted(k for k, v in report.items() if v != "openmp") == ["fileout_t", "tdma_0", "tdma_1", "tdma_mp", "tdma_p"]
    racy_mt = ("psor", "psor_maf", "err_t", "jacobi", "jacobi_maf", "pcr", "pcr_eda", "pcr_esa", "pcr_maf", "pcr_eda_maf", "pcr_esa_maf")
    prev = gomp.omp_get_max_threads()
    try:
        for threads in (4, 1):
            gomp.omp_set_num_threads(threads)
            racy = () if threads == 1 else racy_mt
            for name in mapped:
                if name in racy or name in H.SKIP:
                    continue
                for sz in ([12, 7, 23], [20, 33, 64]):
                    if not MG.in_reference_domain(name, sz):
                        continue
                    a = MG.run_routine(ser, SIGS[name], real, sz)
                    b = MG.run_routine(omp, SIGS[name], real, sz)
                    for nm in H.compared(SIGS[name]):
                        if nm in ("res", "r"):      # (a reduction starts from 0 and is added to the incoming value at the end)
                            assert b[nm][0] == pytest.approx(a[nm][0], rel=5e-5 if real == "f32" else 1e-12), (name, nm)   # REAL accumulators, other order
                        else:
                            assert np.array_equal(a[nm].view(np.uint8), b[nm].view(np.uint8)), (name, sz, nm, threads)
    finally:
        gomp.omp_set_num_threads(prev)


def test_translator_implements_fortran_semantics_on_a_known_answer_program(tmp_path):
    """The translator on a small Fortran program whose answers follow from the Fortran standard alone (nothing of CubeZ):
    operator precedence and associativity, integer division, integer power, mixed-type promotion, REAL vs double
    literals, intrinsics by type, array lower bounds and column-major order, do-loop trip counts (bounds evaluated once,
    zero-trip, negative stride, value after the loop), cycle / exit, one-line if, else if, allocate with bounds,
    preprocessor branches, continuation lines and `;`."""
    sys.path.insert(0, os.path.join(ou.ROOT, "oracle"))
    import f90_to_c
    src = tmp_path / "known.f90"
    src.write_text("""
subroutine known (n, iv, rv, dv, a, b)
implicit none
integer                       :: n, i, j, k, m, cnt
integer, dimension(20)        :: iv
real, dimension(20)           :: rv
double precision, dimension(20) :: dv
real, dimension(-1:n, 0:2)    :: a
real, dimension(:), allocatable :: w
real                          :: x, y
double precision              :: b

iv(1) = 7/2*2            ! (7/2)*2 = 6
iv(2) = -2**2            ! -(2**2) = -4
iv(3) = 2**3**2          ! 2**(3**2) = 512
iv(4) = mod(-7, 2)       ! sign of the dividend: -1
iv(5) = 2 + 3 * 4 - 10 / 4   ! 2 + 12 - 2 = 12
iv(6) = max(3, 9, 5) - min(4, 2, 8)
iv(7) = abs(-5) + 2**(n-8)   ! n = 10: 5 + 4
iv(8) = 17 - 5 - 3       ! left to right: 9

rv(1) = 7/2*2.0          ! integer division first: 6.0
rv(2) = real(7)/2        ! 3.5
rv(3) = 1.0/3.0          ! REAL division
rv(4) = -1.5*2.0*3.0     ! -((1.5*2.0)*3.0)
rv(5) = 0.1              ! REAL literal: float(0.1)
rv(6) = sqrt(2.0) * &
        sqrt(2.0)        ! REAL sqrt, REAL product
rv(7) = abs(-2.5) + max(1.5, 0.5)
rv(8) = 16777216.0 + 1.0 ! REAL: 2^24 + 1 rounds to 2^24
x = 3.0 ; y = 4.0
rv(9) = x - y - 1.0      ! (3-4)-1 = -2

dv(1) = 1.0/3.0          ! REAL quotient converted on assignment
dv(2) = 1.0d0/3.0d0      ! double
dv(3) = dble(1.0)/3.0    ! REAL 3.0 promoted: double quotient
dv(4) = 0.1d0
dv(5) = 16777216.0 + 1.0d0   ! promoted to double: exact
dv(6) = real(0.1d0, kind=8) + real(n)  ! stays double; real(n) REAL
b = b + 18.0d0*0.5d0 * dble(n)

! column-major with declared lower bounds: a(i,j) at (i+1) + (n+2)*j
do j = 0, 2
do i = -1, n
a(i, j) = real(100*j + i)
end do
end do

! trip count fixed at entry; value of the variable after the loop
m = 3
cnt = 0
do i = 1, m
m = 10
cnt = cnt + 1
end do
iv(9) = cnt              ! 3
iv(10) = i               ! 4
cnt = 0
do i = 5, 1              ! zero trip
cnt = cnt + 1
end do
iv(11) = cnt + 10*i      ! 0 + 50
cnt = 0
do i = 10, 1, -3         ! 10 7 4 1
cnt = cnt + i
end do
iv(12) = cnt             ! 22
iv(13) = i               ! -2
cnt = 0
do i = 1, 10
if (mod(i, 2) == 0) cycle
if (i > 7) exit
cnt = cnt + i            ! 1 + 3 + 5 + 7
end do
iv(14) = cnt

if (n < 5) then
iv(15) = 1
else if (n .ge. 10 .and. .not. (n == 11)) then
iv(15) = 2
else
iv(15) = 3
endif

k = 4
allocate(w(-k:k))
do i = -k, k
w(i) = real(i*i)
end do
rv(10) = w(-4) + w(0) + w(3)   ! 16 + 0 + 9
deallocate(w)

#ifdef _NOT_DEFINED
iv(16) = 111
#else
iv(16) = 222
#endif
#ifndef _NOT_DEFINED
iv(17) = 333
#endif
return
end subroutine known
""")
    text, names = f90_to_c.translate_file(str(src), "f32")
    assert names == ["known"]
    (tmp_path / "known.c").write_text(text)
    so = tmp_path / "known.so"
    subprocess.check_call(["gcc", "-O2", "-ffp-contract=off", "-fno-builtin", "-fPIC", "-shared", "-w", "-o", str(so),
                           str(tmp_path / "known.c"), "-lm"])
    lib = C.CDLL(str(so))
    n = np.array([10], dtype=np.int32)
    iv = np.zeros(20, dtype=np.int32); rv = np.zeros(20, dtype=np.float32); dv = np.zeros(20, dtype=np.float64)
    a = np.zeros((3, 12), dtype=np.float32)          # Fortran a(-1:10, 0:2), first index contiguous
    b = np.array([1.0], dtype=np.float64)
    lib.known_(*[C.c_void_p(v.ctypes.data) for v in (n, iv, rv, dv, a, b)])
    f = np.float32
    assert list(iv[:17]) == [6, -4, 512, -1, 12, 7, 9, 9, 3, 4, 50, 22, -2, 16, 2, 222, 333]
    want = [f(6.0), f(3.5), f(1.0) / f(3.0), f(-9.0), f(0.1), np.sqrt(f(2.0)) * np.sqrt(f(2.0)), f(4.0), f(16777216.0), f(-2.0), f(25.0)]
    assert [x.tobytes() for x in rv[:10]] == [np.float32(x).tobytes() for x in want]
    assert dv[0] == float(f(1.0) / f(3.0)) and dv[1] == 1.0 / 3.0 and dv[2] == 1.0 / 3.0 and dv[3] == 0.1
    assert dv[4] == 16777217.0 and dv[5] == 0.1 + 10.0
    assert b[0] == 1.0 + 9.0 * 10.0
    assert np.array_equal(a, np.array([[100 * j + i for i in range(-1, 11)] for j in range(3)], dtype=np.float32))
    # double-precision build (-fdefault-real-8 -fdefault-double-8): REAL literals and variables become double
    text64, _ = f90_to_c.translate_file(str(src), "f64")
    (tmp_path / "known64.c").write_text(text64)
    subprocess.check_call(["gcc", "-O2", "-ffp-contract=off", "-fno-builtin", "-fPIC", "-shared", "-w", "-o", str(tmp_path / "known64.so"),
                           str(tmp_path / "known64.c"), "-lm"])
    lib64 = C.CDLL(str(tmp_path / "known64.so"))
    rv64 = np.zeros(20); a64 = np.zeros((3, 12)); iv[:] = 0
    lib64.known_(*[C.c_void_p(v.ctypes.data) for v in (n, iv, rv64, dv, a64, b)])
    assert rv64[2] == 1.0 / 3.0 and rv64[4] == 0.1 and rv64[7] == 16777217.0 and dv[0] == 1.0 / 3.0


def test_reference_program_converges_like_the_oracle_at_the_baseline_configs():
    """BASELINE.json configs[0..3] to convergence, twice: by the reference PROGRAM as far as this image can build it (its
    unmodified driver over its own Fortran + OpenMP directives translated to C, tests/golden/baseline_configs_f90c.json)
    and by the same driver over the oracle port in its deterministic double-sum mode (baseline_configs.json, what the
    GPU tests and bench.py compare with).  Stationary solvers: the same iteration count and -- the fields being
    bit-identical (same Error max and location), only the REAL residual reduction over 16.6 M cells differing --
    histories equal to 2e-3 (float: a REAL accumulator over millions of squares drops low bits) / 1e-9 (double).
    PBiCGstab: its scalars come out of REAL dot products over 1.3e8 cells whose summation order depends on the team
    size, and the 512^3 run creeps towards eps on a plateau (residual 1.15e-5, 1.15e-5, 1.45e-5, 1.06e-5, 9.98e-6 over
    the last five iterations of the oracle's run): the reference program itself stops at 67 with 8 threads where the
    double-accumulating oracle (and the GPU) stop at 65 -- a spread of the REFERENCE across reduction orders, which is
    what "+- 1 iteration" has to be read against; asserted here as <= 3 with histories equal to 1e-2 over the first 18."""
    ref = json.load(open(os.path.join(HERE, "golden", "baseline_configs_f90c.json")))
    port = json.load(open(os.path.join(HERE, "golden", "baseline_configs.json")))
    assert {"f32 64 jacobi 0.8 None", "f32 256 sor2sma 1.95 None", "f64 256 sor2sma 1.95 None", "f32 256 pcr_rb 1.9 None"} <= set(ref)
    for k, r in ref.items():
        p = port[k]
        assert r["itr"] > 0 and r["itr"] < r["itr_max"], k
        if r["solver"] == "pbicgstab":
            if r["threads"] > 1:
                assert abs(r["itr"] - p["itr"]) <= 3, k
            else:
                # one thread (Jacobi preconditioner: the reference's OpenMP Jacobi races): ONE REAL accumulator over
                # 1.3e8 products per dot product -- 0.3 % off in the very first residual, 12 % by iteration 20 as
                # BiCGstab's scalars drift -- and the reference program stops at 73 where the double-accumulating
                # oracle needs 83: at 512^3 in float the reference's count is a property of its summation order
                assert r["itr"] < p["itr"] and abs(r["itr"] - p["itr"]) <= 0.15 * p["itr"], k
                continue
            common = [i for i in r["history_at"] if i in p["history_at"] and int(i) < 18]
            tol = 1e-2      # the very first residual already differs by 6e-4: one REAL dot product over 1.3e8 cells
        else:
            common = [i for i in r["history_at"] if i in p["history_at"]]
            tol = 2e-3 if r["real"] == "f32" else 1e-9
            # Where the residual CREEPS to eps (the relaxation-factor sweep of config 2 at w <= 1.5: it falls by 2e-4 ..
            # 5e-4 per iteration at the end), the 3e-4 by which the reference's REAL sum over 16.6 M squares sits below
            # the double-accumulated one is worth one or two iterations: 5735 / 5640 / 4234 against the oracle's 5737 /
            # 5641 / 4235 at w = 1.0 / 1.1 / 1.5.  Window = that offset / the decay per iteration, rounded up; 0 (the
            # same iteration) wherever the residual falls by more than 1e-3 per iteration (w >= 1.8: 2214, 1301, 734, 443).
            tail = sorted(int(i) for i in p["history_at"])[-2:]
            decay = 1.0 - float(p["history_at"][str(tail[1])]) / float(p["history_at"][str(tail[0])])
            last = str(max(int(i) for i in common))
            offset = abs(1.0 - float(r["history_at"][last]) / float(p["history_at"][last]))
            slack = int(np.ceil(offset / decay)) if decay < 1e-3 else 0
            assert slack <= 2 and abs(r["itr"] - p["itr"]) <= slack, (k, r["itr"], p["itr"], slack)
            if r["itr"] == p["itr"]:
                assert r["errmax"] == p["errmax"] and r["errloc"] == p["errloc"], k
            else:
                assert float(r["errmax"]) == pytest.approx(float(p["errmax"]), rel=1e-3), k
        assert len(common) >= 15, k
        for i in common:
            assert float(r["history_at"][i]) == pytest.approx(float(p["history_at"][i]), rel=tol), (k, i)
