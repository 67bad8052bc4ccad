"""TEST INFRASTRUCTURE: calls ONE reference subroutine through two libraries that export the reference's Fortran
symbols -- the hand-written C++ port (oracle/liboracle_*.so) and the reference's own Fortran text translated
mechanically to C (oracle/_ref/libczf90c_*.so, oracle/f90_to_c.py) -- on identical seeded inputs, and returns every
argument after the call.  Argument lists, types and array ranks come from the DECLARATIONS of the Fortran source
(f90_to_c.signatures) or from the committed copy of them (tests/golden/f90c_signatures.json), never from a table
written by hand."""
import ctypes as C
import json
import os

import numpy as np

import oracle_util as ou

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_F90 = "/root/reference/src/cz_f90"
F90_FILES = ["cz_solver.f90", "cz_blas.f90", "cz_utility.f90", "cz_maf.f90", "obsolete.f90"]   # FCORE, src/cz_f90/CMakeLists.txt
SIG_JSON = os.path.join(ROOT, "tests", "golden", "f90c_signatures.json")
# work vectors the reference's line routines use as scratch (the port keeps its own) and outputs nobody reads
SCRATCH_1D = {"a", "c", "d", "a1", "c1", "d1", "aw", "cw", "dw", "tmp", "w"}
SKIP = {"fileout_t": "no-op off Aurora (#if _aurora_==1)", "tdma_mp": "not in the port (obsolete.f90, unused)",
        "tdma_1": "not in the port (obsolete.f90, unused)"}


def f90c_path(real, checked=False):
    return os.path.join(ROOT, "oracle", "_ref", "libczf90c_%s%s.so" % ("chk_" if checked else "", real))


def load_f90c(real, checked=False):
    p = f90c_path(real, checked)
    if not os.path.exists(p):
        return None
    return C.CDLL(p)


def signatures():
    if os.path.isdir(REF_F90):
        import sys
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import f90_to_c
        out = []
        for f in F90_FILES:
            out.extend(f90_to_c.signatures(os.path.join(REF_F90, f)))
        return out
    with open(SIG_JSON) as f:
        return json.load(f)


def make_args(sig, real, sz, color=0, seed=1, uniform_mesh=False):
    """name -> numpy array (scalars are 1-element arrays), in the argument order of `sig`."""
    dt = np.float32 if real == "f32" else np.float64
    rng = np.random.default_rng(seed)
    g = 2
    idx = [2, sz[0] - 1, 2, sz[1] - 1, 2, sz[2] - 1]
    n = idx[5] - idx[4] + 1
    pn = ou.get_num_stage(n)
    shape = (sz[1] + 2 * g, sz[0] + 2 * g, sz[2] + 2 * g)        # numpy C order of Fortran (k, i, j)
    name = sig["name"]
    vals = {}
    for a in sig["args"]:
        nm, ty, dims = a["name"], a["type"], a["dims"]
        rank = len(dims) if dims else 0
        if ty == "int":
            if rank == 1:
                v = {"sz": sz, "idx": idx, "nid": [-1, 3, -1, -1, 0, -1], "loc": [7, 7, 7]}[nm]
            else:
                v = [{"g": g, "pn": pn, "s": 2 ** (pn - 2), "color": color, "ofst": 0, "nx": n}[nm]]
            vals[nm] = np.array(v, dtype=np.int32)
        elif ty == "double":
            vals[nm] = np.array([{"res": 0.5, "flop": 0.25, "d": 0.125}[nm]], dtype=np.float64)
        elif ty == "real" and rank == 3:
            if nm == "msk":
                m = np.zeros(shape, dtype=dt)
                m[idx[2] + 1:idx[3] + 2, idx[0] + 1:idx[1] + 2, idx[4] + 1:idx[5] + 2] = 1     # = imask_k
                vals[nm] = m
            elif nm == "pvt":
                vals[nm] = rng.uniform(0.5, 1.5, size=shape).astype(dt)
            else:
                vals[nm] = rng.uniform(-1.0, 1.0, size=shape).astype(dt)
        elif ty == "real" and rank == 1:
            if name.startswith("tdma"):
                v = rng.uniform(-1.0, 1.0, size=n).astype(dt)
                if nm == "b":
                    v = (v + dt(4.0)).astype(dt)
                vals[nm] = v
            elif nm in ("x", "y", "z", "xx", "yy", "zz"):
                ax = {"x": 0, "y": 1, "z": 2}[nm[0]]
                i = np.arange(sz[ax] + 2 * g)
                h = dt(1.0) / dt(sz[2] - 1)
                v = ((i - 1).astype(dt) * h).astype(dt)
                if not uniform_mesh:
                    v = (v + dt(0.3) * h * np.sin(dt(0.37 + ax) * i.astype(dt))).astype(dt)
                vals[nm] = np.ascontiguousarray(v)
            elif nm == "cf":
                vals[nm] = np.array([1.03, 0.97, 1.01, 0.99, 1.05, 0.95, 6.1], dtype=dt)
            elif nm == "org":
                vals[nm] = np.array([0.0, 0.03125, 0.0625], dtype=dt)
            elif nm in SCRATCH_1D:
                # zero, as the reference driver hands them over (czAllocR, cz.h:236-253; SA/SC/SD are cleared again at
                # the top of LSOR_PCR_*ESA, cz_Poisson.cpp:646-661) -- pcr_j_esa relies on it (`a(kst) = 0.0` is
                # commented out, cz_solver.f90:1536).  Generous size: the 4x4 variants write d1(k+3*2^(pn-2)) beyond the
                # declared extent when n is far from 2^pn (oracle/f90_to_c.py bounds check).
                vals[nm] = np.zeros(8 * (sz[2] + 2 * g) + 64, dtype=dt)
            else:
                raise KeyError("%s: rank-1 REAL argument %s" % (name, nm))
        elif ty == "real" and rank == 0:
            v = {"omg": 1.3, "dh": 1.0 / (sz[2] - 1), "a": -0.37, "b": 0.81, "beta": 0.23, "r": 0.0}[nm]
            vals[nm] = np.array([v], dtype=dt)
        else:
            raise KeyError("%s: argument %s of type %s rank %d" % (name, nm, ty, rank))
    return vals


def bounds_report(lib):
    """violations recorded by a libczf90c_chk_*.so since the last reset: [(file:line, array, dim, index, lo, hi, count)]"""
    buf = C.create_string_buffer(1 << 16)
    lib.f2c_bounds_report(buf, len(buf))
    out = []
    for line in buf.value.decode().splitlines():
        loc, a, d, i, lo, hi, cnt = line.split()
        out.append((loc, a, int(d), int(i), int(lo), int(hi), int(cnt)))
    return out


def call(lib, sig, vals):
    fn = getattr(lib, sig["name"] + "_")
    fn.restype = None
    fn.argtypes = [C.c_void_p] * len(sig["args"])
    fn(*[C.c_void_p(vals[a["name"]].ctypes.data) for a in sig["args"]])


def compared(sig):
    """arguments whose value after the call is part of the routine's result"""
    return [a["name"] for a in sig["args"]
            if not (a["type"] == "real" and a["dims"] and len(a["dims"]) == 1 and a["name"] in SCRATCH_1D
                    and not sig["name"].startswith("tdma"))]


def check_library_against_fortran_vectors(lib, real, sizes, digests, sigs, run_routine, record):
    """A library that exports the reference's symbols (the product, or its CPU-emulated build) against the committed
    outputs of the reference's translated Fortran: every array argument bit for bit; double-accumulated sums (the
    library sums REAL products in double, the 1-thread reference in REAL) and the analytic field within a tolerance.
    Not compared: `src` of pcr_j_esa (a work field the library does not fill), tdma_* (not on the path)."""
    sum_tol = 3e-6 if real == "f32" else 1e-13
    n = 0
    for case in digests:
        name = case["routine"]
        if case["real"] != real or case["sz"] not in sizes or name.startswith("tdma"):
            continue
        sig = sigs[name]
        vals = run_routine(lib, sig, real, case["sz"])
        got = record(sig, vals)
        for k, v in got.items():
            want = case["out"][k]
            if k in ("res", "r"):
                assert abs(float.fromhex(v) - float.fromhex(want)) <= sum_tol * max(abs(float.fromhex(want)), 1e-30), (name, case["sz"], k)
            elif k == "flop":
                assert abs(float.fromhex(v) - float.fromhex(want)) <= 1e-12 * abs(float.fromhex(want)), (name, k)
            elif k == "src" and name == "pcr_j_esa":
                continue
            elif k == "e" and name == "exact_t":
                continue      # device sin / sinh: <= 8 ulp, tests/test_gpu_kernels.py::test_bc_exact_err
            else:
                assert v == want, (name, case["sz"], k)
        n += 1
    return n
