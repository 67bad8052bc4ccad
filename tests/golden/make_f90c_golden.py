#!/usr/bin/env python
"""Writes the golden vectors that come from EXECUTING THE REFERENCE'S OWN FORTRAN (this container only: needs
/root/reference and `make -C oracle f90c`, i.e. oracle/f90_to_c.py's mechanical translation of
src/cz_f90/{cz_solver,cz_blas,cz_utility,cz_maf,obsolete}.f90 compiled to oracle/_ref/libczf90c_*.so and the
reference's unmodified driver linked to it, oracle/_ref/cz_f90c_*):

  tests/golden/f90c_signatures.json  argument lists of the 39 subroutines as DECLARED in the Fortran text
  tests/golden/f90c_digests.json     per (precision, routine, grid): sha256 of every array argument and the exact value
                                     (hex float) of every scalar argument after 4 consecutive calls on seeded inputs
                                     (tests/f90c_harness.py builds the inputs from the declarations)
  tests/golden/f90c_histories.json   residual histories + console lines (Iter/Res, Error max + location) of the
                                     reference driver over its translated Fortran, 37 solver configurations

    python tests/golden/make_f90c_golden.py
"""
import json
import os
import re
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import f90c_harness as H   # noqa: E402
import oracle_util as ou   # noqa: E402

SIZES = [[12, 7, 23], [7, 5, 31], [5, 5, 50], [6, 6, 64], [5, 6, 18], [4, 5, 130]]   # line lengths 21 29 48 62 16 128
CALLS = 4


def in_reference_domain(name, sz):
    """pcr_eda / pcr_eda_maf read a(k+2^(pn-1)) of an ALLOCATED a(kst-2^(pn-2):ked+2^(pn-2)) (cz_solver.f90:919-924,
    1005): inside the allocation only for n >= 3*2^(pn-2) -- true for the driver's 2^m-2 lines; elsewhere the
    reference reads heap memory it does not own (found by the F2C_BOUNDS build) and there is no defined answer."""
    n = sz[2] - 2
    pn = ou.get_num_stage(n)
    return not (name.startswith("pcr_eda") and n < 3 * 2 ** (pn - 2))


def run_routine(lib, sig, real, sz):
    vals = H.make_args(sig, real, sz, 0)
    for call in range(CALLS):
        if "color" in vals:
            vals["color"][0] = call % 2
        H.call(lib, sig, vals)
    return vals


def record(sig, vals):
    out = {}
    for nm in H.compared(sig):
        v = vals[nm]
        if v.size == 1 and v.dtype.kind == "f":
            out[nm] = float(v[0]).hex()
        elif v.size == 1:
            out[nm] = int(v[0])
        else:
            import hashlib
            out[nm] = hashlib.sha256(np.ascontiguousarray(v).tobytes()).hexdigest()
    return out


def driver_case(exe, real, n, solver, itr_max, coef, precond, env_extra=None):
    args = [os.path.join(ROOT, "oracle", "_ref", "%s_%s" % (exe, real)), str(n), str(n), str(n), solver, str(itr_max), coef]
    if precond:
        args.append(precond)
    env = dict(os.environ, OMP_NUM_THREADS="1", **(env_extra or {}))
    with tempfile.TemporaryDirectory() as td:
        out = subprocess.run(args, cwd=td, env=env, capture_output=True, text=True, check=False).stdout
        with open(os.path.join(td, solver + ".txt")) as f:
            hist = [line.split(",")[1].strip() for line in list(f)[1:]]
    console = [l.strip() for l in out.splitlines() if re.search(r"Iter =|Error max", l)]
    return {"real": real, "n": n, "solver": solver, "itr_max": itr_max, "coef": coef, "precond": precond,
            "history": hist, "console": console}


def main():
    sigs = H.signatures()
    with open(H.SIG_JSON, "w") as f:
        json.dump(sigs, f, indent=0)
    digests = []
    for real in ("f32", "f64"):
        lib = H.load_f90c(real)
        assert lib is not None, "run `make -C oracle f90c` first"
        for sig in sigs:
            if sig["name"] in H.SKIP:
                continue
            for sz in SIZES:
                if not in_reference_domain(sig["name"], sz):
                    continue
                if sig["name"].startswith("tdma") and sz != SIZES[0]:
                    continue
                digests.append({"real": real, "routine": sig["name"], "sz": sz, "calls": CALLS,
                                "out": record(sig, run_routine(lib, sig, real, sz))})
    with open(os.path.join(HERE, "f90c_digests.json"), "w") as f:
        json.dump(digests, f, indent=0)
    import importlib.util
    spec = importlib.util.spec_from_file_location("mg", os.path.join(HERE, "make_golden.py"))
    mg = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mg)
    hist = [driver_case("cz_f90c", *c) for c in mg.CASES]
    with open(os.path.join(HERE, "f90c_histories.json"), "w") as f:
        json.dump(hist, f, indent=0)
    print("%d signatures, %d digests, %d driver runs" % (len(sigs), len(digests), len(hist)))


if __name__ == "__main__":
    main()
