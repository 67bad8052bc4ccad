#!/usr/bin/env python
"""Writes tests/golden/ref_pcr_stages.json: outputs of the REFERENCE's own C++ PCR (src/cz_cpp/cz_pcr.cpp, CZ::pcr,
compiled unmodified by oracle/Makefile behind oracle/ref_pcr_harness.cpp -> oracle/_ref/ref_pcr_{f32,f64}) for a few
tridiagonal systems, as exact C99 hex floats.  Needs /root/reference (this container); the committed file lets the
check run anywhere.    python tests/golden/make_ref_pcr.py"""
import json
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_util as ou  # noqa: E402


def run_reference(real, n, stages, a, c, d):
    exe = ou.ref_binary("ref_pcr_%s" % real)
    assert exe, "build oracle/_ref first (make -C oracle all, needs /root/reference)"
    inp = "%d %d\n" % (n, stages) + "\n".join(" ".join(float(v).hex() for v in arr) for arr in (a, c, d)) + "\n"
    p = subprocess.run([exe], input=inp, capture_output=True, text=True, timeout=60)
    assert p.returncode == 0, p.stderr
    return [line.split() for line in p.stdout.strip().split("\n")]


def cases(real):
    dt = np.float64 if real == "f64" else np.float32
    rng = np.random.default_rng(20240607)
    for n in (7, 23, 62, 130):
        pn = int(np.floor(np.log2(n))) + 1          # cz.h:293-300
        for kind in ("laplace", "random"):
            if kind == "laplace":                   # the line coefficients of cz_solver.f90:556-564, r = 1/6
                r = dt(1.0) / dt(6.0)
                a = np.full(n, -r, dt); c = np.full(n, -r, dt)
            else:                                   # diagonally dominant, line dependent (as the stretched-grid lines)
                a = -rng.uniform(0.05, 0.45, n).astype(dt); c = -rng.uniform(0.05, 0.45, n).astype(dt)
            a[0] = 0; c[-1] = 0
            d = rng.uniform(-1, 1, n).astype(dt)
            for stages in (sorted({1, pn - 2, pn - 1, pn}) if n < 100 else [pn - 1]):     # 130: the solver's stage count only
                if stages >= 1 and (n < 100 or kind == "laplace"):
                    yield n, stages, kind, a, c, d


def main():
    out = {"made_by": "oracle/_ref/ref_pcr_* = /root/reference/src/cz_cpp/cz_pcr.cpp (unmodified) + oracle/ref_pcr_harness.cpp",
           "cases": []}
    for real in ("f64", "f32"):
        for n, stages, kind, a, c, d in cases(real):
            ra, rc, rd = run_reference(real, n, stages, a, c, d)
            out["cases"].append({"real": real, "n": n, "stages": stages, "kind": kind,
                                 "a": [float(v).hex() for v in a], "c": [float(v).hex() for v in c],
                                 "d": [float(v).hex() for v in d], "ref_a": ra, "ref_c": rc, "ref_d": rd})
    with open(os.path.join(HERE, "ref_pcr_stages.json"), "w") as f:
        json.dump(out, f, separators=(",", ":"))
    print("wrote %d cases" % len(out["cases"]))


if __name__ == "__main__":
    main()
