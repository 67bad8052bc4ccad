#!/usr/bin/env python
"""Generates tests/golden/baseline_configs.json: BASELINE.json configs[0..3] solved TO CONVERGENCE
(eps = 1e-5, cz.h:162) by oracle/_ref/cz_ref_{f32,f64} -- the reference's UNMODIFIED C++ driver
(src/main.cpp + src/cz_cpp/*.cpp compiled where they lie under /root/reference) linked to the C++
restatement of its Fortran kernels (oracle/cz_f90_port.cpp), ORACLE_SUM=double (deterministic for
any thread count; only deterministic solvers are listed).  The solver-loop seconds are the
reference's own "JACOBI"-section timing region (cz_Evaluate.cpp:419-421) measured by wall clock
around the run minus nothing -- they are recorded for orientation only (this container, 8 cores);
bench.py times the CPU arm on the GPU box's own cores.

    python tests/golden/make_golden_baseline.py [case-index ...]

Needs /root/reference-built oracle/_ref (this container only); the JSON is committed so that the
GPU box and the CPU test-suite never read the reference.  History is stored subsampled (first 20,
every 50th, last 5) next to a SHA-256 of the full text history.
"""
import hashlib
import json
import os
import re
import subprocess
import sys
import tempfile
import time

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
OUT = os.path.join(HERE, "baseline_configs.json")

CASES = [
    # (real, N, solver, itr_max, coef, precond)
    ("f32", 64, "jacobi", 4000, "0.8", None),                  # configs[0]
    ("f32", 256, "sor2sma", 20000, "1.95", None),              # configs[1]
    ("f32", 512, "pbicgstab", 400, "1.1", "sor2sma"),          # configs[2] float
    ("f64", 512, "pbicgstab", 400, "1.1", "sor2sma"),          # configs[2] double
    ("f32", 512, "pcr_rb", 8000, "1.9", None),                 # configs[3]
    ("f64", 256, "sor2sma", 20000, "1.95", None),
    ("f32", 512, "pbicgstab", 400, "0.8", "jacobi"),
    ("f32", 256, "pcr_rb", 8000, "1.9", None),
    # configs[1]'s relaxation-factor sweep (SURVEY section 8(d) cfg 2: w in {1.0, 1.1, 1.5, 1.8, 1.9, 1.95, 1.97}; 1.95 is above)
    ("f32", 256, "sor2sma", 100000, "1.0", None),
    ("f32", 256, "sor2sma", 100000, "1.1", None),
    ("f32", 256, "sor2sma", 100000, "1.5", None),
    ("f32", 256, "sor2sma", 100000, "1.8", None),
    ("f32", 256, "sor2sma", 100000, "1.9", None),
    ("f32", 256, "sor2sma", 100000, "1.97", None),
]


def key(c):
    return "%s %d %s %s %s" % (c[0], c[1], c[2], c[4], c[5])


def run_case(real, n, solver, itr_max, coef, precond):
    exe = os.path.join(ROOT, "oracle", "_ref", "cz_ref_%s" % real)
    args = [exe, str(n), str(n), str(n), solver, str(itr_max), coef] + ([precond] if precond else [])
    env = dict(os.environ, OMP_NUM_THREADS=str(os.cpu_count()), OMP_PROC_BIND="close", ORACLE_SUM="double")
    with tempfile.TemporaryDirectory() as td:
        t0 = time.time()
        out = subprocess.run(args, cwd=td, env=env, capture_output=True, text=True, check=False).stdout
        wall = time.time() - t0
        hist = []
        with open(os.path.join(td, solver + ".txt")) as f:
            next(f)
            for line in f:
                it, r = line.split(",")
                hist.append(r.strip())
    m = re.search(r"Iter = (\d+)\s+Res = (\S+)", out)
    e = re.search(r"Error max = (\S+) at \((-?\d+) (-?\d+) (-?\d+)\)", out)
    keep = sorted(set(list(range(min(20, len(hist)))) + list(range(0, len(hist), 50)) +
                      list(range(max(0, len(hist) - 5), len(hist)))))
    return dict(real=real, n=n, solver=solver, itr_max=itr_max, coef=coef, precond=precond,
                itr=int(m.group(1)) if m else 0, res=m.group(2) if m else None,
                errmax=e.group(1) if e else None, errloc=[int(e.group(i)) for i in (2, 3, 4)] if e else None,
                history_len=len(hist), history_sha256=hashlib.sha256("\n".join(hist).encode()).hexdigest(),
                history_at={str(i): hist[i] for i in keep},
                wall_seconds_this_container=round(wall, 2), threads=os.cpu_count())


def main():
    done = {}
    if os.path.exists(OUT):
        done = {k: v for k, v in json.load(open(OUT)).items()}
    sel = [int(a) for a in sys.argv[1:]] or range(len(CASES))
    for i in sel:
        c = CASES[i]
        r = run_case(*c)
        done[key(c)] = r
        print(key(c), "itr", r["itr"], "res", r["res"], "err", r["errmax"], "wall", r["wall_seconds_this_container"], flush=True)
        with open(OUT, "w") as f:
            json.dump(done, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
