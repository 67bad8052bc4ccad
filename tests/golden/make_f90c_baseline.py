#!/usr/bin/env python
"""Writes tests/golden/baseline_configs_f90c.json: BASELINE.json configs[0..3] solved TO CONVERGENCE (eps = 1e-5) by THE
REFERENCE PROGRAM ITSELF as far as this image can build it -- its unmodified C++ driver (src/main.cpp + src/cz_cpp/*.cpp,
compiled where they lie) over its own Fortran kernels with their own OpenMP directives, translated mechanically to C
(oracle/f90_to_c.py --openmp -> oracle/_ref/cz_f90c_omp_{f32,f64}; no Fortran compiler exists here).  Same cases and
record layout as make_golden_baseline.py (which runs the driver over the hand-written port): the two files side by
side are the iteration counts the GPU results are compared with.

Threads: all cores, except where the reference's OpenMP code races (jacobi: END DO NOWAIT before the copy-back,
cz_solver.f90:355-372 -- also as preconditioner): 1 thread there, so that the record is reproducible.
The float residual reduction is REAL and its order depends on the team size.

    python tests/golden/make_f90c_baseline.py [case-index ...]      (this container only; ~1 h for all cases)
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
OUT = os.path.join(HERE, "baseline_configs_f90c.json")
sys.path.insert(0, HERE)
from make_golden_baseline import CASES, key   # noqa: E402


def run_case(real, n, solver, itr_max, coef, precond):
    exe = os.path.join(ROOT, "oracle", "_ref", "cz_f90c_omp_%s" % real)
    args = [exe, str(n), str(n), str(n), solver, str(itr_max), coef] + ([precond] if precond else [])
    racy = solver == "jacobi" or precond == "jacobi"
    threads = 1 if racy else os.cpu_count()
    env = dict(os.environ, OMP_NUM_THREADS=str(threads), OMP_PROC_BIND="close")
    with tempfile.TemporaryDirectory() as td:
        t0 = time.time()
        out = subprocess.run(args, cwd=td, env=env, capture_output=True, text=True, check=False).stdout
        wall = time.time() - t0
        with open(os.path.join(td, solver + ".txt")) as f:
            hist = [line.split(",")[1].strip() for line in list(f)[1:]]
    m = re.search(r"Iter = (\d+)\s+Res = (\S+)", out)
    e = re.search(r"Error max = (\S+) at \((-?\d+) (-?\d+) (-?\d+)\)", out)
    keep = sorted(set(list(range(min(20, len(hist)))) + list(range(0, len(hist), 50)) + list(range(max(0, len(hist) - 5), len(hist)))))
    return dict(real=real, n=n, solver=solver, itr_max=itr_max, coef=coef, precond=precond, threads=threads,
                itr=int(m.group(1)) if m else 0, res=m.group(2) if m else None, errmax=e.group(1) if e else None,
                errloc=[int(e.group(i)) for i in (2, 3, 4)] if e else None, history_len=len(hist),
                history_sha256=hashlib.sha256("\n".join(hist).encode()).hexdigest(),
                history_at={str(i): hist[i] for i in keep}, wall_seconds_this_container=round(wall, 2),
                made_by="oracle/_ref/cz_f90c_omp_%s (reference driver + translated reference Fortran, OpenMP %d threads)" % (real, threads))


def main():
    which = [int(a) for a in sys.argv[1:]] or range(len(CASES))
    data = json.load(open(OUT)) if os.path.exists(OUT) else {}
    for i in which:
        c = CASES[i]
        r = run_case(*c)
        data[key(c)] = r
        print(key(c), "itr", r["itr"], "res", r["res"], "err", r["errmax"], "wall %.1f s" % r["wall_seconds_this_container"], flush=True)
        with open(OUT, "w") as f:
            json.dump(data, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
