#!/usr/bin/env python
"""Generates tests/golden/bench_parity.json with the CPU ORACLE (oracle/liboracle_f32.so): the single-domain
result of bench.py's parity leg -- PAR_ITERS red-black SOR iterations (omega = bench.OMEGA) from the
global-index start field of cubez_b200/synth.py -- as residual history + exact field checksum.  bench.py
compares every run (any number of j-slabs) with it and prints `parity_vs_n1`.

    python tests/golden/make_bench_parity.py [NY ...]        # default: 1024 (the cube) ; 2048 = --weak at 2 GPUs

Needs ~9 GB of host memory per 1024 planes; runs on the host cores only.
"""
import json
import os
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import bench  # noqa: E402
import oracle_util as ou  # noqa: E402
from cubez_b200 import synth  # noqa: E402

OUT = os.path.join(HERE, "bench_parity.json")


def main():
    n = int(os.environ.get("CZB_BENCH_N", "1024"))
    nys = [int(a) for a in sys.argv[1:]] or [n]
    done = json.load(open(OUT)) if os.path.exists(OUT) else {}
    for ny in nys:
        t0 = time.time()
        o = ou.OracleCZ([n, ny, n], "f32", eps=0.0)
        X, B = o.setup_laplace()
        synth.fill(X, o.sz)
        o.bc_k(X)
        itr, res, hist = o.stationary(bench.SOLVER, X, B, bench.PAR_ITERS, bench.OMEGA)
        done["%dx%dx%d" % (n, ny, n)] = {
            "solver": bench.SOLVER, "omega": bench.OMEGA, "iterations": bench.PAR_ITERS,
            "history": [float(h) for h in hist], "field_checksum": synth.checksum_owned(X, o.idx),
            "made_by": "oracle/cz_f90_port.cpp (psor2sma_core, ORACLE_SUM=double), tests/golden/make_bench_parity.py"}
        print(ny, hist, done["%dx%dx%d" % (n, ny, n)]["field_checksum"], "%.1f s" % (time.time() - t0), flush=True)
        with open(OUT, "w") as f:
            json.dump(done, f, indent=1, sort_keys=True)
        del o, X, B


if __name__ == "__main__":
    main()
