#!/usr/bin/env python
"""Generates tests/golden/histories.json by running oracle/_ref/cz_ref_{f32,f64}: the reference's
UNMODIFIED C++ driver (src/main.cpp + src/cz_cpp/*.cpp compiled where they lie under
/root/reference) linked to the C++ restatement of its Fortran kernels (oracle/cz_f90_port.cpp),
OMP_NUM_THREADS=1, ORACLE_SUM=double.  Needs /root/reference (this container only); the JSON is
committed so the GPU box and the CPU test-suite never read the reference.

    python tests/golden/make_golden.py
"""
import json
import os
import re
import subprocess
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))

CASES = [
    # (real, N, solver, itr_max, coef, precond)
    ("f32", 64, "jacobi", 4000, "0.8", None),          # BASELINE.json configs[0]
    ("f32", 32, "jacobi", 2000, "0.8", None),
    ("f32", 32, "sor2sma", 2000, "1.5", None),
    ("f32", 32, "sor2sma", 2000, "1.8", None),
    ("f32", 32, "psor", 2000, "1.5", None),
    ("f32", 32, "pcr_rb", 2000, "1.5", None),
    ("f32", 32, "pcr_rb_esa", 2000, "1.5", None),
    ("f32", 32, "pcr_j_esa", 3000, "0.9", None),
    ("f32", 32, "pcr", 2000, "1.5", None),
    ("f32", 32, "pcr_eda", 2000, "1.5", None),
    ("f32", 32, "pcr_esa", 2000, "1.5", None),
    ("f32", 32, "pbicgstab", 200, "1.1", "sor2sma"),
    ("f32", 32, "pbicgstab", 200, "0.8", "jacobi"),
    ("f32", 32, "pbicgstab", 200, "1.1", "none"),
    ("f32", 32, "pbicgstab", 200, "1.1", "pcr_rb"),
    ("f64", 32, "jacobi", 2000, "0.8", None),
    ("f64", 32, "sor2sma", 2000, "1.5", None),
    ("f64", 32, "pcr_rb", 2000, "1.5", None),
    ("f64", 32, "pbicgstab", 200, "1.1", "sor2sma"),
    ("f64", 32, "pbicgstab", 200, "1.1", "none"),
    ("f32", 64, "sor2sma", 4000, "1.9", None),
    ("f32", 64, "pbicgstab", 400, "1.1", "sor2sma"),
    ("f32", 64, "pcr_rb", 4000, "1.8", None),
    # stretched-grid (MAF) twins on the driver's uniform mesh
    ("f32", 32, "jacobi_maf", 2000, "0.8", None),
    ("f32", 32, "sor2sma_maf", 2000, "1.5", None),
    ("f32", 32, "psor_maf", 2000, "1.5", None),
    ("f32", 32, "pcr_rb_maf", 2000, "1.5", None),
    ("f32", 32, "pcr_rb_esa_maf", 2000, "1.5", None),
    ("f32", 32, "pcr_maf", 2000, "1.5", None),
    ("f32", 32, "pcr_eda_maf", 2000, "1.5", None),
    ("f32", 32, "pcr_esa_maf", 2000, "1.5", None),
    ("f32", 32, "pbicgstab_maf", 200, "1.1", "sor2sma_maf"),
    ("f32", 32, "pbicgstab_maf", 200, "1.1", "pcr_rb_maf"),
    ("f32", 32, "pbicgstab_maf", 200, "1.1", "none"),
    ("f64", 32, "sor2sma_maf", 2000, "1.5", None),
    ("f64", 32, "pcr_rb_maf", 2000, "1.5", None),
    ("f64", 32, "pbicgstab_maf", 200, "1.1", "sor2sma_maf"),
]


def run_case(real, n, solver, itr_max, coef, precond):
    exe = os.path.join(ROOT, "oracle", "_ref", "cz_ref_%s" % real)
    args = [exe, str(n), str(n), str(n), solver, str(itr_max), coef] + ([precond] if precond else [])
    env = dict(os.environ, OMP_NUM_THREADS="1", ORACLE_SUM="double")
    with tempfile.TemporaryDirectory() as td:
        out = subprocess.run(args, cwd=td, env=env, capture_output=True, text=True, check=False).stdout
        hist = []
        with open(os.path.join(td, solver + ".txt")) as f:
            next(f)
            for line in f:
                it, r = line.split(",")
                hist.append(r.strip())
    m = re.search(r"Iter = (\d+)\s+Res = (\S+)", out)
    e = re.search(r"Error max = (\S+) at \((-?\d+) (-?\d+) (-?\d+)\)", out)
    return dict(real=real, n=n, solver=solver, itr_max=itr_max, coef=coef, precond=precond,
                itr=int(m.group(1)) if m else 0, res=m.group(2) if m else None,
                errmax=e.group(1) if e else None, errloc=[int(e.group(i)) for i in (2, 3, 4)] if e else None,
                history=hist)


def main():
    out = [run_case(*c) for c in CASES]
    with open(os.path.join(HERE, "histories.json"), "w") as f:
        json.dump(out, f, indent=0)
    for c in out:
        print(c["real"], c["n"], c["solver"], c["precond"], "itr", c["itr"], "res", c["res"], "err", c["errmax"])


if __name__ == "__main__":
    main()
