"""Multi-GPU parity (needs >= 2 CUDA devices; skipped on the 1-GPU box): slabs over NCCL must
reproduce the single-domain result, see tests/mgpu_worker.py."""
import os
import re
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def test_slabs_over_nccl_reproduce_single_domain():
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    world = 4 if n >= 4 else 2
    p = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
                        "--master-addr", "127.0.0.1", "--master-port", "29533", os.path.join(HERE, "mgpu_worker.py")],
                       capture_output=True, text=True, timeout=900)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-3000:]
    assert "FAIL" not in p.stdout


def test_cz_driver_multi_gpu_matches_the_serial_history():
    """The `cz` driver is the multi-rank entry like the reference binary (main.cpp:31-36, cz_Evaluate.cpp:72-159):
    `--gpus N` (j-slabs) and the reference's own `gdv_x gdv_y gdv_z` arguments (bricks; 1 N 1 = j-slabs), one process per GPU, must
    write the history file of the serial run (sweeps are bit-identical, sums differ by reduction order only)."""
    import json
    import tempfile

    import numpy as np
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    exe = os.path.join(os.path.dirname(HERE), "cubez_b200", "cz_f32")
    golden = json.load(open(os.path.join(HERE, "golden", "histories.json")))
    g = [c for c in golden if c["real"] == "f32" and c["n"] == 64 and c["solver"] == "sor2sma"][0]
    for extra in (["--gpus", "2"], ["1", "2", "1"], ["2", "1", "1"], ["1", "1", "2"]):   # j-slabs twice, then i and k bricks
        with tempfile.TemporaryDirectory() as td:
            p = subprocess.run([exe, "64", "64", "64", "sor2sma", "4000", "1.9"] + extra, cwd=td, capture_output=True,
                               text=True, timeout=600)
            assert p.returncode == 0, p.stdout + p.stderr
            txt = open(os.path.join(td, "sor2sma.txt")).read().split("\n")
        div = " x ".join(extra) if extra[0] != "--gpus" else "1 x 2 x 1"
        assert "Iter = %d" % g["itr"] in p.stdout and ("Division = " + div) in p.stdout, p.stdout[-600:]
        hist = np.array([float(l.split(",")[1]) for l in txt[1:] if l.strip()])
        np.testing.assert_allclose(hist, [float(h) for h in g["history"]], rtol=2e-6)
        m = re.search(r"Error max = (\S+) at", p.stdout)
        assert m and abs(float(m.group(1)) - float(g["errmax"])) <= 1e-6, p.stdout[-400:]
    # line solvers need undivided k-lines
    p = subprocess.run([exe, "64", "64", "64", "pcr_rb", "10", "1.5", "1", "1", "2"], capture_output=True, text=True, timeout=300)
    assert "gdv_z = 1" in p.stdout and "Solver error" in p.stdout
