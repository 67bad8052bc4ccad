"""GPU parity of the stretched-grid (MAF) twins, SURVEY.md §8f-2: runs tests/gpu_maf_cases.py (42 cases: every `*_maf_`
entry point against the oracle bit for bit, the `_maf` solve loops against the reference driver's golden histories,
the unmodified reference driver on libczb200) in a CHILD process.

STATUS: maf.cu was written after the round-1 GPU budget was spent and has not run on hardware yet (on the CPU the same
source is green against the oracle, tests/test_maf_kernels_emulated.py).  Hence two precautions, both to be removed after
the first green hardware run: the child process (a device fault or a hang in new kernels must not take the main GPU
suite down) and a non-strict xfail mark (XPASS = the 42 cases passed, XFAIL = see the captured output)."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
@pytest.mark.xfail(strict=False, reason="MAF kernels: first hardware run pending (round-1 GPU budget exhausted)")
def test_maf_gpu_cases_in_child_process():
    p = subprocess.run([sys.executable, "-m", "pytest", os.path.join("tests", "gpu_maf_cases.py"), "-q", "-m", "gpu",
                        "-p", "no:cacheprovider"], cwd=ROOT, capture_output=True, text=True, timeout=1500)
    tail = (p.stdout[-6000:] + "\n" + p.stderr[-2000:])
    print(tail)
    assert p.returncode == 0, "tests/gpu_maf_cases.py: rc %d\n%s" % (p.returncode, tail)
    assert " passed" in p.stdout and "failed" not in p.stdout.splitlines()[-1]
