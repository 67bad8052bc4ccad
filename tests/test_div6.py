"""CPU test of the exact division-by-6 used by the fused kernels (fused.cu div_dd<true>):
q' = fma(fma(-6, a*y, a), y, a*y) equals IEEE a/6 for every normal float in the guarded range.
The exhaustive run (tests/tools/div6_check.c full) was done once: 0 mismatches of 3.1e9."""
import os
import subprocess
import tempfile

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


def test_markstein_division_by_six_is_ieee_exact_sampled():
    if "fma" not in open("/proc/cpuinfo").read():
        pytest.skip("host CPU has no FMA")
    with tempfile.TemporaryDirectory() as td:
        exe = os.path.join(td, "div6_check")
        subprocess.check_call(["gcc", "-O2", "-mfma", "-fopenmp", "-ffp-contract=off",
                               os.path.join(HERE, "tools", "div6_check.c"), "-o", exe, "-lm"])
        p = subprocess.run([exe], capture_output=True, text=True)
        assert p.returncode == 0, p.stdout
        assert "mismatches 0" in p.stdout
