"""CPU checks of bench.py's host-side plumbing (no GPU work): the optional time-to-convergence child process can only
ever add a key to the headline JSON line, never cost it."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def test_time_to_convergence_child_is_fail_safe(monkeypatch):
    class Ok:
        returncode = 0; stdout = 'import noise\n{"jacobi 64^3 w=0.8": {"iterations": 2742, "converged": true}}\n'; stderr = ""

    class Bad:
        returncode = 134; stdout = ""; stderr = "abort"

    monkeypatch.setattr(bench.subprocess, "run", lambda *a, **k: Ok)
    assert bench.time_to_convergence()["jacobi 64^3 w=0.8"]["iterations"] == 2742
    monkeypatch.setattr(bench.subprocess, "run", lambda *a, **k: Bad)
    assert bench.time_to_convergence() is None

    def hang(*a, **k):
        raise subprocess.TimeoutExpired("bench.py", 1)
    monkeypatch.setattr(bench.subprocess, "run", hang)
    assert bench.time_to_convergence() is None


def test_ttc_cases_are_the_baseline_configs():
    labels = [c[0] for c in bench.TTC_CASES]
    assert any("jacobi 64^3" in x for x in labels) and any("sor2sma 256^3" in x for x in labels)
    assert any("pbicgstab+sor2sma 512^3" in x for x in labels) and any("pcr_rb 512^3" in x for x in labels)
    for label, n, solver, itmax, omg, pre in bench.TTC_CASES:
        assert solver in ("jacobi", "sor2sma", "pbicgstab", "pcr_rb") and 0 < omg < 2 and itmax > 0
