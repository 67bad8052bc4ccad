"""CPU checks of bench.py's host-side plumbing (no GPU work): the optional time-to-convergence child process can only
ever add a key to the headline JSON line, never cost it."""
import json
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
    for label, n, solver, itmax, omg, pre, real in bench.TTC_CASES + bench.TTC_BIG:
        assert solver in ("jacobi", "sor2sma", "pbicgstab", "pcr_rb") and 0 < omg < 2 and itmax > 0 and real in ("f32", "f64")
    assert any("f64" in x for x in labels)                      # configs[2] asks for float AND double
    assert any("1024^3" in c[0] for c in bench.TTC_BIG)         # configs[4]


def test_cpu_arm_runs_with_an_explicit_openmp_team(monkeypatch):
    """under torchrun OMP_NUM_THREADS=1 is exported; the CPU arm must run in a child whose environment sets
    the team size, and must report omp_get_max_threads(), not os.cpu_count()."""
    monkeypatch.setenv("OMP_NUM_THREADS", "1")
    monkeypatch.setenv("CZB_BENCH_N", "48")
    r = bench.run_cpu_child("sample", steps=2, warmup=1, timeout=300)
    assert r is not None and r["value"] > 0
    assert r["cores"] == bench.host_cores()
    assert "omp_get_max_threads" in r["sample"]
    # the reference's own Fortran (translated, OpenMP mapped) when it was built here, else the hand-written port
    have = os.path.exists(os.path.join(os.path.dirname(os.path.abspath(bench.__file__)), "oracle", "_ref", "libczf90c_omp_f32.so"))
    assert r["kind"] == ("reference" if have else "port")
    monkeypatch.setenv("CZB_BENCH_CPU", "port")
    assert bench.run_cpu_child("sample", steps=1, warmup=1, timeout=300)["kind"] == "port"


def test_golden_keys_of_the_ttc_cases_exist():
    gold = bench.load_golden_configs()
    ref = bench.load_golden_configs("f90c")      # the reference program's own runs of the same configurations
    for label, n, solver, itmax, omg, pre, real in bench.TTC_CASES:
        k = bench.golden_key(real, n, solver, omg, pre)
        assert k in gold, label
        assert k in ref and int(ref[k]["itr"]) > 0 and float(ref[k]["res"]) < 1.0e-5, label
        assert {"wall_seconds_this_container", "threads"} <= set(ref[k])


def test_start_field_is_decomposition_independent():
    import numpy as np
    from cubez_b200 import decomp, synth
    gsz = [20, 37, 18]
    full = np.zeros(decomp.field_shape(gsz), np.float32)
    synth.fill(full, gsz)
    tot = 0
    for nr in (2, 3, 5):
        tot = 0
        for r in range(nr):
            d = decomp.slab_decompose(gsz, nr, r)
            f = np.zeros(decomp.field_shape(d["sz"]), np.float32)
            synth.fill(f, d["sz"], head_j=d["head"][1])
            j0 = d["head"][1]                         # local plane 1 = global plane head
            assert np.array_equal(f[2:-2, 2:-2, 2:-2], full[j0 + 1:j0 + 1 + d["sz"][1], 2:-2, 2:-2])
            tot += synth.checksum_owned(f, d["idx"])
        assert tot == synth.checksum_owned(full, [2, gsz[0] - 1, 2, gsz[1] - 1, 2, gsz[2] - 1])


def test_golden_records_are_attached_to_the_time_to_convergence_entries():
    labels = [c[0] for c in bench.TTC_CASES + bench.TTC_BIG]
    ttc = {lab: {"iterations": 1} for lab in labels}
    ttc["also_f64"] = {"x": 1}
    cpu = json.load(open(os.path.join(os.path.dirname(os.path.abspath(bench.__file__)), "profiles", "r02_cpu_ttc_reference.json")))
    out = bench.attach_golden(ttc, cpu)
    for c in bench.TTC_CASES:
        e = out[c[0]]
        assert e["oracle_iterations"] > 0 and e["reference_iterations"] > 0 and e["cpu"]["code"] == "reference", c[0]
        assert abs(e["oracle_iterations"] - e["reference_iterations"]) <= 2, c[0]
    assert out["also_f64"] == {"x": 1}
    json.dumps(out)
