#!/usr/bin/env python
"""bench.py -- headline benchmark of cubez_b200 (BASELINE.json: "Stencil GLUPS & time-to-convergence at
1/2/4/8 B200; HBM GB/s vs peak").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--weak] [--impl reference]
    torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N --steps K --warmup W

Workload (config.workload): red-black SOR ("sor2sma", both colours = one step) on the 1024^3 float cube of
BASELINE.json configs[4]; it fits one B200, so it is also the N=1 workload.  N>1 = STRONG scaling: the same cube
cut into N j-slabs, one process per GPU (cubez_b200/csrc/comm.cpp); --weak extends the j axis with N instead.
The start field is a hash of the GLOBAL node index (cubez_b200/synth.py), i.e. the same problem for every N.

One JSON line on rank 0:
  value        GLUPS (1e9 lattice updates/s, whole job), fields resident in HBM.  NWIN windows of exactly K
               iterations each are timed with CUDA events on the launching stream between barriers, max over
               ranks per window; value / ms_per_step come from the MEDIAN window, all windows are listed
  parity_vs_n1 PAR_ITERS iterations from the start field before the timed region: residual history and an exact
               order-independent checksum of the field against tests/golden/bench_parity.json, which was written
               by the CPU ORACLE on the single domain (tests/golden/make_bench_parity.py)
  e2e          the same K iterations through the public call CZ.solve() with HOST (pinned) arrays:
               H2D of X and B, the solve loop, D2H of X and of the residual history, all timed;
               e2e.to_convergence (N=1): the same call run until res < 1e-5 (256^3 and 1024^3 sor2sma)
  roofline     dominant kernel (fused two-colour sweep k_rb2_pk): algorithmic bytes (4R = 16 B per
               cell.iteration) / measured time per iteration, against the measured HBM copy bandwidth
  cpu_baseline the reference's own psor2sma_core -- its Fortran text with its OpenMP directives, translated mechanically
               to C (oracle/f90_to_c.py --openmp -> oracle/_ref/libczf90c_omp_f32.so, gcc -O3 -fopenmp; kind "reference")
               -- or, where that prebuilt file is absent, the hand-written port (oracle/cz_f90_port.cpp, kind "port"),
               on all host cores, bounded sample, in a CHILD process whose OpenMP team size is set explicitly and
               reported from omp_get_max_threads()
  also         PBiCGstab+sor2sma, Jacobi and line SOR on the same grid; f64 sor2sma / PBiCGstab at 512^3 (N=1)
  time_to_convergence  BASELINE.json configs[0..3] (+ the 1024^3 solvers of configs[4]) solved to res < 1e-5:
               iterations and device time of the solver loop, next to the oracle's iteration count (committed
               golden, tests/golden/baseline_configs.json) and the CPU port's seconds on this box
--impl reference times the CPU implementation alone (all host threads) on the same workload, same line.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_CUBE = int(os.environ.get("CZB_BENCH_N", "1024"))
N_Y = int(os.environ.get("CZB_BENCH_NY", str(N_CUBE)))     # experiments only: shorter j extent (slab axis)
OMEGA = 1.5
SOLVER = "sor2sma"
ALGO_BYTES_PER_CELL = 16.0      # red-black SOR, float, 4R per cell.iteration
NWIN = 5                        # timed windows of --steps iterations each (median reported)
PAR_ITERS = 8                   # iterations of the parity leg
GOLDEN_PARITY = os.path.join(ROOT, "tests", "golden", "bench_parity.json")
GOLDEN_CONFIGS = os.path.join(ROOT, "tests", "golden", "baseline_configs.json")
# dram__bytes_read.sum + dram__bytes_write.sum of ONE k_rb2_pk launch at N=1, 1024^3 (profiles/*_rb2_summary.md);
# null for any other N / grid: the figure is a property of that capture, not of this run
NCU_TRAFFIC_N1 = {"bytes": 14.25e9, "source": "profiles/r02_rb2_summary.md"}


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed windows, in a thread, through NVML (no fork near
    the timed region); falls back to one `nvidia-smi -lms` child started before the first barrier."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, torch_index):
        self.index, self.rows, self.proc, self.h, self.stop_flag, self.th = torch_index, [], None, None, False, None
        self.active = False      # rows are kept only while a timed window is open
        self.mx = None

    def start(self):
        try:
            import pynvml
            import torch
            pynvml.nvmlInit()
            uuid = str(torch.cuda.get_device_properties(self.index).uuid)
            uuid = uuid if uuid.startswith("GPU-") else "GPU-" + uuid
            self.h = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode() if hasattr(uuid, "encode") else uuid)
            self.nv = pynvml
            self.mx = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.th = threading.Thread(target=self._poll, daemon=True)
            self.th.start()
            return
        except Exception:
            self.h = None
        try:
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = vis.split(",")[self.index] if vis else str(self.index)
            self.proc = subprocess.Popen(["nvidia-smi", "-i", phys, "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _poll(self):
        nv = self.nv
        bits = [nv.nvmlClocksEventReasonHwSlowdown, nv.nvmlClocksEventReasonHwThermalSlowdown,
                nv.nvmlClocksEventReasonSwThermalSlowdown, nv.nvmlClocksEventReasonSwPowerCap] \
            if hasattr(nv, "nvmlClocksEventReasonHwSlowdown") else \
            [nv.nvmlClocksThrottleReasonHwSlowdown, nv.nvmlClocksThrottleReasonHwThermalSlowdown,
             nv.nvmlClocksThrottleReasonSwThermalSlowdown, nv.nvmlClocksThrottleReasonSwPowerCap]
        get_reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
        while not self.stop_flag:
            if self.active:
                try:
                    sm = float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                    r = int(get_reasons(self.h))
                    self.rows.append([sm] + [bool(r & b) for b in bits])
                except Exception:
                    pass
            time.sleep(0.004)

    def _read(self):
        for line in self.proc.stdout:
            x = [v.strip() for v in line.split(",")]
            if self.active and len(x) >= 6 and x[0].replace(".", "").isdigit():
                self.mx = float(x[1]) if x[1].replace(".", "").isdigit() else self.mx
                self.rows.append([float(x[0])] + [x[2 + i].startswith("Active") for i in range(4)])

    def stop(self):
        self.stop_flag = True
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()
        if self.h is None and self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no NVML / nvidia-smi"], "samples": 0}
        sm = sorted(r[0] for r in self.rows)
        reasons = sorted({self.NAMES[i] for r in self.rows for i in range(4) if r[1 + i]})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": self.mx, "reasons": reasons,
                "samples": len(sm), "source": "nvml thread" if self.h is not None else "nvidia-smi -lms 20"}


# ======================================================================================= CPU arm (child process)
# BASELINE.json configs[0..3] to convergence (eps = 1e-5, cz.h:162): (label, N, solver, ItrMax, omega, precond, real)
TTC_CASES = (("jacobi 64^3 w=0.8", 64, "jacobi", 4000, 0.8, None, "f32"),
             ("sor2sma 256^3 w=1.95", 256, "sor2sma", 20000, 1.95, None, "f32"),
             ("pbicgstab+sor2sma 512^3 w=1.1", 512, "pbicgstab", 400, 1.1, "sor2sma", "f32"),
             ("pbicgstab+sor2sma 512^3 w=1.1 f64", 512, "pbicgstab", 400, 1.1, "sor2sma", "f64"),
             ("pcr_rb 512^3 w=1.9", 512, "pcr_rb", 8000, 1.9, None, "f32"))
# configs[4]: the 1024^3 solvers of the north-star target, to convergence, at every N
TTC_BIG = (("sor2sma 1024^3 w=1.99", 1024, "sor2sma", 20000, 1.99, None, "f32"),
           ("pbicgstab+sor2sma 1024^3 w=1.1", 1024, "pbicgstab", 600, 1.1, "sor2sma", "f32"))


def golden_key(real, n, solver, omg, pre):
    return "%s %d %s %s %s" % (real, n, solver, ("%g" % omg), pre)


def load_golden_configs(which=None):
    """which = "f90c": the same configurations run by the reference PROGRAM (its unmodified driver over its own Fortran +
    OpenMP, translated to C; tests/golden/make_f90c_baseline.py) instead of the driver over the oracle port."""
    try:
        return json.load(open(GOLDEN_CONFIGS if which is None else GOLDEN_CONFIGS.replace(".json", "_%s.json" % which)))
    except Exception:
        return {}


def cpu_worker(mode, steps, warmup):
    """Child process of the CPU legs (the ONLY code of this file that executes oracle/).  Prints one JSON object."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import numpy as np
    import oracle_util as ou
    cores = host_cores()
    lib = ou.load("f32")
    lib.oracle_set_threads(cores)
    threads = int(lib.oracle_max_threads())
    n = N_CUBE
    if mode in ("sample", "full"):
        # psor2sma_core (cz_solver.f90:404-493), both colours per step.  Preferred: the reference's OWN Fortran text
        # with its OpenMP directives, translated mechanically to C (oracle/f90_to_c.py --openmp) and compiled with the
        # reference's GNU flags (-O3 -fopenmp) -> oracle/_ref/libczf90c_omp_f32.so, kind "reference".  Without that
        # prebuilt file (it needs /root/reference once): the hand-written port oracle/cz_f90_port.cpp, kind "port".
        ny = n if mode == "full" else min(128, n)
        o = ou.OracleCZ([n, n, n], "f32", sz=[n, ny, n])
        kind, what_code = "port", "oracle port of psor2sma_core"
        ref_so = os.path.join(ROOT, "oracle", "_ref", "libczf90c_omp_f32.so")
        if os.path.exists(ref_so) and os.environ.get("CZB_BENCH_CPU", "") != "port":
            import cubez_b200.ffi as ffi
            o.dll = ffi.bind_reference_symbols(C.CDLL(ref_so))     # same libgomp, same team size as reported below
            kind, what_code = "reference", ("the reference's psor2sma_core, cz_solver.f90:404-493 with its !$OMP "
                                            "directives, translated to C by oracle/f90_to_c.py, gcc -O3 -fopenmp")
        X, B = o.setup_laplace()
        from cubez_b200 import synth
        synth.fill(X, o.sz)
        o.bc_k(X)
        o.ac1 = np.float32(OMEGA)
        cells = (n - 2) * (ny - 2) * (n - 2)
        res = C.c_double(0)
        for _ in range(max(warmup, 1)):
            o.sweep_rb(X, B, res)
        t0 = time.perf_counter()
        for _ in range(steps):
            o.sweep_rb(X, B, res)
        dt = time.perf_counter() - t0
        what = ("the whole %d^3 grid" % n) if ny == n else ("a %dx%dx%d j-slab of the %d^3 grid" % (n, ny, n, n))
        print(json.dumps({"value": cells * steps / dt / 1e9, "ms": dt / steps * 1e3, "cores": threads, "kind": kind,
                          "sample": "%d iterations of sor2sma (%s; OpenMP %d threads = "
                                    "omp_get_max_threads(), %d usable cores) on %s per step"
                                    % (steps, what_code, threads, cores, what)}),
              flush=True)
        return
    # mode == "ttc": the reference's timed region is the solver loop (cz_Evaluate.cpp:419-421).  Small configs run to
    # convergence; the 512^3 ones are timed over a bounded window and extrapolated with the oracle's committed
    # iteration count (the full CPU runs take minutes: see wall_seconds in tests/golden/baseline_configs.json).
    gold = load_golden_configs()
    out = {}
    # one-thread cases (the reference's racy OpenMP Jacobi) last: the team size is only ever lowered
    for label, nn, solver, itmax, omg, pre, real in sorted(TTC_CASES, key=lambda c: c[2] == "jacobi" or c[5] == "jacobi"):
        g = gold.get(golden_key(real, nn, solver, omg, pre))
        lr = ou.load(real)
        lr.oracle_set_threads(cores)
        o = ou.OracleCZ([nn] * 3, real)
        ref_so = os.path.join(ROOT, "oracle", "_ref", "libczf90c_omp_%s.so" % real)
        kind_cpu, nthr = "port", threads
        if os.path.exists(ref_so) and os.environ.get("CZB_BENCH_CPU", "") != "port":
            # the reference's own kernels (Fortran + OpenMP directives translated to C); its Jacobi races with more than
            # one thread (END DO NOWAIT before the copy-back, cz_solver.f90:355-372): one thread there
            import cubez_b200.ffi as ffi
            o.dll = ffi.bind_reference_symbols(C.CDLL(ref_so))
            kind_cpu = "reference"
            if solver == "jacobi" or pre == "jacobi":
                lr.oracle_set_threads(1)
                nthr = 1
        X, B = o.setup_laplace()
        full = nn <= 256
        window = itmax if full else (3 if solver == "pbicgstab" else 6)
        t0 = time.perf_counter()
        if solver == "pbicgstab":
            itr, res, hist = o.pbicgstab(X, B, window + 1 if not full else itmax, omg, pre)
            done = len(hist)
        else:
            itr, res, hist = o.stationary(solver, X, B, window, omg)
            done = len(hist)
        dt = time.perf_counter() - t0
        e = {"threads": nthr, "code": kind_cpu, "iterations_timed": int(done), "seconds_timed": dt}
        if full:
            e.update({"kind": "measured to convergence", "iterations": int(itr), "seconds": dt, "residual": float(res)})
        elif g:
            e.update({"kind": "extrapolated: seconds per iteration of a %d-iteration window x the oracle's iteration count" % done,
                      "iterations": int(g["itr"]), "seconds": dt / max(done, 1) * int(g["itr"])})
        out[label] = e
        del o, X, B
    print(json.dumps(out), flush=True)


def run_cpu_child(mode, steps=3, warmup=1, timeout=600):
    """The CPU legs run in a child whose environment sets the OpenMP team explicitly: under torchrun the parent has
    OMP_NUM_THREADS=1 exported and libgomp has already read it."""
    cores = host_cores()
    env = dict(os.environ, OMP_NUM_THREADS=str(cores), OMP_PROC_BIND="close", OMP_DYNAMIC="FALSE", CUDA_VISIBLE_DEVICES="")
    try:
        p = subprocess.run([sys.executable, os.path.abspath(__file__), "--cpu-worker", mode, "--steps", str(steps),
                            "--warmup", str(warmup)], capture_output=True, text=True, timeout=timeout, env=env)
        if p.returncode != 0:
            sys.stderr.write("bench.py: cpu child (%s) rc=%d\n%s\n" % (mode, p.returncode, p.stderr[-2000:]))
            return None
        for ln in reversed(p.stdout.strip().splitlines()):
            if ln.startswith("{"):
                return json.loads(ln)
    except Exception as ex:
        sys.stderr.write("bench.py: cpu child (%s) failed: %r\n" % (mode, ex))
    return None


def workload_string(n, ny, weak):
    if weak:
        return "sor2sma %dx%dx%d f32 (j extended with the GPU count), omega=%.2f, both colours per step" % (n, ny, n, OMEGA)
    if ny != n:
        return "sor2sma %dx%dx%d f32, omega=%.2f, both colours per step" % (n, ny, n, OMEGA)
    return "sor2sma %d^3 f32, omega=%.2f, both colours per step" % (n, OMEGA)


def run_reference(args, rank):
    if rank != 0:
        return
    # the whole cube per step when the run stays within ~2 minutes at ~1 GLUPS, else a j-slab sample
    cells = float(N_CUBE - 2) ** 3
    mode = "full" if cells * (args.steps + args.warmup) / 1.0e9 < 120.0 else "sample"
    r = run_cpu_child(mode, args.steps, args.warmup, timeout=1500)
    if r is None:
        print(json.dumps({"impl": "reference", "unavailable": "the CPU oracle child process failed (see stderr)"}), flush=True)
        return
    print(json.dumps({
        "impl": "reference", "metric": "stencil_glups", "value": r["value"], "unit": "GLUPS", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms"], "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_string(N_CUBE, N_Y, False)},
        "cpu_baseline": {"value": r["value"], "unit": "GLUPS", "cores": r["cores"], "kind": r.get("kind", "port"), "sample": r["sample"]},
        "e2e": {"value": r["value"], "unit": "GLUPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }), flush=True)


# ======================================================================================= GPU arm
def median(v):
    s = sorted(v)
    return s[len(s) // 2]


def run_gpu(args, rank, world, local_rank):
    import numpy as np
    import torch
    import cubez_b200
    from cubez_b200 import synth

    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    n = N_CUBE
    # --weak (BASELINE.json configs[4] "weak": 1024 x (1024*G) x 1024, the j axis extended): per-GPU work fixed
    weak = bool(getattr(args, "weak", False) or os.environ.get("CZB_BENCH_WEAK"))
    ny = N_Y * world if weak else N_Y

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def min_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        return float(t.item())

    def sum_over_ranks_int(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.int64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)      # wraps modulo 2^64 like the single-domain checksum
        return int(t.item())

    def make_cz(gsz, real="f32", eps=0.0):
        cz = cubez_b200.CZ(gsz, real, device=local_rank, nranks=world, rank=rank, eps=eps)
        if world > 1:
            idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
            if rank == 0:
                idt.copy_(torch.frombuffer(bytearray(cz.comm_unique_id()), dtype=torch.uint8))
            dist.broadcast(idt, 0)
            cz.comm_init(bytes(idt.cpu().numpy().tobytes()))
        return cz

    def start_state(cz):
        P, RHS = cz.setup_laplace()
        synth.fill(P, cz.sz, head_j=cz.head[1], gny=cz.gsz[1])
        cz.bc_k(P)
        return P, RHS

    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()                    # before any barrier; rows are recorded only inside the windows
    cz = make_cz([n, ny, n])               # eps = 0: never stops early

    # ---- parity leg: same problem at every N, compared with the ORACLE's single-domain result -------------
    P, RHS = start_state(cz)
    outp = cz.solve(SOLVER, P, RHS, PAR_ITERS, OMEGA, want_history=True)
    chk = sum_over_ranks_int(synth.checksum_owned(P, cz.idx))
    parity = {"iterations": PAR_ITERS, "history": [float(x) for x in outp["history"]], "field_checksum": chk}
    gold = {}
    try:
        gold = json.load(open(GOLDEN_PARITY)).get("%dx%dx%d" % (n, ny, n), {})
    except Exception:
        pass
    if gold:
        gh = np.array(gold["history"][:PAR_ITERS], dtype=np.float64)
        rel = float(np.max(np.abs(np.array(parity["history"]) - gh) / gh))
        parity.update({"max_rel_diff_history": rel, "checksum_equal": chk == int(gold["field_checksum"]),
                       "ok": bool(rel <= 1e-12 and chk == int(gold["field_checksum"])), "golden": gold.get("made_by", "oracle")})
    else:
        parity.update({"ok": None, "note": "no committed single-domain golden for this grid"})
    del parity["history"]
    parity["history_last"] = float(outp["history"][-1])

    # ---- device-resident: fields stay in HBM ---------------------------------------------------
    P, RHS = start_state(cz)
    warm = max(args.warmup, 3)
    cz.solve(SOLVER, P, RHS, warm, OMEGA, want_history=False)     # warm-up (>= 3 steps)
    windows, loops, loops_min = [], [], []
    launches = 0
    out = None
    for w in range(NWIN):
        barrier()
        l0 = cz.launch_count()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        cz.sync_stream()
        barrier()                          # second barrier + synchronize immediately before the start event
        if sampler:
            sampler.active = True
        ev0.record()
        out = cz.solve(SOLVER, P, RHS, args.steps, OMEGA, want_history=False)
        ev1.record()
        barrier()
        if sampler:
            sampler.active = False
        windows.append(max_over_ranks(ev0.elapsed_time(ev1)))
        loops.append(max_over_ranks(out["ms"]))        # the library's own events around the iteration loop
        loops_min.append(min_over_ranks(out["ms"]))
        launches = cz.launch_count() - l0
    clocks = sampler.stop() if sampler else None
    ms_total = median(windows)
    ms_loop = median(loops)
    cells = float(cz.inner_cells_global)
    glups = cells * args.steps / (ms_total * 1e-3) / 1e9
    ms_iter = ms_loop / args.steps

    # ---- the other solvers of the north-star target on the same grid (device resident, after the
    # headline region; reported under "also", not part of value/e2e) -------------------------------
    also = {}
    if not args.no_also:
        for name, pre, omg, bpc, k in (("pbicgstab", "sor2sma", OMEGA, 340.0, 6), ("jacobi", None, 0.8, 12.0, 10),
                                       ("pcr_rb", None, OMEGA, 20.0, 8)):
            P2, R2 = start_state(cz)
            cz.solve(name, P2, R2, 3, omg, precond=pre, want_history=False)
            ms2 = []
            for w in range(3):
                barrier()
                o2 = cz.solve(name, P2, R2, k, omg, precond=pre, want_history=False)
                barrier()
                done = (k - 1) if name == "pbicgstab" else k
                ms2.append(max_over_ranks(o2["ms"]) / done)
            m2 = median(ms2)
            gl = cells / (m2 * 1e-3) / 1e9
            also[name + ("+" + pre if pre else "")] = {
                "value": gl, "unit": "G cell-iterations/s", "ms_per_iteration": m2,
                "algorithmic_bytes_per_cell": bpc, "hbm_frac_of_measured": gl * bpc / world / measured_peak()[0]}
            del P2, R2
        torch.cuda.empty_cache()

    # ---- end to end: host arrays through the public call ------------------------------------------
    Ph = torch.empty(cz.shape, dtype=torch.float32).pin_memory()
    Bh = torch.empty(cz.shape, dtype=torch.float32).pin_memory()
    Ph.copy_(P.cpu()); Bh.copy_(RHS.cpu())
    Pn, Bn = Ph.numpy(), Bh.numpy()
    del P, RHS
    torch.cuda.empty_cache()
    barrier()
    t0 = time.perf_counter()
    out2 = cz.solve(SOLVER, Pn, Bn, args.steps, OMEGA, want_history=True)         # H2D + loop + D2H inside
    torch.cuda.synchronize()
    t_e2e = max_over_ranks(time.perf_counter() - t0)
    e2e_glups = cells * args.steps / t_e2e / 1e9
    field_bytes = float(np.prod(cz.shape)) * 4.0
    h2d_per_step = 2.0 * field_bytes / args.steps
    d2h_per_step = (field_bytes + 8.0 * args.steps) / args.steps
    del Ph, Bh, Pn, Bn
    cz.close()

    peak, peak_src = measured_peak()
    achieved = (cells / world) * ALGO_BYTES_PER_CELL / (ms_iter * 1e-3) / 1e9      # GB/s per GPU
    traffic = NCU_TRAFFIC_N1["bytes"] if (world == 1 and n == 1024 and ny == 1024) else None
    line = {
        "metric": "stencil_glups", "value": glups, "unit": "GLUPS", "n_gpus": world, "steps": args.steps,
        "warmup": warm, "ms_per_step": ms_total / args.steps, "higher_is_better": True,
        "scaling": "weak" if weak else "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_string(n, ny, weak),
                   "parallelism": "j-slabs x%d" % world, "l2": "fields (%.1f GB each) far larger than the 126 MB L2"
                   % (field_bytes * world / 1e9), "residual_last": out["res"], "itr": out["itr"],
                   "start_field": "hash of the global node index (cubez_b200/synth.py): the same problem at every N"},
        "windows_ms": windows, "windows": NWIN, "window_stat": "median",
        "e2e": {"value": e2e_glups, "unit": "GLUPS", "h2d_bytes_per_step": h2d_per_step,
                "d2h_bytes_per_step": d2h_per_step, "seconds": t_e2e,
                "note": "CZ.solve() with pinned host X,B: 2 fields H2D, K iterations, X + history D2H"},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic, "traffic_source": NCU_TRAFFIC_N1["source"] if traffic else None,
                     "peak_source": peak_src,
                     "kernel": "k_rb2_pk (fused two-colour sweep, TMA-staged, packed f32x2)",
                     "algorithmic_bytes_per_cell": ALGO_BYTES_PER_CELL, "ms_per_launch": ms_iter,
                     "ms_per_launch_fastest_rank": median(loops_min) / args.steps,
                     "frac_of_8000": achieved / 8000.0},
        "parity_vs_n1": parity,
        "clocks": clocks,
    }
    if also:
        line["also"] = also

    # ---- everything below can only ADD keys: a watchdog prints the line as it stands if a later leg hangs ----
    printed = threading.Event()

    def emit():
        if rank == 0 and not printed.is_set():
            printed.set()
            print(json.dumps(line), flush=True)

    def bail():
        emit()
        os._exit(0)
    dog = threading.Timer(float(os.environ.get("CZB_BENCH_TAIL_S", "420")), bail)
    dog.daemon = True
    dog.start()
    try:
        if world == 1 and not args.no_cpu and rank == 0:
            r = run_cpu_child("sample", 3, 1)
            if r:
                line["cpu_baseline"] = {"value": r["value"], "unit": "GLUPS", "cores": r["cores"],
                                        "kind": r.get("kind", "port"), "sample": r["sample"]}
        if not args.no_ttc and not args.no_also:
            if world == 1:
                ttc = time_to_convergence()
                cpu = run_cpu_child("ttc", timeout=400) if (ttc and not args.no_cpu) else None
            else:
                ttc = ttc_cases(make_cz, barrier, max_over_ranks, TTC_CASES[1:] + TTC_BIG, torch)
                cpu = None
            if ttc:
                attach_golden(ttc, cpu)
                if "also_f64" in ttc:
                    line["also_f64"] = ttc.pop("also_f64")
                if "e2e_to_convergence" in ttc:
                    line["e2e"]["to_convergence"] = ttc.pop("e2e_to_convergence")
                line["time_to_convergence"] = ttc
    except Exception as ex:      # never lose the headline
        sys.stderr.write("bench.py: optional legs failed: %r\n" % (ex,))
    dog.cancel()
    emit()
    if dist is not None:
        dist.destroy_process_group()


def attach_golden(ttc, cpu):
    """Beside every time-to-convergence entry: the oracle's iteration count, the reference program's own (both committed
    goldens) and the CPU child's record for the same configuration."""
    gold = load_golden_configs()
    gold_ref = load_golden_configs("f90c")
    for label, nn, solver, itmax, omg, pre, real in TTC_CASES + TTC_BIG:
        if label not in ttc:
            continue
        g = gold.get(golden_key(real, nn, solver, omg, pre))
        if g:
            ttc[label]["oracle_iterations"] = int(g["itr"])
            ttc[label]["oracle_residual"] = float(g["res"])
        g = gold_ref.get(golden_key(real, nn, solver, omg, pre))
        if g:       # the reference program's own run (made in the build container)
            ttc[label]["reference_iterations"] = int(g["itr"])
            ttc[label]["reference_residual"] = float(g["res"])
            ttc[label]["reference_seconds_build_container"] = {"seconds": g["wall_seconds_this_container"], "threads": g["threads"]}
        if cpu and label in cpu:
            ttc[label]["cpu"] = cpu[label]
    return ttc


def ttc_cases(make_cz, barrier, max_over_ranks, cases, torch):
    """Solve each case to res < 1e-5 from the reference's start field (P = 0 + Dirichlet faces, cz_Evaluate.cpp:375-390);
    time = the library's own CUDA events around the solver loop, max over ranks."""
    out = {}
    for label, n, solver, itmax, omg, pre, real in cases:
        cz = make_cz([n] * 3, real, eps=1.0e-5)
        P, RHS = cz.setup_laplace()
        cz.solve(solver, P, RHS, 3, omg, precond=pre, want_history=False)          # warm-up (tables, work fields)
        P, RHS = cz.setup_laplace()
        barrier()
        o = cz.solve(solver, P, RHS, itmax, omg, precond=pre, want_history=False)
        barrier()
        ms = max_over_ranks(float(o["ms"]))
        it = int(o["itr"])
        out[label] = {"n": n, "solver": solver, "precond": pre, "omega": omg, "real": real, "iterations": it,
                      "residual": float(o["res"]), "converged": bool(0 < it and o["res"] < 1.0e-5),
                      "solver_loop_ms": ms, "ms_per_iteration": ms / max(it, 1)}
        cz.close()
        del P, RHS
        torch.cuda.empty_cache()
    return out


# double precision on configs[2]'s grid (BASELINE.json: "512^3 float and double"): (label, solver, precond, omega, B/cell, k)
F64_ALSO = (("sor2sma", "sor2sma", None, 1.5, 32.0, 12), ("pbicgstab+sor2sma", "pbicgstab", "sor2sma", 1.5, 680.0, 6),
            ("jacobi", "jacobi", None, 0.8, 24.0, 12), ("pcr_rb", "pcr_rb", None, 1.5, 40.0, 8))


def ttc_worker():
    """Child process of time_to_convergence() at N = 1: prints one JSON object."""
    import torch
    import cubez_b200
    from cubez_b200 import synth

    def make_cz(gsz, real="f32", eps=1.0e-5):
        return cubez_b200.CZ(gsz, real, eps=eps)

    def barrier():
        torch.cuda.synchronize()
    out = ttc_cases(make_cz, barrier, lambda x: x, TTC_CASES + TTC_BIG, torch)
    # f64 throughput at 512^3, device resident
    f64 = {}
    n = 512
    cz = cubez_b200.CZ([n] * 3, "f64", eps=0.0)
    cells = float(cz.inner_cells_global)
    peak = measured_peak()[0]
    for label, solver, pre, omg, bpc, k in F64_ALSO:
        P, RHS = cz.setup_laplace()
        synth.fill(P, cz.sz)
        cz.bc_k(P)
        cz.solve(solver, P, RHS, 3, omg, precond=pre, want_history=False)
        ms = []
        for w in range(3):
            torch.cuda.synchronize()
            o = cz.solve(solver, P, RHS, k, omg, precond=pre, want_history=False)
            ms.append(float(o["ms"]) / ((k - 1) if solver == "pbicgstab" else k))
        m = sorted(ms)[1]
        gl = cells / (m * 1e-3) / 1e9
        f64[label] = {"value": gl, "unit": "G cell-iterations/s", "n": n, "ms_per_iteration": m,
                      "algorithmic_bytes_per_cell": bpc, "hbm_frac_of_measured": gl * bpc / peak}
        del P, RHS
    cz.close()
    out["also_f64"] = f64
    # end to end TO CONVERGENCE: the public call with pinned HOST arrays (H2D of X and B, the solver loop until
    # res < 1e-5, D2H of X and of the history) timed on the host clock -- the API-level figure at an iteration count a
    # user runs, next to the 20-step `e2e` that mostly measures the PCIe link
    e2c = {}
    for label, n, solver, itmax, omg, pre, real in (TTC_CASES[1], TTC_BIG[0]):
        cz = cubez_b200.CZ([n] * 3, real, eps=1.0e-5)
        P, RHS = cz.setup_laplace()
        cz.solve(solver, P, RHS, 3, omg, precond=pre, want_history=False)          # warm-up (work fields, tables)
        P, RHS = cz.setup_laplace()
        Ph, Bh = P.cpu().pin_memory(), RHS.cpu().pin_memory()
        del P, RHS
        torch.cuda.empty_cache()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        o = cz.solve(solver, Ph.numpy(), Bh.numpy(), itmax, omg, precond=pre, want_history=True)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        it = int(o["itr"])
        fb = float(Ph.numel()) * Ph.element_size()
        e2c[label] = {"iterations": it, "seconds": dt, "solver_loop_ms": float(o["ms"]), "residual": float(o["res"]),
                      "value": float(cz.inner_cells_global) * it / dt / 1e9, "unit": "GLUPS",
                      "h2d_bytes": 2.0 * fb, "d2h_bytes": fb + 8.0 * it}
        cz.close()
        del Ph, Bh
    out["e2e_to_convergence"] = e2c
    print(json.dumps(out), flush=True)


def time_to_convergence(timeout=300):
    """Runs ttc_worker() in a child process; None on any failure (the headline line must not depend on it)."""
    try:
        p = subprocess.run([sys.executable, os.path.abspath(__file__), "--ttc-worker"], capture_output=True, text=True,
                           timeout=timeout, env=dict(os.environ, CUDA_VISIBLE_DEVICES=os.environ.get("CUDA_VISIBLE_DEVICES", "0")))
        if p.returncode != 0:
            return None
        for ln in reversed(p.stdout.strip().splitlines()):
            if ln.startswith("{"):
                d = json.loads(ln)
                return d if isinstance(d, dict) and d else None
    except Exception:
        pass
    return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="cubez_b200")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-also", action="store_true", help="skip the PBiCGstab / Jacobi / line-SOR side measurements")
    ap.add_argument("--no-ttc", action="store_true", help="skip the time-to-convergence leg")
    ap.add_argument("--weak", action="store_true", help="weak scaling: 1024 x (1024*N) x 1024 instead of the fixed cube")
    ap.add_argument("--ttc-worker", action="store_true", help=argparse.SUPPRESS)
    ap.add_argument("--cpu-worker", default=None, help=argparse.SUPPRESS)
    args = ap.parse_args()
    if args.ttc_worker:
        ttc_worker()
        return
    if args.cpu_worker:
        cpu_worker(args.cpu_worker, args.steps, args.warmup)
        return
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    if world == 1 and args.gpus > 1:
        # convenience: re-launch under torchrun, one process per GPU
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(args.gpus),
               "--master-addr", "127.0.0.1", "--master-port", os.environ.get("MASTER_PORT", "29511"),
               os.path.abspath(__file__), "--gpus", str(args.gpus), "--steps", str(args.steps), "--warmup", str(args.warmup)]
        cmd += ["--weak"] if args.weak else []
        cmd += ["--no-also"] if args.no_also else []
        cmd += ["--no-ttc"] if args.no_ttc else []
        sys.exit(subprocess.call(cmd))
    run_gpu(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
