#!/usr/bin/env python
"""bench.py -- headline benchmark of cubez_b200 (BASELINE.json: "Stencil GLUPS ... HBM GB/s vs peak").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N --steps K --warmup W

Workload (config.workload): red-black SOR ("sor2sma", both colours = one step) on the 1024^3
float cube of BASELINE.json configs[4]; it fits one B200, so it is also the N=1 workload.
N>1 = STRONG scaling: the same cube cut into N j-slabs, one process per GPU, halo planes
exchanged with NCCL, residual all-reduced (cubez_b200/csrc/comm.cpp).

One JSON line on rank 0:
  value        GLUPS (1e9 lattice updates/s, whole job), fields resident in HBM, K iterations timed
               with CUDA events on the launching stream, max over ranks
  e2e          the same K iterations through the public call CZ.solve() with HOST (pinned) arrays:
               H2D of X and B, the solve loop, D2H of X and of the residual history, all timed
  roofline     dominant kernel (fused two-colour sweep k_rb2_pk): algorithmic bytes
               (4R = 16 B per cell.iteration) / measured time per iteration,
               against the measured HBM copy bandwidth of MEASURED_PEAKS.json
  cpu_baseline the CPU oracle port of the reference kernel (oracle/, "port") on all host cores,
               bounded sample, rank 0 / N=1 only
  time_to_convergence  (N=1 only, optional) the metric's second half: BASELINE.json configs[0..3] solved to
               res < 1e-5 in a CHILD process after the timed regions (iterations, device time of the solver loop);
               left out if the child fails or times out -- it can never cost the headline line
--impl reference times that CPU implementation alone (all host threads) and prints the same line.
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
NCU_TRAFFIC_PER_ITER = 14.24e9     # dram read+write bytes per k_rb2_pk launch (profiles/r01_rb2_summary.md, 1 GPU)


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm = sorted(float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit())
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 6 for i in range(4) if r[2 + i].startswith("Active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def cpu_port_glups(steps, warmup, sample_planes=128, n=N_CUBE):
    """Time the oracle port of psor2sma_core (the reference's CPU kernel restated in C++,
    oracle/cz_f90_port.cpp) on all host cores: a j-slab of the workload per step."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import numpy as np
    import oracle_util as ou
    cores = os.cpu_count() or 1
    os.environ.setdefault("OMP_NUM_THREADS", str(cores))
    os.environ.setdefault("OMP_PROC_BIND", "close")
    sz = [n, min(sample_planes, n), n]
    o = ou.OracleCZ([n, n, n], "f32", sz=sz)
    X, B = o.setup_laplace()
    rng = np.random.default_rng(1)
    X[2:-2, 2:-2, 2:-2] = rng.uniform(-1, 1, size=X[2:-2, 2:-2, 2:-2].shape).astype(np.float32)
    o.ac1 = np.float32(OMEGA)
    cells = (sz[0] - 2) * (sz[1] - 2) * (sz[2] - 2)
    res = C.c_double(0)
    for _ in range(max(warmup, 1)):
        o.sweep_rb(X, B, res)
    t0 = time.perf_counter()
    for _ in range(steps):
        o.sweep_rb(X, B, res)
    dt = time.perf_counter() - t0
    return cells * steps / dt / 1e9, dt / steps * 1e3, cores, \
        "%d iterations of sor2sma on a %dx%dx%d slab of the %d^3 grid (oracle port, OpenMP %d threads)" % (
            steps, sz[0], sz[1], sz[2], n, cores)


def run_reference(args, rank):
    if rank != 0:
        return
    v, ms, cores, sample = cpu_port_glups(args.steps, args.warmup)
    print(json.dumps({
        "impl": "reference", "metric": "stencil_glups", "value": v, "unit": "GLUPS", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "sor2sma %d^3 f32, omega=%.2f (CPU: j-slab sample per step)" % (N_CUBE, OMEGA)},
        "cpu_baseline": {"value": v, "unit": "GLUPS", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "GLUPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }), flush=True)


def run_gpu(args, rank, world, local_rank):
    import numpy as np
    import torch
    import cubez_b200

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
    cz = cubez_b200.CZ([n, ny, n], "f32", device=local_rank, nranks=world, rank=rank, eps=0.0)  # eps=0: never stops early
    if world > 1:
        idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            idt.copy_(torch.frombuffer(bytearray(cz.comm_unique_id()), dtype=torch.uint8))
        dist.broadcast(idt, 0)
        cz.comm_init(bytes(idt.cpu().numpy().tobytes()))

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

    # ---- device-resident: fields stay in HBM ---------------------------------------------------
    P, RHS = cz.setup_laplace()
    gen = torch.Generator(device="cuda").manual_seed(1234 + rank)
    P[2:-2, 2:-2, 2:-2] = torch.rand(P[2:-2, 2:-2, 2:-2].shape, generator=gen, device="cuda") * 2 - 1   # synthetic state
    cz.bc_k(P)
    cz.solve(SOLVER, P, RHS, max(args.warmup, 3), OMEGA, want_history=False)     # warm-up (>= 3 steps)
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
    l0 = cz.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    cz.sync_stream()
    ev0.record()
    out = cz.solve(SOLVER, P, RHS, args.steps, OMEGA, want_history=False)
    ev1.record()
    barrier()
    ms_total = max_over_ranks(ev0.elapsed_time(ev1))
    ms_loop = max_over_ranks(out["ms"])           # the library's own events around the iteration loop
    launches = cz.launch_count() - l0
    clocks = sampler.stop() if sampler else None
    cells = float(cz.inner_cells_global)
    glups = cells * args.steps / (ms_total * 1e-3) / 1e9
    ms_iter = ms_loop / args.steps

    # ---- the other solvers of the north-star target on the same grid (device resident, after the
    # headline region; reported under "also", not part of value/e2e) -------------------------------
    also = {}
    if not args.no_also:
        for name, pre, omg, bpc, k in (("pbicgstab", "sor2sma", OMEGA, 340.0, 6), ("jacobi", None, 0.8, 12.0, 10),
                                       ("pcr_rb", None, OMEGA, 20.0, 8)):
            P2, R2 = cz.setup_laplace()
            P2[2:-2, 2:-2, 2:-2] = torch.rand(P2[2:-2, 2:-2, 2:-2].shape, generator=gen, device="cuda") * 2 - 1
            cz.bc_k(P2)
            cz.solve(name, P2, R2, 3, omg, precond=pre, want_history=False)
            barrier()
            o2 = cz.solve(name, P2, R2, k, omg, precond=pre, want_history=False)
            barrier()
            done = (k - 1) if name == "pbicgstab" else k
            ms2 = max_over_ranks(o2["ms"]) / done
            gl = cells / (ms2 * 1e-3) / 1e9
            also[name + ("+" + pre if pre else "")] = {
                "value": gl, "unit": "G cell-iterations/s", "ms_per_iteration": ms2,
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

    if rank != 0:
        return
    peak, peak_src = measured_peak()
    achieved = (cells / world) * ALGO_BYTES_PER_CELL / (ms_iter * 1e-3) / 1e9      # GB/s per GPU
    cpu = None
    if world == 1 and not args.no_cpu:
        v, cms, cores, sample = cpu_port_glups(3, 1)
        cpu = {"value": v, "unit": "GLUPS", "cores": cores, "kind": "port", "sample": sample}
    line = {
        "metric": "stencil_glups", "value": glups, "unit": "GLUPS", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_total / args.steps, "higher_is_better": True,
        "scaling": "weak" if weak else "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": ("sor2sma %dx%dx%d f32 (j extended with the GPU count), omega=%.2f, both colours per step" % (n, ny, n, OMEGA))
                   if weak else "sor2sma %d^3 f32, omega=%.2f, both colours per step" % (n, OMEGA),
                   "parallelism": "j-slabs x%d" % world, "l2": "fields (%.1f GB each) far larger than the 126 MB L2"
                   % (field_bytes * world / 1e9), "residual_last": out["res"], "itr": out["itr"]},
        "e2e": {"value": e2e_glups, "unit": "GLUPS", "h2d_bytes_per_step": h2d_per_step,
                "d2h_bytes_per_step": d2h_per_step, "seconds": t_e2e,
                "note": "CZ.solve() with pinned host X,B: 2 fields H2D, K iterations, X + history D2H"},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": NCU_TRAFFIC_PER_ITER, "peak_source": peak_src,
                     "kernel": "k_rb2_pk (fused two-colour sweep, TMA-staged, packed f32x2)",
                     "algorithmic_bytes_per_cell": ALGO_BYTES_PER_CELL, "ms_per_launch": ms_iter,
                     "frac_of_8000": achieved / 8000.0},
        "clocks": clocks,
    }
    if cpu:
        line["cpu_baseline"] = cpu
    if also:
        line["also"] = also
    if world == 1 and not args.no_ttc and not args.no_also:
        ttc = time_to_convergence()
        if ttc:
            line["time_to_convergence"] = ttc
    print(json.dumps(line), flush=True)
    cz.close()
    if dist is not None:
        dist.destroy_process_group()


# BASELINE.json configs[0..3] to convergence (eps = 1e-5, cz.h:162): (label, N, solver, ItrMax, omega, precond)
TTC_CASES = (("jacobi 64^3 w=0.8", 64, "jacobi", 4000, 0.8, None),
             ("sor2sma 256^3 w=1.95", 256, "sor2sma", 20000, 1.95, None),
             ("pbicgstab+sor2sma 512^3 w=1.1", 512, "pbicgstab", 400, 1.1, "sor2sma"),
             ("pcr_rb 512^3 w=1.9", 512, "pcr_rb", 8000, 1.9, None))


def ttc_worker():
    """Child process of time_to_convergence(): prints one JSON object."""
    import torch
    import cubez_b200
    out = {}
    for label, n, solver, itmax, omg, pre in TTC_CASES:
        cz = cubez_b200.CZ([n] * 3, "f32")
        P, RHS = cz.setup_laplace()
        cz.solve(solver, P, RHS, 3, omg, precond=pre, want_history=False)          # warm-up (tables, work fields)
        P, RHS = cz.setup_laplace()                                                # the reference's start: P = 0 + Dirichlet faces
        torch.cuda.synchronize()
        o = cz.solve(solver, P, RHS, itmax, omg, precond=pre, want_history=False)
        out[label] = {"n": n, "solver": solver, "precond": pre, "omega": omg, "iterations": int(o["itr"]),
                      "residual": float(o["res"]), "converged": bool(0 < o["itr"] and o["res"] < 1.0e-5),
                      "solver_loop_ms": float(o["ms"])}
        cz.close()
        del P, RHS
        torch.cuda.empty_cache()
    print(json.dumps(out), flush=True)


def time_to_convergence(timeout=180):
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
    ap.add_argument("--no-ttc", action="store_true", help="skip the time-to-convergence child process")
    ap.add_argument("--weak", action="store_true", help="weak scaling: 1024 x (1024*N) x 1024 instead of the fixed cube")
    ap.add_argument("--ttc-worker", action="store_true", help=argparse.SUPPRESS)
    args = ap.parse_args()
    if args.ttc_worker:
        ttc_worker()
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
        sys.exit(subprocess.call(cmd))
    run_gpu(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
