#!/usr/bin/env python
"""scratch/call_overhead.py NX NY NZ SOLVER -- wall time of czb_solve calls with K iterations against the library's own
loop time: what one call costs beyond its iterations (host enqueue latency, initial exchanges, final synchronisation)."""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, cubez_b200
from cubez_b200 import synth
nx, ny, nz = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]); solver = sys.argv[4]
cz = cubez_b200.CZ([nx, ny, nz], "f32", device=0, eps=0.0)
P, RHS = cz.setup_laplace(); synth.fill(P, cz.sz); cz.bc_k(P)
cz.solve(solver, P, RHS, 4, 1.5, want_history=False)
out = {}
for k in (1, 2, 20, 40):
    torch.cuda.synchronize(); t0 = time.perf_counter(); n = 30; loop = 0.0
    for _ in range(n):
        o = cz.solve(solver, P, RHS, k, 1.5, want_history=False); loop += o["ms"]
    torch.cuda.synchronize(); wall = (time.perf_counter() - t0) / n * 1e3
    out[k] = {"wall_ms_per_call": round(wall, 4), "loop_ms_per_call": round(loop / n, 4), "outside_loop_ms": round(wall - loop / n, 4)}
print(json.dumps({"grid": [nx, ny, nz], "solver": solver, "calls": out}))
