#!/usr/bin/env python
"""scratch/solver_bench.py SOLVER N ITERS [PRECOND] [REAL] [OMEGA] -- device-resident timing of any solver
(random interior state so that no iteration converges early; eps = 0)."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import cubez_b200

solver = sys.argv[1]; n = int(sys.argv[2]); iters = int(sys.argv[3])
precond = sys.argv[4] if len(sys.argv) > 4 and sys.argv[4] != "-" else None
real = sys.argv[5] if len(sys.argv) > 5 else "f32"
omega = float(sys.argv[6]) if len(sys.argv) > 6 else 1.5
cz = cubez_b200.CZ([n] * 3, real, device=0, eps=0.0)
P, RHS = cz.setup_laplace()
gen = torch.Generator(device="cuda").manual_seed(1234)
P[2:-2, 2:-2, 2:-2] = torch.rand(P[2:-2, 2:-2, 2:-2].shape, generator=gen, device="cuda", dtype=P.dtype) * 2 - 1
cz.bc_k(P)
cz.solve(solver, P, RHS, 3, omega, precond=precond, want_history=False)
torch.cuda.synchronize()
l0 = cz.launch_count()
out = cz.solve(solver, P, RHS, iters, omega, precond=precond, want_history=False)
torch.cuda.synchronize()
done = out["itr"] if solver != "pbicgstab" else iters - 1
done = max(min(done, iters), 1)
ms = out["ms"] / done
cells = float(cz.inner_cells_global)
R = 4 if real == "f32" else 8
print(json.dumps({"solver": solver, "precond": precond, "n": n, "real": real, "iters_run": done, "ms_per_iter": ms,
                  "glups": cells / ms / 1e6, "res": out["res"], "launches_per_iter": (cz.launch_count() - l0) / done}))
