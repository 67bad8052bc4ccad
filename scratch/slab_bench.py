#!/usr/bin/env python
"""scratch/slab_bench.py NX NY NZ SOLVER ITERS -- single-rank timing of a non-cubic grid (emulates the per-rank
slab of a multi-GPU run without communication)."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, cubez_b200
nx, ny, nz = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]); solver = sys.argv[4]; iters = int(sys.argv[5])
pre = sys.argv[6] if len(sys.argv) > 6 else None
cz = cubez_b200.CZ([nx, ny, nz], "f32", device=0, eps=0.0)
P, RHS = cz.setup_laplace()
gen = torch.Generator(device="cuda").manual_seed(1)
P[2:-2, 2:-2, 2:-2] = torch.rand(P[2:-2, 2:-2, 2:-2].shape, generator=gen, device="cuda") * 2 - 1
cz.bc_k(P)
cz.solve(solver, P, RHS, 4, 1.5, precond=pre, want_history=False)
out = cz.solve(solver, P, RHS, iters, 1.5, precond=pre, want_history=False)
done = iters - 1 if solver == "pbicgstab" else iters
ms = out["ms"] / done
print(json.dumps({"grid": [nx, ny, nz], "solver": solver, "ms_per_iter": ms, "glups": cz.inner_cells_global / ms / 1e6,
                  "env": {k: v for k, v in os.environ.items() if k.startswith("CZB_")}}))
