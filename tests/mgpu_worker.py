"""Multi-GPU parity worker (run under torchrun, one process per GPU):
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 tests/mgpu_worker.py
Every rank solves its j-slab with libczb200 (NCCL halo exchange + all-reduced residual); rank 0
also solves the single-domain problem on its GPU and checks that the gathered slabs reproduce it:
bit-exact fields for the stationary solvers (updates are order-independent within a colour),
residual histories to reduction-order tolerance, iteration counts equal."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cubez_b200  # noqa: E402


def main():
    rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    n = int(os.environ.get("CZB_MGPU_N", "64"))
    n2 = int(os.environ.get("CZB_MGPU_N2", "200"))   # long lines: tensor-memory line-SOR kernels on slabs
    # last field: start from the global-index hash field (cubez_b200/synth.py) instead of P = 0 -> the halos matter
    # from the first sweep on; eps = 0 there, so every iteration runs
    cases = [(n, "sor2sma", None, 1.5, 60, 0, 0), (n, "sor2sma", None, 1.5, 60, 1, 0), (n, "jacobi", None, 0.8, 40, 0, 0),
             (n, "pcr_rb", None, 1.5, 30, 0, 0), (n, "pbicgstab", "sor2sma", 1.1, 30, 0, 0),
             (n2, "pcr_rb", None, 1.5, 8, 0, 0), (n2, "pcr_rb_esa", None, 1.5, 8, 0, 0), (n2, "sor2sma", None, 1.5, 12, 0, 0),
             (n, "sor2sma", None, 1.5, 37, 0, 1), (n2, "sor2sma", None, 1.8, 21, 0, 1), (n2, "jacobi", None, 0.8, 9, 0, 1),
             (n2, "pcr_rb", None, 1.5, 5, 0, 1), (n, "sor2sma", None, 1.5, 400, 0, 2)]
    ok = True
    from cubez_b200 import synth
    for n, solver, pc, omega, iters, variant, start in cases:
        cz = cubez_b200.CZ([n, n, n], "f32", device=lr, nranks=world, rank=rank, eps=1e-5 if solver == "pbicgstab" else 0.0)
        cz.variant(variant)
        idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            idt.copy_(torch.frombuffer(bytearray(cz.comm_unique_id()), dtype=torch.uint8))
        dist.broadcast(idt, 0)
        cz.comm_init(idt.cpu().numpy().tobytes())
        P, RHS = cz.setup_laplace()
        if start == 1:
            synth.fill(P, cz.sz, head_j=cz.head[1], gny=n)
            cz.bc_k(P)
        if start == 2:          # converges inside the run: the ranks must agree on the iteration it stops at
            cz.dll.czb_ctx_set_eps(cz.ctx, 1.0e-5)
        out = cz.solve(solver, P, RHS, iters, omega, pc)
        if start == 1:          # a second solve on the same context continues from the first (inbox / counters reused)
            cz.solve(solver, P, RHS, 3, omega, pc)
        own = P[cz.idx[2] + 1:cz.idx[3] + 2].contiguous()          # owned inner planes (local C indices)
        sizes = [None] * world
        dist.all_gather_object(sizes, own.shape[0])
        if rank == 0:
            parts = [own]
            for r in range(1, world):
                t = torch.empty((sizes[r],) + tuple(own.shape[1:]), dtype=own.dtype, device="cuda")
                dist.recv(t, r); parts.append(t)
            full = torch.cat(parts, 0).cpu().numpy()
            c1 = cubez_b200.CZ([n, n, n], "f32", device=lr, eps=1e-5 if solver == "pbicgstab" else 0.0)
            c1.variant(variant)
            P1, R1 = c1.setup_laplace()
            if start == 1:
                synth.fill(P1, c1.sz)
                c1.bc_k(P1)
            if start == 2:
                c1.dll.czb_ctx_set_eps(c1.ctx, 1.0e-5)
            o1 = c1.solve(solver, P1, R1, iters, omega, pc)
            if start == 1:
                c1.solve(solver, P1, R1, 3, omega, pc)
            ref = P1[3:-3].cpu().numpy()
            same_itr = o1["itr"] == out["itr"]
            m = min(len(o1["history"]), len(out["history"]))
            if solver == "pbicgstab":
                good = abs(o1["itr"] - out["itr"]) <= 1 and np.allclose(full, ref, rtol=0, atol=1e-5 * np.abs(ref).max())
                good = good and np.allclose(out["history"][:m - 1], o1["history"][:m - 1], rtol=1e-3)
            else:
                good = same_itr and np.array_equal(full, ref) and np.allclose(out["history"][:m], o1["history"][:m], rtol=1e-12)
            print("mgpu n=%d %-10s pc=%-8s variant=%d start=%d ranks=%d itr %d/%d maxdiff %.3e %s" % (
                n, solver, pc, variant, start, world, out["itr"], o1["itr"], float(np.abs(full - ref).max()),
                "OK" if good else "FAIL"), flush=True)
            ok = ok and good
            c1.close()
        else:
            dist.send(own, 0)
        dist.barrier()
        cz.close()
    # ---- 3-D bricks: i / k faces travel packed, per-colour sweeps with an exchange after each colour,
    # Jacobi and the BiCGstab operator on the fused one-plane kernels with guard rows / cells from the neighbours ----
    from cubez_b200 import decomp
    divisions = {2: [(2, 1, 1), (1, 1, 2)], 4: [(2, 1, 2), (2, 2, 1)], 8: [(2, 2, 2), (4, 1, 2)]}.get(world, [])
    bcases = [(n, "sor2sma", None, 1.5, 60), (n, "jacobi", None, 0.8, 60), (n, "pbicgstab", "sor2sma", 1.1, 30),
              (n2, "jacobi", None, 0.8, 12), (n2, "sor2sma", None, 1.7, 12)]
    for gdv in divisions:
        for n, solver, pc, omega, iters in bcases:
            eps = 1e-5 if solver == "pbicgstab" else 0.0
            cz = cubez_b200.CZ([n, n, n], "f32", device=lr, nranks=world, rank=rank, eps=eps, gdv=gdv)
            idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
            if rank == 0:
                idt.copy_(torch.frombuffer(bytearray(cz.comm_unique_id()), dtype=torch.uint8))
            dist.broadcast(idt, 0)
            cz.comm_init(idt.cpu().numpy().tobytes())
            P, RHS = cz.setup_laplace()
            out = cz.solve(solver, P, RHS, iters, omega, pc)
            ix = cz.idx
            own = P[ix[2] + 1:ix[3] + 2, ix[0] + 1:ix[1] + 2, ix[4] + 1:ix[5] + 2].contiguous()
            if rank == 0:
                full = torch.full((n - 2, n - 2, n - 2), float("nan"), dtype=own.dtype, device="cuda")
                for r in range(world):
                    d = decomp.brick_decompose([n, n, n], list(gdv), r)
                    shp = (d["idx"][3] - d["idx"][2] + 1, d["idx"][1] - d["idx"][0] + 1, d["idx"][5] - d["idx"][4] + 1)
                    if r == 0:
                        t = own
                    else:
                        t = torch.empty(shp, dtype=own.dtype, device="cuda"); dist.recv(t, r)
                    j0, i0, k0 = d["head"][1] + d["idx"][2] - 3, d["head"][0] + d["idx"][0] - 3, d["head"][2] + d["idx"][4] - 3
                    full[j0:j0 + shp[0], i0:i0 + shp[1], k0:k0 + shp[2]] = t
                full = full.cpu().numpy()
                c1 = cubez_b200.CZ([n, n, n], "f32", device=lr, eps=eps)
                P1, R1 = c1.setup_laplace()
                o1 = c1.solve(solver, P1, R1, iters, omega, pc)
                ref = P1[3:-3, 3:-3, 3:-3].cpu().numpy()
                m = min(len(o1["history"]), len(out["history"]))
                if solver == "pbicgstab":
                    good = abs(o1["itr"] - out["itr"]) <= 1 and np.allclose(full, ref, rtol=0, atol=1e-5 * np.abs(ref).max())
                    good = good and np.allclose(out["history"][:m - 1], o1["history"][:m - 1], rtol=1e-3)
                else:
                    good = o1["itr"] == out["itr"] and np.array_equal(full, ref) and np.allclose(out["history"][:m], o1["history"][:m], rtol=1e-12)
                print("mgpu n=%d %-10s pc=%-8s bricks %dx%dx%d itr %d/%d maxdiff %.3e %s" % (
                    n, solver, pc, gdv[0], gdv[1], gdv[2], out["itr"], o1["itr"], float(np.nanmax(np.abs(full - ref))),
                    "OK" if good else "FAIL"), flush=True)
                ok = ok and good
                c1.close()
            else:
                dist.send(own, 0)
            dist.barrier()
            cz.close()
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.broadcast(flag, 0)
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) == 1 else 1)


if __name__ == "__main__":
    main()
