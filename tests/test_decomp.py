"""CPU tests of the multi-rank host logic: j-slab decomposition (C++ vs its
Python mirror), and a world_size-2 gloo run of red-black SOR / Jacobi over slabs with halo
exchange + all-reduced residual that must reproduce the single-domain oracle bit for bit.
The compute in the gloo run is the CPU oracle (test infrastructure); what is under test is the
decomposition, the exchange pattern (one plane after each colour) and the global colouring."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

import oracle_util as ou
from cubez_b200 import decomp, lib

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.parametrize("gsz", [[64, 64, 64], [16, 9, 12], [1024, 1024, 1024], [32, 35, 32]])
def test_slab_decompose_cpp_matches_python_and_tiles_the_cube(gsz):
    d = lib.load("f32")
    sz, head, nid, idx = (C.c_int * 3)(), (C.c_int * 3)(), (C.c_int * 6)(), (C.c_int * 6)()
    for R in (1, 2, 3, 4, 7, 8):
        if gsz[1] - 2 < R:
            assert d.czb_slab_decompose((C.c_int * 3)(*gsz), R, 0, sz, head, nid, idx) == -1
            with pytest.raises(ValueError):
                decomp.slab_decompose(gsz, R, 0)
            continue
        covered, cells = [], 0
        for r in range(R):
            assert d.czb_slab_decompose((C.c_int * 3)(*gsz), R, r, sz, head, nid, idx) == 0
            p = decomp.slab_decompose(gsz, R, r)
            assert (list(sz), list(head), list(nid), list(idx)) == (p["sz"], p["head"], p["nID"], p["idx"])
            # owned inner planes in global numbering
            g0, g1 = p["head"][1] + p["idx"][2] - 1, p["head"][1] + p["idx"][3] - 1
            covered += list(range(g0, g1 + 1))
            cells += decomp.inner_cells(p["idx"])
            assert p["nID"][2] == (r - 1 if r > 0 else -1) and p["nID"][3] == (r + 1 if r < R - 1 else -1)
        assert covered == list(range(2, gsz[1]))            # every inner plane exactly once, in order
        assert cells == (gsz[0] - 2) * (gsz[1] - 2) * (gsz[2] - 2)
        sizes = [decomp.slab_decompose(gsz, R, r)["idx"][3] - decomp.slab_decompose(gsz, R, r)["idx"][2] for r in range(R)]
        assert max(sizes) - min(sizes) <= 1                  # balanced


@pytest.mark.parametrize("gsz,gdv", [([64, 64, 64], [2, 2, 2]), ([16, 9, 12], [3, 1, 2]), ([1024, 1024, 1024], [1, 8, 1]),
                                     ([32, 35, 32], [2, 3, 1]), ([20, 20, 20], [1, 1, 4])])
def test_brick_decompose_cpp_matches_python_and_tiles_the_cube(gsz, gdv):
    """3-D bricks (reference: CBrick division, cz_Evaluate.cpp:103-159): C++ == Python mirror, the bricks
    tile the inner nodes exactly once, neighbour tables are mutual, gdv = {1, N, 1} is the j-slab decomposition."""
    _check_bricks(gsz, gdv)


def test_brick_decompose_on_random_grids_and_divisions():
    """The same properties on 40 seeded random (grid, division) pairs, incl. divisions that leave one inner node per brick."""
    rng = np.random.default_rng(11)
    for _ in range(40):
        gsz = [int(rng.integers(4, 40)) for _ in range(3)]
        gdv = [int(rng.integers(1, min(5, gsz[a] - 2) + 1)) for a in range(3)]
        _check_bricks(gsz, gdv)
    _check_bricks([6, 5, 7], [4, 3, 5])                        # one inner node per brick along every axis


def _check_bricks(gsz, gdv):
    d = lib.load("f32")
    sz, head, nid, idx = (C.c_int * 3)(), (C.c_int * 3)(), (C.c_int * 6)(), (C.c_int * 6)()
    R = gdv[0] * gdv[1] * gdv[2]
    seen = np.zeros((gsz[0] + 1, gsz[1] + 1, gsz[2] + 1), np.int32)
    parts = []
    for r in range(R):
        assert d.czb_brick_decompose((C.c_int * 3)(*gsz), (C.c_int * 3)(*gdv), r, sz, head, nid, idx) == 0
        p = decomp.brick_decompose(gsz, gdv, r)
        assert (list(sz), list(head), list(nid), list(idx)) == (p["sz"], p["head"], p["nID"], p["idx"])
        parts.append(p)
        g = [(p["head"][a] + p["idx"][2 * a] - 1, p["head"][a] + p["idx"][2 * a + 1] - 1) for a in range(3)]
        seen[g[0][0]:g[0][1] + 1, g[1][0]:g[1][1] + 1, g[2][0]:g[2][1] + 1] += 1
        if gdv[0] == 1 and gdv[2] == 1:
            assert p == decomp.slab_decompose(gsz, R, r)
    assert (seen[2:gsz[0], 2:gsz[1], 2:gsz[2]] == 1).all() and seen.sum() == (gsz[0] - 2) * (gsz[1] - 2) * (gsz[2] - 2)
    for r, p in enumerate(parts):
        for f in range(6):
            nb = p["nID"][f]
            if nb >= 0:
                assert parts[nb]["nID"][f ^ 1] == r        # my X+ neighbour has me as its X-
                a = f // 2                                  # adjacent along axis a: its first node follows my last
                lo, hi = (parts[nb], p) if f % 2 == 0 else (p, parts[nb])
                assert lo["head"][a] + lo["idx"][2 * a + 1] == hi["head"][a] + hi["idx"][2 * a] - 1
    assert d.czb_brick_decompose((C.c_int * 3)(*gsz), (C.c_int * 3)(gsz[0], 1, 1), 0, sz, head, nid, idx) == -1
    with pytest.raises(ValueError):
        decomp.brick_decompose(gsz, [gsz[0], 1, 1], 0)


def _worker(rank, world, port, n, solver, iters, omega, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import torch
    p = decomp.slab_decompose([n, n, n], world, rank)
    o = ou.OracleCZ([n, n, n], "f32", sz=p["sz"], idx=p["idx"], nID=p["nID"], head=p["head"])
    X, B = o.setup_laplace()
    rng = np.random.default_rng(7)
    full = rng.uniform(-1, 1, size=(n + 4, n + 4, n + 4)).astype(np.float32)
    j0 = p["head"][1] - 1                     # global C index of local C index 0
    # Dirichlet data from the single-domain bc_k (the reference's origin + dh*(j_local-1) rounds
    # differently per slab; the product uses global j for the same reason, czb_ctx_bc_k)
    og = ou.OracleCZ([n, n, n], "f32"); og.bc_k(full)
    X[...] = full[j0:j0 + X.shape[0]]
    o.ac1 = np.float32(omega)
    plane = X.shape[1] * X.shape[2]
    lo, hi = p["nID"][2], p["nID"][3]
    jx = p["sz"][1]

    def halo(F, w=1):   # Comm_S: guard planes <- neighbour's edge planes (cubez_b200/csrc/comm.cpp)
        reqs = []
        if lo >= 0:
            reqs.append(dist.isend(torch.from_numpy(F[2:2 + w].copy()), lo))
            rl = torch.empty((w,) + F.shape[1:], dtype=torch.float32); reqs.append(dist.irecv(rl, lo))
        if hi >= 0:
            reqs.append(dist.isend(torch.from_numpy(F[jx + 2 - w:jx + 2].copy()), hi))
            rh = torch.empty((w,) + F.shape[1:], dtype=torch.float32); reqs.append(dist.irecv(rh, hi))
        for r in reqs:
            r.wait()
        if lo >= 0:
            F[2 - w:2] = rl.numpy()
        if hi >= 0:
            F[jx + 2:jx + 2 + w] = rh.numpy()

    halo(X)
    ip = (p["head"][0] + p["head"][1] + p["head"][2] + 1) % 2          # cz_Poisson.cpp:181
    hist = []
    cf = o.cf
    for it in range(iters):
        res = C.c_double(0)
        if solver == "sor2sma":
            for color in (0, 1):
                flop = C.c_double(0); om = C.c_float(omega); cip = C.c_int(ip); cc = C.c_int(color)
                o.dll.psor2sma_core_(ou.ptr(X), o._sz, o._idx, o.g(), ou.ptr(cf), C.byref(cip), C.byref(cc), C.byref(om),
                                     ou.ptr(B), C.byref(res), C.byref(flop))
                halo(X)                       # exchange after EACH colour
        else:
            o.sweep_jacobi(X, B, res); halo(X)
        t = torch.tensor([res.value], dtype=torch.float64)
        dist.all_reduce(t)
        hist.append(float(np.sqrt(t.item() * o.res_normal)))
    q.put((rank, j0, p["idx"][2], p["idx"][3], X.copy(), hist))
    dist.destroy_process_group()


@pytest.mark.parametrize("solver", ["sor2sma", "jacobi"])
def test_two_rank_gloo_slabs_reproduce_single_domain(solver):
    import torch.multiprocessing as mp
    n, iters, omega, world = 20, 6, 1.3 if solver == "sor2sma" else 0.8, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + (os.getpid() % 200)
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, solver, iters, omega, q)) for r in range(world)]
    for p_ in procs:
        p_.start()
    outs = [q.get(timeout=120) for _ in range(world)]
    for p_ in procs:
        p_.join(timeout=60)
    # single-domain oracle on the same data
    o = ou.OracleCZ([n, n, n], "f32")
    X, B = o.setup_laplace()
    rng = np.random.default_rng(7)
    X[...] = rng.uniform(-1, 1, size=X.shape).astype(np.float32); o.bc_k(X)
    itr, res, hist = (o.stationary(solver, X, B, iters, omega))
    for rank, j0, jst, jed, Xl, h in outs:
        own = slice(jst + 1, jed + 2)                       # local C indices of owned inner planes
        glob = slice(j0 + jst + 1, j0 + jed + 2)
        assert np.array_equal(Xl[own, 2:-2, 2:-2], X[glob, 2:-2, 2:-2]), "rank %d field differs" % rank
        np.testing.assert_allclose(h, hist[:iters], rtol=1e-12)
