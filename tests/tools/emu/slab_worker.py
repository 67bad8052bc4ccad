"""TEST INFRASTRUCTURE: one rank of a j-slab run of the CPU-emulated library (tests/tools/emu), the CPU counterpart of
tests/mgpu_worker.py.  One OS process per rank; NCCL is the file-based stand-in nccl_emu.cpp, selected through
comm.cpp's CZB_NCCL_LIB switch.

    python slab_worker.py RANK NRANKS WORKDIR N SOLVER PRECOND|- OMEGA ITERS EPS [GX,GY,GZ]
writes WORKDIR/out_RANK.npz (owned inner planes -- with a division GX,GY,GZ the owned inner box and its global C offsets
-- iteration count, history)."""
import ctypes as C
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
import emu_build  # noqa: E402
from cubez_b200 import decomp, ffi  # noqa: E402


def main():
    rank, nranks, work, n = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3], int(sys.argv[4])
    solver, pre, omega, iters, eps = sys.argv[5], (None if sys.argv[6] == "-" else sys.argv[6]), float(sys.argv[7]), int(sys.argv[8]), float(sys.argv[9])
    os.environ["CZB_NCCL_LIB"] = emu_build.build_nccl()
    os.environ["CZB_NCCL_EMU_DIR"] = work
    dll = C.CDLL(emu_build.build_lib()["f32"])
    ffi.bind_reference_symbols(dll); ffi.bind_czb_symbols(dll, C.c_float)
    dll.czb_init(0)
    gsz = [n, n, n]
    gdv = [int(v) for v in sys.argv[10].split(",")] if len(sys.argv) > 10 else None
    d = decomp.brick_decompose(gsz, gdv, rank) if gdv else decomp.slab_decompose(gsz, nranks, rank)
    sz, head, nID, idx = d["sz"], d["head"], d["nID"], d["idx"]
    dt = np.float32
    pitch = dt(1.0) / dt(gsz[2] - 1)
    origin = [dt(0) + dt(head[a] - 1) * pitch for a in range(3)]
    ia = lambda v: (C.c_int * len(v))(*[int(x) for x in v])
    ctx = dll.czb_ctx_create(ia(sz), ia(idx), ia(nID), C.c_float(pitch), (C.c_float * 3)(*origin))
    dll.czb_ctx_set_eps(ctx, eps)
    dll.czb_ctx_set_head(ctx, ia(head))
    dll.czb_ctx_set_res_normal(ctx, 1.0 / float((n - 2) ** 3))
    if nranks > 1:
        uid_path = os.path.join(work, "uid")
        if rank == 0:
            buf = C.create_string_buffer(128)
            dll.czb_comm_unique_id(buf)
            with open(uid_path + ".tmp", "wb") as f:
                f.write(buf.raw)
            os.rename(uid_path + ".tmp", uid_path)
        t0 = time.time()
        while not os.path.exists(uid_path):
            time.sleep(0.01)
            assert time.time() - t0 < 60, "no unique id"
        buf = C.create_string_buffer(open(uid_path, "rb").read(), 128)
        dll.czb_ctx_comm_init(ctx, nranks, rank, buf)
    shape = decomp.field_shape(sz)
    P, RHS = np.zeros(shape, dt), np.zeros(shape, dt)
    dll.czb_ctx_bc_k(ctx, C.c_void_p(P.ctypes.data)); dll.czb_ctx_bc_k(ctx, C.c_void_p(RHS.ctypes.data))
    hist = (C.c_double * (iters + 2))()
    res = C.c_double(0); flop = C.c_double(0)
    itr = dll.czb_solve(ctx, solver.encode(), pre.encode() if pre else None, C.c_void_p(P.ctypes.data), C.c_void_p(RHS.ctypes.data),
                        iters, C.c_float(omega), C.byref(res), hist, iters + 2, C.byref(flop))
    nh = min(itr, iters - 1 if solver.startswith("pbicgstab") else iters)
    own = P[idx[2] + 1:idx[3] + 2].copy()          # owned inner planes (local C indices)
    extra = {}
    if gdv:   # owned inner box and where it sits in the global (j, i, k) array of inner + boundary + guard cells
        own = P[idx[2] + 1:idx[3] + 2, idx[0] + 1:idx[1] + 2, idx[4] + 1:idx[5] + 2].copy()
        extra["at"] = np.array([head[1] + idx[2], head[0] + idx[0], head[2] + idx[4]])   # global C index of the first owned cell
    np.savez(os.path.join(work, "out_%d.npz" % rank), own=own, itr=itr, hist=np.array(hist[:max(nh, 0)]), res=res.value, **extra)
    dll.czb_ctx_destroy(ctx)


if __name__ == "__main__":
    main()
