"""czb_dot2_noalloc (include/czb200.h; cubez_b200/csrc/noalloc.cu): the allocation-free variant of blas_dot2_ whose
reduction scratch is a file-scope static __device__ array.  No counterpart in the reference.
Called through the C ABI and checked against the existing entry point blas_dot2_ (and, on the
small grid, against the CPU oracle) within the tolerances tests/test_gpu_kernels.py uses for the dot products, on two
calls in a row (the second call finds the first call's contents in the shared scratch).  This file sorts last on purpose:
a failure here cannot hide another test."""
import ctypes as C

import numpy as np
import pytest

import oracle_util as ou



def _dot2_existing(cz, Xg, Yg):
    r, flop = cz.creal(0), C.c_double(0)
    cz.dll.blas_dot2_(C.byref(r), cz._ptr(Xg), cz._ptr(Yg), cz._sz, cz._idx, C.byref(cz._g), C.byref(flop))
    return float(r.value)


def _dot2_noalloc(cz, Xg, Yg, out, stream=None):
    rc = cz.dll.czb_dot2_noalloc(cz._ptr(Xg), cz._ptr(Yg), cz._sz, cz._idx, int(cz._g.value), C.c_void_p(out.data_ptr()),
                                 C.c_void_p(stream))
    return rc


@pytest.mark.gpu
@pytest.mark.parametrize("real", ["f32", "f64"])
@pytest.mark.parametrize("sz", [[16, 16, 16], [37, 21, 50], [130, 96, 258]])
def test_dot2_noalloc_equals_blas_dot2_on_two_calls_in_a_row(real, sz):
    import torch
    import cubez_b200
    cz = cubez_b200.CZ(sz, real)
    o = ou.OracleCZ(sz, real)
    X, Y, Z = ou.random_fields(o.shape, o.dt, n=3)
    Xg, Yg, Zg = cz.from_numpy(X), cz.from_numpy(Y), cz.from_numpy(Z)
    out = torch.full((1,), float("nan"), dtype=torch.float64, device=cz.device)     # caller-owned output
    s = torch.cuda.current_stream(cz.device).cuda_stream
    rel = 2e-6 if real == "f32" else 1e-11           # tests/test_gpu_kernels.py::test_blas_and_operator (dot2)
    ab = 1e-3 if real == "f32" else 1e-10
    # two calls in a row with DIFFERENT operands: the second must not see anything of the first in the shared scratch
    for A, B in ((Xg, Yg), (Yg, Zg), (Xg, Yg)):
        out.fill_(float("nan"))
        assert _dot2_noalloc(cz, A, B, out, s) == 0
        torch.cuda.synchronize()
        got = float(out.item())
        ref = _dot2_existing(cz, A, B)
        got_real = float(cz.ndtype(got))              # blas_dot2_ returns the double sum rounded to REAL
        assert got_real == pytest.approx(ref, rel=rel, abs=ab), (got, ref)
    if max(sz) <= 50:                                 # and the oracle's blas_dot2 itself
        ro, flop = cz.creal(0), C.c_double(0)
        o.dll.blas_dot2_(C.byref(ro), ou.ptr(X), ou.ptr(Y), o._sz, o._idx, o.g(), C.byref(flop))
        assert float(cz.ndtype(got)) == pytest.approx(float(ro.value), rel=rel, abs=ab)
        # the double sum before rounding: identical REAL products, only the order of the double additions differs
        exact = float(np.sum((X * Y).astype(np.float64)[o.idx[2] + 1:o.idx[3] + 2, o.idx[0] + 1:o.idx[1] + 2, o.idx[4] + 1:o.idx[5] + 2]))
        assert got == pytest.approx(exact, rel=1e-12, abs=1e-12)
    cz.close()


@pytest.mark.gpu
def test_dot2_noalloc_allocates_nothing_and_rejects_what_it_cannot_hold():
    import torch
    import cubez_b200
    cz = cubez_b200.CZ([64, 64, 64], "f32")
    o = ou.OracleCZ([64, 64, 64], "f32")
    X, Y = ou.random_fields(o.shape, o.dt, n=2)
    Xg, Yg = cz.from_numpy(X), cz.from_numpy(Y)
    out = torch.zeros(1, dtype=torch.float64, device=cz.device)
    assert _dot2_noalloc(cz, Xg, Yg, out) == 0        # first call on this device: symbol look-up, no allocation
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info(cz.device)[0]
    vals = []
    for _ in range(50):
        assert _dot2_noalloc(cz, Xg, Yg, out) == 0
        vals.append(out.clone())
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info(cz.device)[0] >= free0 - (1 << 20)     # nothing but the 50 tiny clones above
    assert all(float(v.item()) == float(vals[0].item()) for v in vals)    # deterministic: fixed summation order
    # host pointers are refused (staging would allocate); nothing is launched
    assert cz.dll.czb_dot2_noalloc(C.c_void_p(X.ctypes.data), cz._ptr(Yg), cz._sz, cz._idx, 2, C.c_void_p(out.data_ptr()), None) == -2
    # a (k,i) plane beyond CZB_NOALLOC_MAX_BLOCKS * 256 inner cells is rejected before any pointer is used for a launch
    big = (C.c_int * 3)(3000, 8, 3000)
    bidx = (C.c_int * 6)(2, 2999, 2, 7, 2, 2999)
    assert cz.dll.czb_dot2_noalloc(cz._ptr(Xg), cz._ptr(Yg), big, bidx, 2, C.c_void_p(out.data_ptr()), None) == -1
    # the library still works afterwards
    assert _dot2_noalloc(cz, Xg, Yg, out) == 0
    torch.cuda.synchronize()
    assert float(out.item()) == float(vals[0].item())
    cz.close()


@pytest.mark.parametrize("real", ["f32", "f64"])
def test_dot2_noalloc_host_logic_on_the_emulated_library(real):
    """CPU: noalloc.cu compiled into the CPU-emulated library (tests/tools/emu; the static `__device__` array is a host
    array there): launch geometry, j chunking, ticket handling, argument checks and the size limit -- against the
    oracle's blas_dot2 on two calls in a row.  No GPU involved; the GPU tests above are the parity tests proper."""
    import os
    import sys
    sys.path.insert(0, os.path.join(ou.ROOT, "tests", "tools", "emu"))
    import emu_build
    from cubez_b200 import ffi
    dll = C.CDLL(emu_build.build_lib()[real])
    creal = C.c_float if real == "f32" else C.c_double
    ffi.bind_reference_symbols(dll); ffi.bind_czb_symbols(dll, creal)
    dll.czb_init(0)
    # the emulation runs one block per CUDA THREAD, so the grids here are kept below 8192 threads in total
    for sz in ([12, 9, 20], [20, 6, 40]):
        o = ou.OracleCZ(sz, real)
        X, Y = ou.random_fields(o.shape, o.dt, n=2)
        nb = X.nbytes
        dx, dy, dout = dll.czb_malloc(nb), dll.czb_malloc(nb), dll.czb_malloc(8)
        dll.czb_memcpy_h2d(dx, X.ctypes.data, nb); dll.czb_memcpy_h2d(dy, Y.ctypes.data, nb)
        ro, flop = creal(0), C.c_double(0)
        o.dll.blas_dot2_(C.byref(ro), ou.ptr(X), ou.ptr(Y), o._sz, o._idx, o.g(), C.byref(flop))
        for _ in range(2):
            got = C.c_double(float("nan"))
            dll.czb_memcpy_h2d(dout, C.addressof(got), 8)
            assert dll.czb_dot2_noalloc(dx, dy, o._sz, o._idx, 2, dout, None) == 0
            dll.czb_memcpy_d2h(C.addressof(got), dout, 8)
            assert float(o.dt(got.value)) == pytest.approx(float(ro.value), rel=2e-6 if real == "f32" else 1e-11,
                                                             abs=1e-3 if real == "f32" else 1e-10)
        assert dll.czb_dot2_noalloc(X.ctypes.data, dy, o._sz, o._idx, 2, dout, None) == -2        # host pointer refused
        big, bidx = (C.c_int * 3)(3000, 8, 3000), (C.c_int * 6)(2, 2999, 2, 7, 2, 2999)
        assert dll.czb_dot2_noalloc(dx, dy, big, bidx, 2, dout, None) == -1                       # beyond the static scratch
        for d in (dx, dy, dout):
            dll.czb_free(d)
