"""Host-side mirror of the reference's `class CZ` (src/cz_cpp/cz.h, cz_Evaluate.cpp) for the
iterative-solver hot path, over the C ABI of libczb200.  torch is used only for device memory,
streams and (multi-GPU) rendezvous; every kernel is in the CUDA library."""
import ctypes as C

import numpy as np

from . import decomp, lib

GUIDE = decomp.GUIDE
EPS_DEFAULT = 1.0e-5  # reference src/cz_cpp/cz.h:162

SOLVERS = ("jacobi", "psor", "sor2sma", "pbicgstab", "pcr", "pcr_eda", "pcr_esa", "pcr_rb", "pcr_rb_esa",
           "pcr_j_esa",
           # stretched-grid twins, CZ::setLS cz_Evaluate.cpp:741-797
           "jacobi_maf", "psor_maf", "sor2sma_maf", "pbicgstab_maf", "pcr_maf", "pcr_eda_maf", "pcr_esa_maf",
           "pcr_rb_maf", "pcr_rb_esa_maf")


def _iarr(v):
    return (C.c_int * len(v))(*[int(x) for x in v])


class CZ:
    """One (sub)domain of the CubeZ problem on one GPU.

    gsz: global nodes per axis (the reference's gsz_x gsz_y gsz_z, boundary nodes included);
    real: 'f32' | 'f64' (the reference's real_type build option);
    nranks/rank: j-slab decomposition (one process per GPU); gdv = (gdv_x, gdv_y, gdv_z): general 3-D bricks
    instead (the reference's trailing command-line arguments, main.cpp:19-29), nranks = their product.
    """

    def __init__(self, gsz, real="f32", device=None, nranks=1, rank=0, eps=EPS_DEFAULT, gdv=None):
        import torch
        self.torch = torch
        if not torch.cuda.is_available():
            raise RuntimeError("cubez_b200 needs a CUDA device (no CPU fallback)")
        self.real = real
        self.dll = lib.load(real)
        self.device = torch.device("cuda", torch.cuda.current_device() if device is None else device)
        torch.cuda.set_device(self.device)
        self.dll.czb_init(self.device.index)
        self.dll.czb_set_stream(C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream))
        self.tdtype = torch.float32 if real == "f32" else torch.float64
        self.ndtype = np.float32 if real == "f32" else np.float64
        self.creal = C.c_float if real == "f32" else C.c_double
        self.gsz = [int(x) for x in gsz]
        self.nranks, self.rank = nranks, rank
        self.gdv = [int(v) for v in gdv] if gdv is not None else None
        if self.gdv is not None and self.gdv[0] * self.gdv[1] * self.gdv[2] != nranks:
            raise ValueError("The number of proceees does not agree with the division size.")   # cz_Evaluate.cpp:79-83
        d = decomp.brick_decompose(self.gsz, self.gdv, rank) if self.gdv is not None else decomp.slab_decompose(self.gsz, nranks, rank)
        self.sz, self.head, self.nID, self.idx = d["sz"], d["head"], d["nID"], d["idx"]
        # cz_Evaluate.cpp:88  pitch = 1/(G_size[2]-1) on all axes, computed in REAL
        self.pitch = self.ndtype(1.0) / self.ndtype(self.gsz[2] - 1)
        # cz_Evaluate.cpp:136-138
        self.origin = [self.ndtype(0) + self.ndtype(self.head[a] - 1) * self.pitch for a in range(3)]
        self.shape = decomp.field_shape(self.sz)
        self._sz = _iarr(self.sz); self._idx = _iarr(self.idx); self._nID = _iarr(self.nID)
        self._g = C.c_int(GUIDE)
        self._dh = (self.creal * 3)(self.pitch, self.pitch, self.pitch)
        self._org = (self.creal * 3)(*self.origin)
        self.ctx = self.dll.czb_ctx_create(self._sz, self._idx, self._nID, self.creal(self.pitch), self._org)
        self.dll.czb_ctx_set_eps(self.ctx, eps)
        self.dll.czb_ctx_set_head(self.ctx, _iarr(self.head))
        n_global = (self.gsz[0] - 2) * (self.gsz[1] - 2) * (self.gsz[2] - 2)
        self.res_normal = 1.0 / float(n_global)          # cz_Evaluate.cpp:222-224 (after Comm_SUM_1)
        self.dll.czb_ctx_set_res_normal(self.ctx, self.res_normal)
        self.inner_cells_local = decomp.inner_cells(self.idx)
        self.inner_cells_global = n_global

    # ---- fields -------------------------------------------------------------------------
    def zeros(self):
        return self.torch.zeros(self.shape, dtype=self.tdtype, device=self.device)

    def from_numpy(self, a):
        assert a.shape == self.shape
        return self.torch.from_numpy(np.ascontiguousarray(a, dtype=self.ndtype)).to(self.device)

    def _ptr(self, t):
        if isinstance(t, np.ndarray):
            assert t.dtype == self.ndtype and t.flags["C_CONTIGUOUS"]
            return C.c_void_p(t.ctypes.data)
        assert t.dtype == self.tdtype and t.is_contiguous()
        return C.c_void_p(t.data_ptr())

    def sync_stream(self):
        self.dll.czb_set_stream(C.c_void_p(self.torch.cuda.current_stream(self.device).cuda_stream))

    # ---- problem setup (cz_Evaluate.cpp:375-390) -----------------------------------------
    def bc_k(self, p):
        self.dll.czb_ctx_bc_k(self.ctx, self._ptr(p))

    def setup_laplace(self):
        """P and RHS of the reference problem: zero field, Dirichlet faces from bc_k."""
        P, RHS = self.zeros(), self.zeros()
        self.bc_k(P); self.bc_k(RHS)
        return P, RHS

    def exact_error(self, P):
        """exact_t_ + err_t_ (cz_Evaluate.cpp:550-563): max |P - analytic| over inner cells."""
        E = self.zeros()
        self.dll.exact_t_(self._sz, C.byref(self._g), self._ptr(E), self._dh, self._org)
        d = C.c_double(0.0); loc = (C.c_int * 3)()
        self.dll.err_t_(self._sz, self._idx, C.byref(self._g), C.byref(d), self._ptr(P), self._ptr(E), loc)
        return d.value, tuple(loc)

    # ---- multi-rank ----------------------------------------------------------------------
    def comm_init(self, unique_id_bytes):
        buf = C.create_string_buffer(bytes(unique_id_bytes), 128)
        self.dll.czb_ctx_comm_init(self.ctx, self.nranks, self.rank, buf)

    def comm_unique_id(self):
        buf = C.create_string_buffer(128)
        self.dll.czb_comm_unique_id(buf)
        return buf.raw

    # ---- solve ---------------------------------------------------------------------------
    def solve(self, solver, X, B, itr_max, coef, precond=None, want_history=True):
        """Run CZ::JACOBI / RBSOR / PBiCGSTAB / LSOR_PCR_* (cz_Poisson.cpp) on the device.
        X, B: torch CUDA tensors (zero copy) or numpy arrays (staged H2D/D2H).
        Returns dict(itr, res, history, ms)."""
        if solver.lower() not in SOLVERS:
            raise ValueError("Invalid solver %r" % solver)
        self.sync_stream()
        n = max(int(itr_max), 1) + 2
        hist = (C.c_double * n)() if want_history else None
        res = C.c_double(0.0); flop = C.c_double(0.0)
        itr = self.dll.czb_solve(self.ctx, solver.encode(), precond.encode() if precond else None,
                                 self._ptr(X), self._ptr(B), int(itr_max), self.creal(coef),
                                 C.byref(res), hist, n if want_history else 0, C.byref(flop))
        nh = 0
        if want_history and itr > 0:
            nh = min(itr, itr_max - 1 if solver.lower() in ("pbicgstab", "pbicgstab_maf") else itr_max)
        return dict(itr=itr, res=res.value, history=np.array(hist[:nh]) if want_history else None,
                    ms=self.dll.czb_last_solve_ms(self.ctx), flop=flop.value)

    def set_mesh(self, xc=None, yc=None, zc=None):
        """Coordinate vectors of the `*_maf` solvers (numpy, sz[a]+4 entries = Fortran X(-1:sz+2)); None = the
        reference driver's uniform mesh (cz_Evaluate.cpp:342-362)."""
        keep = []
        for a, v in enumerate((xc, yc, zc)):
            if v is None:
                keep.append(None)
                continue
            v = np.ascontiguousarray(v, dtype=self.ndtype)
            assert v.shape == (self.sz[a] + 2 * GUIDE,)
            keep.append(v)
        self.dll.czb_ctx_set_mesh(self.ctx, *[C.c_void_p(v.ctypes.data) if v is not None else None for v in keep])

    def variant(self, v):
        self.dll.czb_ctx_set_variant(self.ctx, int(v))

    def launch_count(self):
        return int(self.dll.czb_launch_count())

    def close(self):
        if self.ctx:
            self.dll.czb_ctx_destroy(self.ctx)
            self.ctx = None
