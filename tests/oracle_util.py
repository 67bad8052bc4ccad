"""TEST INFRASTRUCTURE: ctypes access to the CPU oracle (oracle/liboracle_*.so, a C++
restatement of the reference's Fortran kernels) and Python restatements of the reference's
iteration loops (src/cz_cpp/cz_Poisson.cpp) over it.  Never imported by the product."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
GUIDE = 2
_LIBS = {}


def build_oracle():
    need = [os.path.join(ORACLE_DIR, "liboracle_f32.so"), os.path.join(ORACLE_DIR, "liboracle_f64.so")]
    if not all(os.path.exists(p) for p in need):
        subprocess.check_call(["make", "-s", "-C", ORACLE_DIR, "libs"])


def load(real="f32"):
    if real in _LIBS:
        return _LIBS[real]
    build_oracle()
    import cubez_b200.ffi as ffi
    dll = C.CDLL(os.path.join(ORACLE_DIR, "liboracle_%s.so" % real))
    ffi.bind_reference_symbols(dll)
    for n in ("tdma_p_", "tdma_0_", "oracle_pcr_line_"):
        getattr(dll, n).restype = None
    _LIBS[real] = dll
    return dll


def ref_binary(name):
    p = os.path.join(ORACLE_DIR, "_ref", name)
    return p if os.path.exists(p) else None


def iarr(v):
    return (C.c_int * len(v))(*[int(x) for x in v])


def ptr(a):
    return C.c_void_p(a.ctypes.data)


def get_num_stage(n):  # cz.h:293-300
    b = 1
    for i in range(1, 20):
        b *= 2
        if n < b:
            return i
    return -1


class OracleCZ:
    """The reference's CZ state + loops (serial, numProc=1) over the oracle kernels."""

    def __init__(self, gsz, real="f32", sz=None, idx=None, nID=None, head=None, eps=1.0e-5):
        self.real = real
        self.dll = load(real)
        self.dt = np.float32 if real == "f32" else np.float64
        self.creal = C.c_float if real == "f32" else C.c_double
        self.gsz = list(gsz)
        self.sz = list(sz) if sz else list(gsz)
        self.nID = list(nID) if nID else [-1] * 6
        self.head = list(head) if head else [1, 1, 1]
        self.idx = list(idx) if idx else [2, self.sz[0] - 1, 2, self.sz[1] - 1, 2, self.sz[2] - 1]
        self.shape = (self.sz[1] + 4, self.sz[0] + 4, self.sz[2] + 4)
        self.pitch = self.dt(1.0) / self.dt(self.gsz[2] - 1)
        self.origin = [self.dt(0) + self.dt(self.head[a] - 1) * self.pitch for a in range(3)]
        self._sz, self._idx, self._nID = iarr(self.sz), iarr(self.idx), iarr(self.nID)
        self._g = C.c_int(GUIDE)
        self._dh = (self.creal * 3)(self.pitch, self.pitch, self.pitch)
        self._org = (self.creal * 3)(*self.origin)
        self.cf = np.array([1, 1, 1, 1, 1, 1, 6], dtype=self.dt)
        self.eps = eps
        n_global = (self.gsz[0] - 2) * (self.gsz[1] - 2) * (self.gsz[2] - 2)
        self.res_normal = 1.0 / float(n_global)
        self.ac1 = self.dt(0)
        self.WRK = None; self.MSK = None
        self.work1d = [np.zeros(self.sz[2] + 4, dtype=self.dt) for _ in range(6)]

    def zeros(self):
        return np.zeros(self.shape, dtype=self.dt)

    def g(self):
        return C.byref(self._g)

    def bc_k(self, p):
        self.dll.bc_k_(self._sz, self.g(), ptr(p), self._dh, self._org, self._nID)

    def setup_laplace(self):
        P, RHS = self.zeros(), self.zeros()
        self.bc_k(P); self.bc_k(RHS)
        return P, RHS

    def mask(self):
        if self.MSK is None:
            self.MSK = self.zeros()
            self.dll.imask_k_(ptr(self.MSK), self._sz, self._idx, self.g())
        return self.MSK

    # ---- single sweeps ------------------------------------------------------------------
    def sweep_jacobi(self, X, B, res):
        if self.WRK is None:
            self.WRK = self.zeros()
        flop = C.c_double(0)
        om = self.creal(self.ac1)
        self.dll.jacobi_(ptr(X), self._sz, self._idx, self.g(), ptr(self.cf), C.byref(om), ptr(B), C.byref(res),
                         ptr(self.WRK), C.byref(flop))

    def sweep_rb(self, X, B, res, ip=0):
        flop = C.c_double(0); om = self.creal(self.ac1); cip = C.c_int(ip)
        for color in (0, 1):
            cc = C.c_int(color)
            self.dll.psor2sma_core_(ptr(X), self._sz, self._idx, self.g(), ptr(self.cf), C.byref(cip), C.byref(cc),
                                    C.byref(om), ptr(B), C.byref(res), C.byref(flop))

    def sweep_psor(self, X, B, res):
        flop = C.c_double(0); om = self.creal(self.ac1)
        self.dll.psor_(ptr(X), self._sz, self._idx, self.g(), ptr(self.cf), C.byref(om), ptr(B), C.byref(res),
                       C.byref(flop))

    def sweep_pcr(self, kind, X, B, res):
        n = self.idx[5] - self.idx[4] + 1
        pn = C.c_int(get_num_stage(n)); ip = C.c_int(0); s = C.c_int(2 ** (pn.value - 2))
        flop = C.c_double(0); om = self.creal(self.ac1)
        M = self.mask(); w = [ptr(a) for a in self.work1d]
        d = self.dll
        if kind in ("pcr_rb", "pcr_rb_esa"):
            for color in (0, 1):
                cc = C.c_int(color)
                if kind == "pcr_rb":
                    d.pcr_rb_(self._sz, self._idx, self.g(), C.byref(pn), C.byref(ip), C.byref(cc), ptr(X), ptr(M),
                              ptr(B), *w, C.byref(om), C.byref(res), C.byref(flop))
                else:
                    d.pcr_rb_esa_(self._sz, self._idx, self.g(), C.byref(pn), C.byref(ip), C.byref(cc), C.byref(s),
                                  ptr(X), ptr(M), ptr(B), *w, C.byref(om), C.byref(res), C.byref(flop))
        elif kind == "pcr":
            d.pcr_(self._sz, self._idx, self.g(), C.byref(pn), ptr(X), ptr(M), ptr(B), *w, C.byref(om), C.byref(res),
                   C.byref(flop))
        elif kind == "pcr_eda":
            d.pcr_eda_(self._sz, self._idx, self.g(), C.byref(pn), ptr(X), ptr(M), ptr(B), *w[:3], C.byref(om),
                       C.byref(res), C.byref(flop))
        elif kind == "pcr_esa":
            d.pcr_esa_(self._sz, self._idx, self.g(), C.byref(pn), C.byref(s), ptr(X), ptr(M), ptr(B), *w, C.byref(om),
                       C.byref(res), C.byref(flop))
        elif kind == "pcr_j_esa":
            if self.WRK is None:
                self.WRK = self.zeros()
            src = self.zeros()
            d.pcr_j_esa_(self._sz, self._idx, self.g(), C.byref(pn), C.byref(s), ptr(X), ptr(M), ptr(B), *w, ptr(src),
                         ptr(self.WRK), C.byref(om), C.byref(res), C.byref(flop))
        else:
            raise ValueError(kind)

    # ---- stretched-grid (MAF) twins, cz_maf.f90 / cz_blas.f90:738-1039 --------------------------------
    def mesh(self, stretch=0.0):
        """xc, yc, zc as the reference driver sets them (cz_Evaluate.cpp:342-362): (REAL)(i-1)*pitch for the C index
        i = 0..size+3 (= Fortran X(-1:sz+2)); stretch > 0 adds a smooth monotone perturbation for kernel tests."""
        out = []
        for a in range(3):
            n = self.sz[a] + 2 * GUIDE
            i = np.arange(n)
            x = ((i - 1).astype(self.dt) * self.pitch).astype(self.dt)
            if stretch:
                x = (x + self.dt(stretch) * self.pitch * np.sin(self.dt(0.37 + a) * i.astype(self.dt))).astype(self.dt)
            out.append(np.ascontiguousarray(x))
        return out

    def pivot(self, mesh):
        pvt = self.zeros()
        self.dll.search_pivot_(ptr(pvt), self._sz, self._idx, self.g(), ptr(mesh[0]), ptr(mesh[1]), ptr(mesh[2]))
        return pvt

    def sweep_maf(self, kind, X, B, res, mesh, ip=0):
        """One iteration of the MAF twin `kind` (jacobi_maf, psor_maf, sor2sma_maf, pcr_rb_maf, pcr_rb_esa_maf,
        pcr_maf, pcr_eda_maf, pcr_esa_maf) with the argument lists of cz_Poisson.cpp:50,112,197,565,689,778,861,966.
        Returns the flop count the kernels added."""
        d = self.dll
        flop = C.c_double(0); om = self.creal(self.ac1); cip = C.c_int(ip)
        xc, yc, zc = [ptr(m) for m in mesh]
        tmp = np.zeros(self.sz[2] + 2 * GUIDE, dtype=self.dt)
        if kind == "jacobi_maf":
            if self.WRK is None:
                self.WRK = self.zeros()
            d.jacobi_maf_(ptr(X), self._sz, self._idx, self.g(), xc, yc, zc, C.byref(om), ptr(B), C.byref(res),
                          ptr(self.WRK), ptr(tmp), C.byref(flop))
            return flop.value
        if kind == "psor_maf":
            d.psor_maf_(ptr(X), self._sz, self._idx, self.g(), xc, yc, zc, C.byref(om), ptr(B), C.byref(res),
                        C.byref(flop))
            return flop.value
        if kind == "sor2sma_maf":
            for color in (0, 1):
                cc = C.c_int(color)
                d.psor2sma_core_maf_(ptr(X), self._sz, self._idx, self.g(), xc, yc, zc, C.byref(cip), C.byref(cc),
                                     C.byref(om), ptr(B), C.byref(res), ptr(tmp), C.byref(flop))
            return flop.value
        n = self.idx[5] - self.idx[4] + 1
        pn = C.c_int(get_num_stage(n)); s = C.c_int(2 ** (pn.value - 2))
        M = self.mask(); w = [ptr(a) for a in self.work1d]
        head = (self._sz, self._idx, self.g(), C.byref(pn))
        tail = (C.byref(om), C.byref(res), ptr(tmp), C.byref(flop))
        if kind == "pcr_rb_maf":
            for color in (0, 1):
                cc = C.c_int(color)
                d.pcr_rb_maf_(*head, C.byref(cip), C.byref(cc), ptr(X), ptr(M), ptr(B), xc, yc, zc, *w, *tail)
        elif kind == "pcr_rb_esa_maf":
            for color in (0, 1):
                cc = C.c_int(color)
                d.pcr_rb_esa_maf_(*head, C.byref(cip), C.byref(cc), C.byref(s), ptr(X), ptr(M), ptr(B), xc, yc, zc, *w,
                                  *tail)
        elif kind == "pcr_maf":
            d.pcr_maf_(*head, ptr(X), ptr(M), ptr(B), xc, yc, zc, *w, *tail)
        elif kind == "pcr_eda_maf":
            d.pcr_eda_maf_(*head, ptr(X), ptr(M), ptr(B), xc, yc, zc, *w[:3], *tail)
        elif kind == "pcr_esa_maf":
            d.pcr_esa_maf_(*head, C.byref(s), ptr(X), ptr(M), ptr(B), xc, yc, zc, *w, *tail)
        else:
            raise ValueError(kind)
        return flop.value

    # ---- loops (cz_Poisson.cpp) ------------------------------------------------------------
    def stationary(self, kind, X, B, itr_max, coef, converge_check=True, mesh=None):
        """CZ::JACOBI/PSOR/RBSOR/LSOR_PCR_*: returns (itr, res, history).  `*_maf` kinds use `mesh` (default: the
        driver's uniform mesh)."""
        self.ac1 = self.dt(coef)
        if kind.endswith("_maf") and mesh is None:
            mesh = self.mesh()
        hist = []
        res_v = 0.0
        itr = 1
        while itr <= itr_max:
            res = C.c_double(0.0)
            if kind.endswith("_maf"):
                self.sweep_maf(kind, X, B, res, mesh)
            elif kind == "jacobi":
                self.sweep_jacobi(X, B, res)
            elif kind == "sor2sma":
                self.sweep_rb(X, B, res)
            elif kind == "psor":
                self.sweep_psor(X, B, res)
            else:
                self.sweep_pcr(kind, X, B, res)
            if converge_check:
                res_v = float(np.sqrt(res.value * self.res_normal))
                hist.append(res_v)
                self.bc_k(X)
                if res_v < self.eps:
                    break
            itr += 1
        return itr, res_v, np.array(hist)

    def dot1(self, x):
        r = self.creal(0); flop = C.c_double(0)
        self.dll.blas_dot1_(C.byref(r), ptr(x), self._sz, self._idx, self.g(), C.byref(flop))
        return self.dt(r.value)

    def dot2(self, x, y):
        r = self.creal(0); flop = C.c_double(0)
        self.dll.blas_dot2_(C.byref(r), ptr(x), ptr(y), self._sz, self._idx, self.g(), C.byref(flop))
        return self.dt(r.value)

    def precond(self, xx, bb, pc, mesh=None):
        if pc in ("jacobi", "psor", "sor2sma", "pcr", "pcr_eda", "pcr_rb", "pcr_rb_esa",
                  "jacobi_maf", "psor_maf", "sor2sma_maf", "pcr_maf", "pcr_eda_maf", "pcr_rb_maf", "pcr_rb_esa_maf"):
            self.stationary(pc, xx, bb, 8, self.ac1, converge_check=False, mesh=mesh)   # cz_Poisson.cpp:282-317
        else:
            xx[...] = bb

    def pbicgstab(self, X, B, itr_max, coef, pc, maf=False, mesh=None):
        """CZ::PBiCGSTAB (cz_Poisson.cpp:332-504): returns (itr, res, history).  maf: pbicgstab_maf, i.e.
        calc_rk_maf / calc_ax_maf scaled by pvt (:350-358, 415-419, 451-455)."""
        self.ac1 = self.dt(coef)
        d = self.dll; dt = self.dt
        flop = C.c_double(0)
        p, p_, r, r0, q, s, s_, t_ = [self.zeros() for _ in range(8)]
        if maf:
            mesh = mesh if mesh is not None else self.mesh()
            pvt = self.pivot(mesh)
            mp = [ptr(m) for m in mesh]

            def apply_ax(out, v):
                d.calc_ax_maf_(ptr(out), ptr(v), self._sz, self._idx, self.g(), *mp, ptr(pvt), C.byref(flop))
            d.calc_rk_maf_(ptr(r), ptr(X), ptr(B), self._sz, self._idx, self.g(), *mp, ptr(pvt), C.byref(flop))
        else:
            def apply_ax(out, v):
                d.blas_calc_ax_(ptr(out), ptr(v), self._sz, self._idx, self.g(), ptr(self.cf), C.byref(flop))
            d.blas_calc_rk_(ptr(r), ptr(X), ptr(B), self._sz, self._idx, self.g(), ptr(self.cf), C.byref(flop))
        r0[...] = r
        rho_old, alpha, omega = dt(1.0), dt(0.0), dt(1.0)
        hist = []; res_v = 0.0
        flt_min = np.finfo(np.float32).tiny
        itr = 1
        while itr < itr_max:
            rho = self.dot2(r, r0)
            if abs(rho) < flt_min:
                return 0, res_v, np.array(hist)
            if itr == 1:
                p[...] = r
            else:
                beta = dt(dt(dt(rho / rho_old) * alpha) / omega)
                cb, co = self.creal(beta), self.creal(omega)
                d.blas_bicg_1_(ptr(p), ptr(r), ptr(q), C.byref(cb), C.byref(co), self._sz, self._idx, self.g(),
                               C.byref(flop))
            p_[...] = 0
            self.precond(p_, p, pc, mesh)
            apply_ax(q, p_)
            alpha = dt(rho / self.dot2(q, r0))
            ra = self.creal(-alpha)
            d.blas_triad_(ptr(s), ptr(q), ptr(r), C.byref(ra), self._sz, self._idx, self.g(), C.byref(flop))
            s_[...] = 0
            self.precond(s_, s, pc, mesh)
            apply_ax(t_, s_)
            omega = dt(self.dot2(t_, s) / self.dot1(t_))
            ca, co = self.creal(alpha), self.creal(omega)
            d.blas_bicg_2_(ptr(X), ptr(p_), ptr(s_), C.byref(ca), C.byref(co), self._sz, self._idx, self.g(),
                           C.byref(flop))
            ro = self.creal(-omega)
            d.blas_triad_(ptr(r), ptr(t_), ptr(s), C.byref(ro), self._sz, self._idx, self.g(), C.byref(flop))
            res_v = float(np.sqrt(float(self.dot1(r)) * self.res_normal))
            hist.append(res_v)
            self.bc_k(X)
            if res_v < self.eps:
                break
            rho_old = rho
            itr += 1
        return itr, res_v, np.array(hist)


def random_fields(shape, dt, n=2, seed=20240607):
    """Kernel-level parity input: uniform(-1,1) on all entries."""
    rng = np.random.default_rng(seed)
    return [rng.uniform(-1.0, 1.0, size=shape).astype(dt) for _ in range(n)]
