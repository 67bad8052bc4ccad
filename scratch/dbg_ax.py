import os, sys, ctypes as C
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import cubez_b200, oracle_util as ou
n = int(sys.argv[1]) if len(sys.argv) > 1 else 16
cz = cubez_b200.CZ([n]*3, "f32"); o = ou.OracleCZ([n]*3, "f32")
X, B, W = ou.random_fields(o.shape, o.dt, n=3)
Xg, Wg = cz.from_numpy(X), cz.from_numpy(W)
cf = np.array([1,1,1,1,1,1,6], dtype=o.dt); flop = C.c_double(0)
Ao = np.zeros_like(X); o.dll.blas_calc_ax_(ou.ptr(Ao), ou.ptr(X), o._sz, o._idx, o.g(), ou.ptr(cf), C.byref(flop))
Ag = cz.from_numpy(np.zeros_like(X)); dw, ds = C.c_double(0), C.c_double(0)
ok = cz.dll.czb_apply_ax(cz.ctx, cz._ptr(Ag), cz._ptr(Xg), cz._ptr(Wg), C.byref(dw), C.byref(ds))
A = Ag.cpu().numpy()
d = np.argwhere(A != Ao)
print("ok", ok, "ndiff", len(d), "of", A.size, "shape", A.shape)
print("first diffs (j,i,k):", d[:12].tolist())
for idx in d[:6]:
    t = tuple(idx); print(t, A[t], Ao[t], A[t]-Ao[t])
if len(d):
    print("j range", d[:,0].min(), d[:,0].max(), "i", d[:,1].min(), d[:,1].max(), "k", d[:,2].min(), d[:,2].max())
    print("k hist", np.bincount(d[:,2]).tolist())
