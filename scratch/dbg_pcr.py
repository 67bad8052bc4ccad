import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, cubez_b200, oracle_util as ou
sz = [int(a) for a in sys.argv[1:4]]; solver = sys.argv[4]
cz = cubez_b200.CZ(sz, "f32", eps=0.0); o = ou.OracleCZ(sz, "f32", eps=0.0)
X, B = ou.random_fields(o.shape, o.dt); o.bc_k(X)
Xg, Bg = cz.from_numpy(X), cz.from_numpy(B)
out = cz.solve(solver, Xg, Bg, 3, 1.5)
itr, res, hist = o.stationary(solver, X, B, 3, 1.5)
got = Xg.cpu().numpy()
print(sz, solver, "equal" if np.array_equal(got, X) else "DIFF max %g" % np.abs(got - X).max(), out["history"], hist)
