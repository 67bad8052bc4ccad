import sys, torch
sys.path.insert(0, '/root/repo')
import cubez_b200
n=1024
cz = cubez_b200.CZ([n]*3, "f32", eps=0.0)
for mode in ("zero", "random", "const"):
    P, RHS = cz.setup_laplace()
    if mode == "random":
        P[2:-2,2:-2,2:-2] = torch.rand(P[2:-2,2:-2,2:-2].shape, device="cuda")*2-1; cz.bc_k(P)
    if mode == "const":
        P[3:-3,3:-3,3:-3] = 0.5
    for K in (20, 100):
        out = cz.solve("sor2sma", P, RHS, K, 1.5, want_history=False)
        print(mode, K, "ms/iter %.3f" % (out["ms"]/K), "GLUPS %.1f" % (cz.inner_cells_global*K/out["ms"]/1e6))
    del P, RHS
