import sys, torch
sys.path.insert(0, '/root/repo')
import cubez_b200
n=1024
cz = cubez_b200.CZ([n]*3, "f32", eps=0.0)
for mode, val in (("1e-30", 1e-30), ("1e-37", 1e-37), ("3e-38(e=2)", 3e-38), ("1e-39", 1e-39), ("1e-44", 1e-44), ("randn*1e-38", None)):
    P = cz.zeros(); RHS = cz.zeros()
    if val is None:
        P[3:-3,3:-3,3:-3] = torch.rand(P[3:-3,3:-3,3:-3].shape, device="cuda")*2e-38
    else:
        P[3:-3,3:-3,3:-3] = val
    for K in (10,):
        out = cz.solve("sor2sma", P, RHS, K, 1.0, want_history=False)
        print(mode, K, "ms/iter %.3f" % (out["ms"]/K))
    out = cz.solve("jacobi", P, RHS, 10, 1.0, want_history=False)
    print("   jacobi", "ms/iter %.3f" % (out["ms"]/10))
    del P, RHS
