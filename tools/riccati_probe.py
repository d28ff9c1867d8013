"""Time the device Riccati comparison solver (rslqr_cuda_riccati_solve_async) beside the nested-dissection solve.
usage: riccati_probe.py [batch] [N]"""
import sys, numpy as np
import os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import rslqr_b200 as R
n, m, B = 12, 4, int(sys.argv[1]) if len(sys.argv) > 1 else 4096
N = int(sys.argv[2]) if len(sys.argv) > 2 else 256
kn, x0 = R.synth_problems(n, m, N, 2024, 0, B)
s = R.CudaSolver(n, m, N, batch=B, flags=0)
s.upload(kn, x0)
for _ in range(3):
    s.riccati()
s.mark(0)
for _ in range(5):
    R.LIB.rslqr_cuda_riccati_solve_async(s.h)
s.mark(1); s.sync()
ms = s.elapsed_ms(0, 1) / 5
zr = s.solution().copy()
s.solve()
zn = s.solution()
print("N", N, "batch", B, "riccati ms/step", ms, "solves/s", B / ms * 1e3, "rel diff vs ND", np.linalg.norm(zr - zn) / np.linalg.norm(zn))
