"""Per-phase device time of the staged path at n=64, m=16, N=1024 (one chunk of 16 problems)."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import rslqr_b200 as R
n, m, N, B = 64, 16, 1024, 16
kn, x0 = R.synth_problems(n, m, N, 2024, 0, B)
s = R.CudaSolver(n, m, N, batch=B, flags=R.KEEP_FACT | R.PROFILE)
s.set_mode(R.MODE_STAGED)
s.upload(kn, x0)
for _ in range(2):
    s.reset(); s.solve()
print("phase ms [total, leaves, products, cholesky, solves, schur, rhs]:", s.profile())
