import sys, time, numpy as np
sys.path.insert(0, "/root/repo")
import rslqr_b200 as R
n, m, N, B = 64, 16, 1024, int(sys.argv[1]) if len(sys.argv) > 1 else 8
kn, x0 = R.synth_problems(n, m, N, 2024, 0, B)
s = R.CudaSolver(n, m, N, batch=B, flags=R.KEEP_FACT | R.PROFILE)
s.set_mode(R.MODE_STAGED)
s.upload(kn, x0)
for it in range(3):
    s.reset(); t0 = time.perf_counter(); s.solve(); dt = time.perf_counter() - t0
    print("staged n=64 N=1024 batch", B, "solve %.1f ms = %.2f ms/problem" % (dt * 1e3, dt * 1e3 / B), s.profile())
