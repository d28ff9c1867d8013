"""Device time of ONE solve (CUDA events around solve_async) for a few shapes; run on the GPU box."""
import sys, time
sys.path.insert(0, "/root/repo")
import numpy as np
import rslqr_b200 as R
for (n, m, N) in [(6, 3, 8), (6, 3, 256), (12, 4, 256), (12, 4, 16384)]:
    kn, x0 = R.synth_problems(n, m, N, 1, 0, 1)
    s = R.CudaSolver(n, m, N, batch=1, flags=0)
    s.upload(kn, x0)
    for _ in range(5):
        s.solve()
    s.mark(0)
    for _ in range(20):
        s.solve_async(True)
    s.mark(1)
    dev = s.elapsed_ms(0, 1) / 20
    t0 = time.perf_counter()
    for _ in range(20):
        s.solve()
    host = (time.perf_counter() - t0) / 20 * 1e3
    print(f"n={n} m={m} N={N}: mode {s.mode}, {s.launches} launches, device {dev*1e3:.1f} us/solve back to back, "
          f"host solve()+sync {host*1e3:.1f} us")
    s.close()
