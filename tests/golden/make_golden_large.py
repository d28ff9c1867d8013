"""Stored outputs of the UNMODIFIED reference at the full sizes of BASELINE configs[3] and configs[4]
(run in the build container, where /root/reference exists and oracle/_ref has been built):

    quad_n12_m4_N16384_id0   single long horizon, n=12, m=4, N=16384   (reference: ~10 s on 8 threads)
    gen_n64_m16_N1024_id0    larger state, n=64, m=16, N=1024           (reference: ~40 s on 8 threads)

For each: the reference's solution vector (ndlqr_Solve, src/solve.c:38) and the KKT residual of that solution,
so that the GPU box can compare at size without /root/reference.   python tests/golden/make_golden_large.py
"""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_lib as O  # noqa: E402
import rslqr_b200 as R  # noqa: E402

assert O.REF is not None, "build oracle/_ref first (make -C oracle ref)"
threads = max(1, len(os.sched_getaffinity(0)))
out = {}
for name, n, m, N, pid in [("quad_n12_m4_N16384", 12, 4, 16384, 0), ("gen_n64_m16_N1024", 64, 16, 1024, 0)]:
    kn, x0 = R.synth_problems(n, m, N, 2024, pid, 1)
    t0 = time.time()
    soln, _, ms = O.ref_solve(n, m, N, kn[0], x0[0], threads=threads, want_fact=False)
    key = f"{name}_id{pid}"
    out[key + "_soln"] = soln
    out[key + "_kkt"] = np.array(O.kkt_residual(n, m, N, kn[0], x0[0], soln))
    print(key, "reference solve %.1f s (solve_time_ms %.0f), KKT residual %.3e, |z| %.3e" %
          (time.time() - t0, ms, float(out[key + "_kkt"]), np.linalg.norm(soln)), flush=True)
np.savez_compressed(os.path.join(HERE, "ref_synth_large.npz"), **out)
print("wrote", len(out), "arrays")
