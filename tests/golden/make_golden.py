"""Regenerates tests/golden/ (run in the build container, where /root/reference exists):

1. copies the reference's DATA fixtures (lqr_prob.json, lqr_prob_256.json, sample_problem.json,
   lqr_data.json - problem data and stored solutions, no source code);
2. runs the UNMODIFIED reference (oracle/_ref/libref_rslqr.so) on synthetic problems of the
   benchmark shapes and stores its outputs (solution, a few factor blocks, KKT residual) as .npz,
   so the GPU box can compare against the real reference without /root/reference.

    python tests/golden/make_golden.py
"""
import os
import shutil
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
REFDIR = "/root/reference"

for f in ["lqr_prob.json", "lqr_prob_256.json", "sample_problem.json", "lqr_data.json"]:
    shutil.copyfile(os.path.join(REFDIR, f), os.path.join(HERE, f))

import oracle_lib as O  # noqa: E402
import rslqr_b200 as R  # noqa: E402

assert O.REF is not None, "build oracle/_ref first (make -C oracle ref)"
cases = [("quad_n12_m4_N256", 12, 4, 256, [0, 1, 4095]), ("quad_n12_m4_N32", 12, 4, 32, [0, 7]),
         ("gen_n6_m3_N64", 6, 3, 64, [3]), ("gen_n16_m5_N16", 16, 5, 16, [2]), ("gen_n64_m16_N8", 64, 16, 8, [0])]
out = {}
for name, n, m, N, ids in cases:
    for pid in ids:
        kn, x0 = R.synth_problems(n, m, N, 2024, pid, 1)
        soln, fact, _ = O.ref_solve(n, m, N, kn[0], x0[0], threads=1)
        key = f"{name}_id{pid}"
        out[key + "_soln"] = soln
        out[key + "_kkt"] = np.array(O.kkt_residual(n, m, N, kn[0], x0[0], soln))
        # top-level column of the finished factorization (what sample_problem.json stores as F*2*)
        out[key + "_fact_top"] = fact[-1].copy()
        out[key + "_fact_l0_k1"] = fact[0, 1].copy()
np.savez_compressed(os.path.join(HERE, "ref_synth.npz"), **out)
print("wrote", len(out), "arrays")
