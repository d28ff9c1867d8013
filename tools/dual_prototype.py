"""numpy prototype of the dual-form nested dissection (the formulation of kernels_dual.cuh), checked against the oracle.
A subtree [a,b] is represented by its Schur complement onto the multipliers (lambda_a, lambda_{b+1}):
  element matrix [[Saa, Sab], [Sab^T, Sbb]], element rhs [sa; sb].
Merging [a,i] and [i+1,b] over the separator lambda_{i+1}:
  S = SbbL + SaaR,  S_L = SabL^T,  S_R = SabR,  s = sbL + saR,  [fL fR zs] = S^-1 [S_L S_R s]
  Saa' = SaaL - S_L^T fL, Sab' = -S_L^T fR, Sbb' = SbbR - S_R^T fR, sa' = saL - S_L^T zs, sb' = sbR - S_R^T zs.
Same separators, same S, same records [fL | fR | zs] as the reference's loop (src/solve.c:68-134)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def dual_solve(n, m, N, knots, x0):
    oA = 2 * n + 2 * m + 1
    Q, R = knots[:, :n], knots[:, n:n + m]
    q, r = knots[:, n + m:2 * n + m], knots[:, 2 * n + m:2 * n + 2 * m]
    A = knots[:, oA:oA + n * n].reshape(N, n, n).transpose(0, 2, 1)
    B = knots[:, oA + n * n:oA + n * n + n * m].reshape(N, m, n).transpose(0, 2, 1)
    d = knots[:, oA + n * n + n * m:]
    ent = {}
    for k in range(N):
        zx = -q[k] / Q[k] if k > 0 else x0.copy()
        zu = -r[k] / R[k] if k < N - 1 else np.zeros(m)
        zl = -(d[k - 1]) if k > 0 else None
        if k == 0:
            Saa = np.zeros((n, n)); Sab = np.zeros((n, n)); sa = np.zeros(n)
            Sbb = B[k] @ np.diag(1 / R[k]) @ B[k].T
        else:
            Saa = np.diag(1 / Q[k]); sa = -zx - zl
            if k < N - 1:
                Sab = -np.diag(1 / Q[k]) @ A[k].T
                Sbb = A[k] @ np.diag(1 / Q[k]) @ A[k].T + B[k] @ np.diag(1 / R[k]) @ B[k].T
            else:
                Sab = np.zeros((n, n)); Sbb = np.zeros((n, n))
        sb = A[k] @ zx + B[k] @ zu if k < N - 1 else np.zeros(n)
        ent[(k, k)] = (Saa, Sab, Sbb, sa, sb)
    rec = {}
    size = 1
    while size < N:
        for a in range(0, N, 2 * size):
            i, b = a + size - 1, a + 2 * size - 1
            SaaL, SabL, SbbL, saL, sbL = ent.pop((a, i))
            SaaR, SabR, SbbR, saR, sbR = ent.pop((i + 1, b))
            S, SL, SR, s = SbbL + SaaR, SabL.T, SabR, sbL + saR
            F = np.linalg.solve(S, np.concatenate([SL, SR, s[:, None]], axis=1))
            fL, fR, zs = F[:, :n], F[:, n:2 * n], F[:, 2 * n]
            rec[i + 1] = (fL, fR, zs, a, b + 1)
            ent[(a, b)] = (SaaL - SL.T @ fL, -SL.T @ fR, SbbR - SR.T @ fR, saL - SL.T @ zs, sbR - SR.T @ zs)
        size *= 2
    lam = np.zeros((N + 1, n))
    for sep in sorted(rec, key=lambda t: -(rec[t][4] - rec[t][3])):   # coarsest separators first
        fL, fR, zs, a, b1 = rec[sep]
        lam[sep] = zs - fL @ lam[a] - fR @ lam[b1]
    z = np.zeros((N, 2 * n + m))
    for k in range(N):
        ln = lam[k + 1] if k < N - 1 else np.zeros(n)
        if k == 0:
            z[k, n:2 * n] = x0
            z[k, :n] = Q[0] * x0 + q[0] + A[0].T @ ln
        else:
            z[k, :n] = lam[k]
            z[k, n:2 * n] = (-q[k] + lam[k] - (A[k].T @ ln if k < N - 1 else 0)) / Q[k]
        if k < N - 1:
            z[k, 2 * n:] = (-r[k] - B[k].T @ ln) / R[k]
    return z.reshape(-1)


if __name__ == "__main__":
    import oracle_lib as O
    import rslqr_b200 as R
    for (n, m, N) in [(12, 4, 2), (12, 4, 8), (12, 4, 64), (6, 3, 32), (5, 2, 16)]:
        kn, x0 = R.synth_problems(n, m, N, 7, 0, 1)
        want, _ = O.port_solve(n, m, N, kn[0], x0[0])
        z = dual_solve(n, m, N, kn[0], x0[0])
        print((n, m, N), "rel err vs oracle: %.2e" % (np.linalg.norm(z - want) / np.linalg.norm(want)))
