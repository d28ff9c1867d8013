"""Shared helpers for the test-suite."""
import json
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def golden_json(name):
    with open(os.path.join(GOLDEN, name)) as f:
        return json.load(f)


def jmat(obj):
    """JSON array-of-columns -> numpy (rows, cols)."""
    return np.array(obj, dtype=np.float64).T


def problem_from_json(name):
    """Pure-python reading of a reference problem file -> (n, m, N, knots[N,stride], x0)."""
    d = golden_json(name)
    N = d["nhorizon"]
    n, m = d["lqrdata"][0]["nstates"], d["lqrdata"][0]["ninputs"]
    st = 2 * n + 2 * m + 1 + n * n + n * m + n
    knots = np.zeros((N, st))
    for kd in d["lqrdata"]:
        k = kd["index"] - 1
        parts = [kd["Q"], kd["R"], kd["q"], kd["r"], [kd["c"]],
                 np.array(kd["A"]).reshape(-1), np.array(kd["B"]).reshape(-1), kd["d"]]
        knots[k] = np.concatenate([np.asarray(p, dtype=np.float64).reshape(-1) for p in parts])
    return n, m, N, knots, np.array(d["x0"], dtype=np.float64), d


def split_xul(z, n, m, N):
    """solution vector -> (lambda[N,n], x[N,n], u[N,m])"""
    z = np.asarray(z).reshape(N, 2 * n + m)
    return z[:, :n], z[:, n:2 * n], z[:, 2 * n:]


def rel_err_classes(z, zref, n, m, N):
    """relative 2-norm error per variable class (lambda, x, u) as the north star states it"""
    out = []
    for a, b in zip(split_xul(z, n, m, N), split_xul(zref, n, m, N)):
        out.append(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))
    return out


def blocks(F, n, m):
    """factor block (fs,) -> (Lambda[n,n], X[n,n], U[m,n]) column-major views"""
    nn = n * n
    return (F[:nn].reshape(n, n, order="F"), F[nn:2 * nn].reshape(n, n, order="F"),
            F[2 * nn:].reshape(m, n, order="F"))
