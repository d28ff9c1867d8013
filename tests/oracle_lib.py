"""Test-side loaders for the parity oracles (TEST INFRASTRUCTURE, never imported by the product):

* ``PORT``  - oracle/liboracle_port.so, this repo's plain-C restatement (oracle/ndlqr_port.c)
* ``REF``   - oracle/_ref/libref_rslqr.so, the unmodified reference sources compiled by oracle/Makefile
              (None when it has not been built; the GPU box receives the prebuilt file).
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
GOLDEN = os.path.join(ROOT, "tests", "golden")
c_dp = C.POINTER(C.c_double)


def _build():
    port = os.path.join(ORACLE_DIR, "liboracle_port.so")
    src = os.path.join(ORACLE_DIR, "ndlqr_port.c")
    if not os.path.exists(port) or os.path.getmtime(port) < os.path.getmtime(src):
        subprocess.run(["make", "-C", ORACLE_DIR, "port"], check=True, capture_output=True)
    ref = os.path.join(ORACLE_DIR, "_ref", "libref_rslqr.so")
    if not os.path.exists(ref) and os.path.isdir("/root/reference/src"):
        subprocess.run(["make", "-C", ORACLE_DIR, "ref"], check=True, capture_output=True)
    return port, ref


_PORT_PATH, _REF_PATH = _build()


class PortSolverStruct(C.Structure):
    _fields_ = [("n", C.c_int), ("m", C.c_int), ("N", C.c_int), ("K", C.c_int), ("fs", C.c_int), ("zs", C.c_int),
                ("data", c_dp), ("fact", c_dp), ("soln", c_dp), ("diag", c_dp), ("chol_status", C.POINTER(C.c_int))]


PORT = C.CDLL(_PORT_PATH)
_PS = C.POINTER(PortSolverStruct)
for name, res, args in [
    ("port_knot_stride", C.c_int, [C.c_int, C.c_int]),
    ("port_new", _PS, [C.c_int, C.c_int, C.c_int]),
    ("port_free", None, [_PS]),
    ("port_reset", None, [_PS]),
    ("port_init", C.c_int, [_PS, C.c_void_p, C.c_void_p]),
    ("port_index_level", C.c_int, [C.c_int]),
    ("port_index_from_leaf", C.c_int, [C.c_int, C.c_int]),
    ("port_index_at_level", C.c_int, [C.c_int, C.c_int]),
    ("port_should_calc_lambda", C.c_int, [C.c_int, C.c_int, C.c_int]),
    ("port_solve_leaf", None, [_PS, C.c_int]),
    ("port_leaves", None, [_PS]),
    ("port_inner_product", None, [_PS, c_dp, C.c_int, C.c_int, C.c_int, C.c_int]),
    ("port_level_products", None, [_PS, C.c_int]),
    ("port_level_cholesky", None, [_PS, C.c_int]),
    ("port_level_solves", None, [_PS, C.c_int]),
    ("port_level_schur", None, [_PS, C.c_int]),
    ("port_rhs_level", None, [_PS, C.c_int]),
    ("port_solve", C.c_int, [_PS, C.c_int]),
    ("port_kkt_residual", C.c_double, [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    ("port_riccati", C.c_int, [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
]:
    fn = getattr(PORT, name)
    fn.restype, fn.argtypes = res, args


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class PortSolver:
    """numpy-facing wrapper of the C port; arrays are views of the C buffers."""

    def __init__(self, n, m, N):
        self.p = PORT.port_new(n, m, N)
        assert self.p, "port_new failed"
        s = self.p.contents
        self.n, self.m, self.N, self.K, self.fs, self.zs = n, m, N, s.K, s.fs, s.zs
        self.data = np.ctypeslib.as_array(s.data, shape=(s.K, N, s.fs))
        self.fact = np.ctypeslib.as_array(s.fact, shape=(s.K, N, s.fs))
        self.soln = np.ctypeslib.as_array(s.soln, shape=(N, s.zs))
        self.diag = np.ctypeslib.as_array(s.diag, shape=(N, n * n + m * m))
        self.status = np.ctypeslib.as_array(s.chol_status, shape=(3 * N,))

    def init(self, knots, x0):
        self._k = np.ascontiguousarray(knots, dtype=np.float64)
        self._x = np.ascontiguousarray(x0, dtype=np.float64)
        PORT.port_reset(self.p)
        return PORT.port_init(self.p, _p(self._k), _p(self._x))

    def leaves(self):
        PORT.port_leaves(self.p)

    def products(self, lv):
        PORT.port_level_products(self.p, lv)

    def cholesky(self, lv):
        PORT.port_level_cholesky(self.p, lv)

    def solves(self, lv):
        PORT.port_level_solves(self.p, lv)

    def schur(self, lv):
        PORT.port_level_schur(self.p, lv)

    def rhs_level(self, lv):
        PORT.port_rhs_level(self.p, lv)

    def solve(self, threads=1):
        return PORT.port_solve(self.p, threads)

    def close(self):
        if self.p:
            PORT.port_free(self.p)
            self.p = None

    def __del__(self):
        self.close()


def port_solve(n, m, N, knots, x0, threads=1):
    """Full oracle solve; returns (soln[N*zs], fact[K,N,fs])."""
    s = PortSolver(n, m, N)
    s.init(knots, x0)
    s.solve(threads)
    out = s.soln.reshape(-1).copy(), s.fact.copy()
    s.close()
    return out


def kkt_residual(n, m, N, knots, x0, z):
    knots = np.ascontiguousarray(knots, dtype=np.float64)
    x0 = np.ascontiguousarray(x0, dtype=np.float64)
    z = np.ascontiguousarray(z, dtype=np.float64)
    return PORT.port_kkt_residual(n, m, N, _p(knots), _p(x0), _p(z))


def riccati(n, m, N, knots, x0):
    knots = np.ascontiguousarray(knots, dtype=np.float64)
    x0 = np.ascontiguousarray(x0, dtype=np.float64)
    z = np.zeros(N * (2 * n + m))
    PORT.port_riccati(n, m, N, _p(knots), _p(x0), _p(z))
    return z


# ---------------------------------------------------------------------------------------------
# the compiled reference (same public C API as the product, reference struct layout)
# ---------------------------------------------------------------------------------------------
REF = None
if os.path.exists(_REF_PATH):
    import sys
    sys.path.insert(0, ROOT)
    from rslqr_b200 import ctypes_api as _pl  # struct definitions + prototype table only (loads no library)
    REF = _pl.bind_host_api(C.CDLL(_REF_PATH), _pl.RefSolverStruct)
    REF.ndlqr_SolveRiccati = getattr(REF, "ndlqr_SolveRiccati")


def ref_problem_from_flat(n, m, N, knots, x0):
    """Build an LQRProblem* inside the reference library from flat knot blocks."""
    st = 2 * n + 2 * m + 1 + n * n + n * m + n
    pp = REF.ndlqr_NewLQRProblem(n, m, N)
    knots = np.ascontiguousarray(knots, dtype=np.float64).reshape(N, st)
    for k in range(N):
        C.memmove(pp.contents.lqrdata[k].contents.Q, knots[k].ctypes.data, st * 8)
    x0 = np.ascontiguousarray(x0, dtype=np.float64)
    C.memmove(pp.contents.x0, x0.ctypes.data, n * 8)
    return pp


def ref_solve(n, m, N, knots, x0, threads=1, want_fact=True):
    """ndlqr_Solve of the unmodified reference; returns (soln[N*zs], fact[K,N,fs] or None, solve_ms)."""
    pp = ref_problem_from_flat(n, m, N, knots, x0)
    sp = REF.ndlqr_NewNdLqrSolver(n, m, N)
    REF.ndlqr_InitializeWithLQRProblem(pp, sp)
    REF.ndlqr_SetNumThreads(sp, threads)
    REF.ndlqr_Solve(sp)
    s = sp.contents
    K = s.depth
    zs, fs = 2 * n + m, (2 * n + m) * n
    soln = np.ctypeslib.as_array(s.soln.contents.data, shape=(N * zs,)).copy()
    fact = np.ctypeslib.as_array(s.fact.contents.data, shape=(K, N, fs)).copy() if want_fact else None
    ms = s.solve_time_ms
    REF.ndlqr_FreeNdLqrSolver(sp)
    REF.ndlqr_FreeLQRProblem(pp)
    return soln, fact, ms
