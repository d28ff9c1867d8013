"""ctypes binding of librslqr.so (include/rslqr_cuda.h, include/solver.h, include/solve.h, ...).

Mirrors the reference's operator surface for the solve path: ``NdLqrSolver`` wraps
ndlqr_NewNdLqrSolver / ndlqr_InitializeWithLQRProblem / ndlqr_Solve / ndlqr_GetSolution
(reference src/solver.h:131-157, src/solve.h:37-72) and ``CudaSolver`` wraps the batched C-ABI.
"""
import ctypes as C
import os

import numpy as np

# RSLQR_LIB: another build of the same library (A/B builds of a kernel variant, tools/build_variant.sh)
LIB_PATH = os.environ.get("RSLQR_LIB") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "librslqr.so")
if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
        "(make -C rslqr_b200/csrc). There is no Python or CPU fallback for the solve path.")
LIB = C.CDLL(LIB_PATH)

MODE_AUTO, MODE_STAGED, MODE_FUSED, MODE_SCAN = 0, 1, 2, 3
KEEP_FACT, PROFILE = 1, 2
STAGE = dict(PACK_RHS=0, LEAVES=1, PRODUCTS=2, CHOLESKY=3, SOLVES=4, SCHUR=5,
             RHS_PRODUCTS=6, RHS_SOLVES=7, RHS_SCHUR=8)

c_dp = C.POINTER(C.c_double)
c_ip = C.POINTER(C.c_int)


class RslqrError(RuntimeError):
    pass


def _sig(name, restype, *argtypes):
    fn = getattr(LIB, name)
    fn.restype = restype
    fn.argtypes = list(argtypes)
    return fn


_sig("rslqr_cuda_last_error", C.c_char_p)
_sig("rslqr_cuda_device_count", C.c_int)
_sig("rslqr_cuda_knot_stride", C.c_int, C.c_int, C.c_int)
_sig("rslqr_cuda_create", C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint, C.POINTER(C.c_void_p))
_sig("rslqr_cuda_destroy", C.c_int, C.c_void_p)
_sig("rslqr_cuda_upload", C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int)
_sig("rslqr_cuda_upload_rhs", C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int)
_sig("rslqr_cuda_reset", C.c_int, C.c_void_p)
_sig("rslqr_cuda_set_mode", C.c_int, C.c_void_p, C.c_int)
_sig("rslqr_cuda_ensure_fact", C.c_int, C.c_void_p)
_sig("rslqr_cuda_zero_fact", C.c_int, C.c_void_p)
_sig("rslqr_cuda_has_factorization", C.c_int, C.c_void_p)
_sig("rslqr_cuda_invalidate_factorization", C.c_int, C.c_void_p)
_sig("rslqr_cuda_get_mode", C.c_int, C.c_void_p)
_sig("rslqr_cuda_solve", C.c_int, C.c_void_p, C.c_int)
_sig("rslqr_cuda_solve_async", C.c_int, C.c_void_p, C.c_int)
_sig("rslqr_cuda_sync", C.c_int, C.c_void_p)
_sig("rslqr_cuda_factor_async", C.c_int, C.c_void_p)
_sig("rslqr_cuda_riccati_solve_async", C.c_int, C.c_void_p)
_sig("rslqr_cuda_solve_rhs_async", C.c_int, C.c_void_p, C.c_int)
_sig("rslqr_cuda_solve_host", C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int)
_sig("rslqr_cuda_solve_static", C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
     C.c_void_p, C.c_void_p)
_sig("rslqr_cuda_stage", C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int)
_sig("rslqr_cuda_schur_task", C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int)
_sig("rslqr_cuda_download_solution", C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int)
_sig("rslqr_cuda_download_fact", C.c_int, C.c_void_p, C.c_void_p, C.c_int)
_sig("rslqr_cuda_upload_fact", C.c_int, C.c_void_p, C.c_void_p, C.c_int)
_sig("rslqr_cuda_download_status", C.c_int, C.c_void_p, C.c_void_p, C.c_int)
_sig("rslqr_cuda_download_flags", C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int)
_sig("rslqr_cuda_download_leaf_factors", C.c_int, C.c_void_p, C.c_void_p, C.c_int)
_sig("rslqr_cuda_get_profile", C.c_int, C.c_void_p, c_dp)
_sig("rslqr_cuda_device_pointers", C.c_int, C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p),
     C.POINTER(C.c_void_p), C.POINTER(C.c_void_p))
_sig("rslqr_cuda_stream", C.c_void_p, C.c_void_p)
_sig("rslqr_cuda_kernel_launches", C.c_int, C.c_void_p)
_sig("rslqr_cuda_kernel_timing", C.c_int, C.c_void_p, C.c_int, C.c_int, c_dp, c_ip)
_sig("rslqr_cuda_mark", C.c_int, C.c_void_p, C.c_int)
_sig("rslqr_cuda_elapsed_ms", C.c_double, C.c_void_p, C.c_int, C.c_int)
_sig("rslqr_cuda_upload_async", C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int)
_sig("rslqr_cuda_download_solution_async", C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int)
_sig("rslqr_cuda_host_alloc", C.c_void_p, C.c_size_t)
_sig("rslqr_cuda_host_free", None, C.c_void_p)
_sig("rslqr_cuda_host_register", C.c_int, C.c_void_p, C.c_size_t)
_sig("rslqr_cuda_host_unregister", C.c_int, C.c_void_p)
_sig("rslqr_cuda_create_shard", C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint, C.POINTER(C.c_void_p))
_sig("rslqr_cuda_shard_payload_doubles", C.c_size_t, C.c_void_p)
_sig("rslqr_cuda_shard_reduce", C.c_int, C.c_void_p, C.c_void_p)
_sig("rslqr_cuda_shard_finish", C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int)
_sig("rslqr_cuda_shard_push", C.c_int, C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.c_int, C.c_int, C.c_uint64)
_sig("rslqr_cuda_shard_finish_wait", C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.c_int)
_sig("rslqr_cuda_shard_exchange", C.c_int, C.POINTER(C.c_void_p), C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p))
_sig("rslqr_cuda_set_stream", C.c_int, C.c_void_p, C.c_void_p)
_sig("rslqr_synth_fill", C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint64, C.c_int, C.c_int, C.c_void_p, C.c_void_p)


def knot_stride(n, m):
    return LIB.rslqr_cuda_knot_stride(n, m)


def _check(rc, what):
    if rc != 0:
        raise RslqrError(f"{what} failed ({rc}): {LIB.rslqr_cuda_last_error().decode()}")


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def pinned_array(shape, dtype=np.float64):
    """numpy array over cudaHostAlloc'ed memory (freed when the array's base is collected)."""
    count = int(np.prod(shape))
    nbytes = max(count * np.dtype(dtype).itemsize, 8)
    p = LIB.rslqr_cuda_host_alloc(nbytes)
    if not p:
        raise RslqrError("pinned allocation failed: " + LIB.rslqr_cuda_last_error().decode())
    buf = (C.c_char * nbytes).from_address(p)
    arr = np.frombuffer(buf, dtype=dtype, count=count).reshape(shape)
    _PINNED[id(buf)] = (buf, p)
    return arr


_PINNED = {}


class CudaSolver:
    """Batched device solver: one handle = `batch` problems of shape (n, m, N) on one GPU."""

    def __init__(self, n, m, N, batch=1, device=0, flags=KEEP_FACT):
        self.n, self.m, self.N, self.batch, self.device = n, m, N, batch, device
        self.K = N.bit_length() - 1
        self.fs, self.zs, self.stride = (2 * n + m) * n, 2 * n + m, knot_stride(n, m)
        self.h = C.c_void_p()
        _check(LIB.rslqr_cuda_create(n, m, N, batch, device, flags, C.byref(self.h)), "rslqr_cuda_create")

    def close(self):
        if self.h:
            LIB.rslqr_cuda_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- data movement
    def upload(self, knots, x0, first=0):
        knots = np.ascontiguousarray(knots, dtype=np.float64)
        x0 = np.ascontiguousarray(x0, dtype=np.float64)
        count = x0.size // self.n
        assert knots.size == count * self.N * self.stride, "knots must be [count][N][stride]"
        _check(LIB.rslqr_cuda_upload(self.h, _ptr(knots), _ptr(x0), first, count), "upload")

    def upload_rhs(self, soln, first=0):
        soln = np.ascontiguousarray(soln, dtype=np.float64)
        _check(LIB.rslqr_cuda_upload_rhs(self.h, _ptr(soln), first, soln.size // (self.N * self.zs)), "upload_rhs")

    def reset(self):
        _check(LIB.rslqr_cuda_reset(self.h), "reset")

    def set_mode(self, mode):
        _check(LIB.rslqr_cuda_set_mode(self.h, mode), "set_mode")

    @property
    def mode(self):
        return LIB.rslqr_cuda_get_mode(self.h)

    # -- compute
    def solve(self, pack_rhs=True):
        _check(LIB.rslqr_cuda_solve(self.h, int(pack_rhs)), "solve")

    def solve_async(self, pack_rhs=True):
        _check(LIB.rslqr_cuda_solve_async(self.h, int(pack_rhs)), "solve_async")

    def sync(self):
        _check(LIB.rslqr_cuda_sync(self.h), "sync")

    # factor once, solve many (staged mode): include/rslqr_cuda.h
    def factor(self):
        _check(LIB.rslqr_cuda_factor_async(self.h), "factor")
        self.sync()

    def riccati(self):
        """Comparison solver: Riccati recursion on the device (n=12, m=4), same output layout as solve()."""
        _check(LIB.rslqr_cuda_riccati_solve_async(self.h), "riccati")
        self.sync()

    def solve_rhs(self, pack_rhs=False):
        _check(LIB.rslqr_cuda_solve_rhs_async(self.h, int(pack_rhs)), "solve_rhs")
        self.sync()

    def solve_host(self, knots, x0, soln, count=None, chunk=0):
        """host buffers in, host solutions out, copies and kernels pipelined; asynchronous (call sync())"""
        count = self.batch if count is None else count
        for a in (knots, x0, soln):
            assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"]
        assert knots.size >= count * self.N * self.stride and x0.size >= count * self.n and soln.size >= count * self.N * self.zs
        _check(LIB.rslqr_cuda_solve_host(self.h, _ptr(knots), _ptr(x0), _ptr(soln), count, chunk), "solve_host")

    def stage(self, name, level=0, item=-1, col=-1, problem=-1):
        _check(LIB.rslqr_cuda_stage(self.h, STAGE[name], level, item, col, problem), "stage " + name)

    def schur_task(self, index, knot, level, upper_level, calc_lambda, use_rhs=False, problem=-1):
        _check(LIB.rslqr_cuda_schur_task(self.h, index, knot, level, upper_level, int(calc_lambda),
                                         int(use_rhs), problem), "schur_task")

    # -- results
    def solution(self, first=0, count=None, out=None):
        count = self.batch - first if count is None else count
        if out is None:
            out = np.empty((count, self.N * self.zs))
        _check(LIB.rslqr_cuda_download_solution(self.h, _ptr(out), first, count), "download_solution")
        return out

    def fact(self, problem=0):
        out = np.empty((self.K, self.N, self.fs))
        _check(LIB.rslqr_cuda_download_fact(self.h, _ptr(out), problem), "download_fact")
        return out

    def status(self, problem=0):
        out = np.zeros(3 * self.N - 1, dtype=np.int32)
        _check(LIB.rslqr_cuda_download_status(self.h, _ptr(out), problem), "download_status")
        return out

    def flags(self):
        out = np.zeros(self.batch, dtype=np.int32)
        _check(LIB.rslqr_cuda_download_flags(self.h, _ptr(out), 0, self.batch), "download_flags")
        return out

    def leaf_factors(self, problem=0):
        out = np.empty((self.N, self.n + self.m))
        _check(LIB.rslqr_cuda_download_leaf_factors(self.h, _ptr(out), problem), "download_leaf_factors")
        return out

    def profile(self):
        ms = (C.c_double * 7)()
        _check(LIB.rslqr_cuda_get_profile(self.h, ms), "get_profile")
        return dict(zip(("total", "leaves", "products", "cholesky", "solves", "schur", "rhs"), list(ms)))

    def device_pointers(self):
        p = [C.c_void_p() for _ in range(4)]
        _check(LIB.rslqr_cuda_device_pointers(self.h, *[C.byref(x) for x in p]), "device_pointers")
        return dict(zip(("knots", "x0", "soln", "fact"), [x.value for x in p]))

    def mark(self, slot):
        _check(LIB.rslqr_cuda_mark(self.h, slot), "mark")

    def elapsed_ms(self, a, b):
        ms = LIB.rslqr_cuda_elapsed_ms(self.h, a, b)
        if ms < 0:
            raise RslqrError("elapsed_ms failed")
        return ms

    def upload_async(self, knots, x0, first=0):
        count = x0.size // self.n
        _check(LIB.rslqr_cuda_upload_async(self.h, _ptr(knots), _ptr(x0), first, count), "upload_async")

    def solution_async(self, out, first=0, count=None):
        count = self.batch - first if count is None else count
        _check(LIB.rslqr_cuda_download_solution_async(self.h, _ptr(out), first, count), "download_solution_async")

    @property
    def stream(self):
        return LIB.rslqr_cuda_stream(self.h)

    @property
    def launches(self):
        return LIB.rslqr_cuda_kernel_launches(self.h)

    def kernel_timing(self, enable=True, reset=False):
        ms, cnt = C.c_double(), C.c_int()
        _check(LIB.rslqr_cuda_kernel_timing(self.h, int(enable), int(reset), C.byref(ms), C.byref(cnt)), "kernel_timing")
        return ms.value, cnt.value


def synth_problems(n, m, N, seed, first_id, count, knots=None, x0=None):
    """Flat synthetic problems: returns (knots[count,N,stride], x0[count,n])."""
    st = knot_stride(n, m)
    if knots is None:
        knots = np.empty((count, N, st))
    if x0 is None:
        x0 = np.empty((count, n))
    rc = LIB.rslqr_synth_fill(n, m, N, seed, first_id, count, _ptr(knots), _ptr(x0))
    if rc != 0:
        raise RslqrError("rslqr_synth_fill failed")
    return knots, x0


from .ctypes_api import (  # noqa: E402,F401  (struct layouts + prototype table, shared with the test-side oracle loader)
    Matrix, LQRData, LQRProblem, NdFactor, NdData, CholeskyInfo, NdLqrCholeskyFactors, UnitRange, BinaryNode,
    OrderedBinaryTree, NdLqrProfile, RefSolverStruct, SolverStruct, bind_host_api)

bind_host_api(LIB, SolverStruct)


def problem_to_flat(prob_ptr, n, m):
    """(knots[N,stride], x0[n]) numpy copies of an LQRProblem* (either library's)."""
    pr = prob_ptr.contents
    N, st = pr.nhorizon, 2 * n + 2 * m + 1 + n * n + n * m + n
    knots = np.empty((N, st))
    for k in range(N):
        knots[k] = np.ctypeslib.as_array(pr.lqrdata[k].contents.Q, shape=(st,))
    x0 = np.ctypeslib.as_array(pr.x0, shape=(n,)).copy()
    return knots, x0


def write_problems_binary(path, n, m, N, knots, x0):
    """Write flat problems (knots[count,N,stride], x0[count,n]) to the binary container of include/rslqr_batch.h."""
    knots = np.ascontiguousarray(knots, dtype=np.float64)
    x0 = np.ascontiguousarray(x0, dtype=np.float64)
    count = x0.size // n
    if LIB.rslqr_WriteProblemsBinary(os.fsencode(path), n, m, N, count, _ptr(knots), _ptr(x0)) != 0:
        raise RslqrError(f"could not write {path}")


def read_problems_binary(path, first=0, count=None, knots=None, x0=None):
    """Read problems [first, first+count) of a binary container. Returns (n, m, N, knots[count,N,stride], x0[count,n]);
    `knots` / `x0` may be preallocated (e.g. pinned) arrays."""
    n, m, N, total = (C.c_int() for _ in range(4))
    if LIB.rslqr_ReadProblemsBinaryHeader(os.fsencode(path), C.byref(n), C.byref(m), C.byref(N), C.byref(total)) != 0:
        raise RslqrError(f"could not read {path}")
    n, m, N, total = n.value, m.value, N.value, total.value
    count = total - first if count is None else count
    if knots is None:
        knots = np.empty((count, N, knot_stride(n, m)))
    if x0 is None:
        x0 = np.empty((count, n))
    if LIB.rslqr_ReadProblemsBinary(os.fsencode(path), first, count, _ptr(knots), _ptr(x0)) != 0:
        raise RslqrError(f"could not read problems [{first}, {first + count}) from {path}")
    return n, m, N, knots, x0


def load_problem_json(path):
    """Read a reference-format problem file through the library's JSON loader.
    Returns (n, m, N, knots[N,stride], x0[n])."""
    pp = LIB.ndlqr_ReadLQRProblemJSONFile(os.fsencode(path))
    if not pp:
        raise RslqrError(f"could not read LQR problem from {path}")
    try:
        n, m = pp.contents.lqrdata[0].contents.nstates, pp.contents.lqrdata[0].contents.ninputs
        knots, x0 = problem_to_flat(pp, n, m)
        return n, m, pp.contents.nhorizon, knots, x0
    finally:
        LIB.ndlqr_FreeLQRProblem(pp)


class NdLqrSolver:
    """The reference's single-problem flow (reference src/solver.h:84-96) through the C host API."""

    def __init__(self, n, m, N, lib=LIB):
        self.lib = lib
        self.p = lib.ndlqr_NewNdLqrSolver(n, m, N)
        if not self.p:
            raise RslqrError("ndlqr_NewNdLqrSolver returned NULL")
        self.n, self.m, self.N = n, m, N

    @property
    def c(self):
        return self.p.contents

    def initialize(self, prob_ptr):
        return self.lib.ndlqr_InitializeWithLQRProblem(prob_ptr, self.p)

    def reset(self):
        self.lib.ndlqr_ResetSolver(self.p)

    def solve(self):
        return self.lib.ndlqr_Solve(self.p)

    def sync_fact(self):
        """solver->fact <- the reference's dense factor array of the initialized problem (include/solve.h)"""
        return self.lib.ndlqr_SyncFactMirror(self.p)

    def solution(self):
        mat = self.lib.ndlqr_GetSolution(self.p)
        return np.ctypeslib.as_array(mat.data, shape=(mat.rows,)).copy()

    def soln_array(self):
        return np.ctypeslib.as_array(self.c.soln.contents.data, shape=(self.N * (2 * self.n + self.m),))

    def nddata_array(self, which):
        nd = getattr(self.c, which).contents
        K = nd.depth
        return np.ctypeslib.as_array(nd.data, shape=(K, self.N, (2 * self.n + self.m) * nd.width))

    def close(self):
        if self.p:
            self.lib.ndlqr_FreeNdLqrSolver(self.p)
            self.p = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class ShardSolver(CudaSolver):
    """One shard (knots [rank*shard_knots, (rank+1)*shard_knots)) of a single long-horizon problem on one GPU."""

    def __init__(self, n, m, N, rank, nranks, device=0):
        assert N % nranks == 0
        Nl = N // nranks
        self.n, self.m, self.N, self.batch, self.device = n, m, Nl, 1, device
        self.Nglobal, self.rank, self.nranks = N, rank, nranks
        self.K = Nl.bit_length() - 1
        self.fs, self.zs, self.stride = (2 * n + m) * n, 2 * n + m, knot_stride(n, m)
        self.h = C.c_void_p()
        _check(LIB.rslqr_cuda_create_shard(n, m, N, rank * Nl, Nl, device, 0, C.byref(self.h)), "rslqr_cuda_create_shard")
        self.payload_doubles = LIB.rslqr_cuda_shard_payload_doubles(self.h)

    def upload_shard(self, knots_full, x0):
        """knots_full: [N, stride] of the whole problem; takes this shard's slice and its boundary vector"""
        lo = self.rank * self.N
        head = x0 if self.rank == 0 else knots_full[lo - 1, -self.n:]
        self.upload(np.ascontiguousarray(knots_full[lo:lo + self.N])[None], np.ascontiguousarray(head)[None])

    def reduce(self, d_payload_ptr):
        _check(LIB.rslqr_cuda_shard_reduce(self.h, C.c_void_p(d_payload_ptr)), "shard_reduce")

    def finish(self, d_gathered_ptr):
        _check(LIB.rslqr_cuda_shard_finish(self.h, C.c_void_p(d_gathered_ptr), self.rank, self.nranks), "shard_finish")

    def set_stream(self, stream_ptr):
        _check(LIB.rslqr_cuda_set_stream(self.h, C.c_void_p(stream_ptr)), "set_stream")

    def push(self, d_payload_ptr, peer_gathered_ptrs, peer_flag_ptrs, epoch):
        """peer stores of this shard's payload into every rank's gathered buffer + flag (NVLink peer mappings)"""
        n = self.nranks
        pg = (C.c_void_p * n)(*peer_gathered_ptrs)
        pf = (C.c_void_p * n)(*peer_flag_ptrs)
        _check(LIB.rslqr_cuda_shard_push(self.h, C.c_void_p(d_payload_ptr), pg, pf, self.rank, n, epoch), "shard_push")

    def finish_wait(self, d_gathered_ptr, d_flags_ptr, epoch):
        _check(LIB.rslqr_cuda_shard_finish_wait(self.h, C.c_void_p(d_gathered_ptr), C.c_void_p(d_flags_ptr), epoch,
                                                self.rank, self.nranks), "shard_finish_wait")


def shard_exchange(shards, payload_ptrs, gathered_ptrs):
    """single-process all-gather of the shard payloads (peer copies)"""
    n = len(shards)
    hs = (C.c_void_p * n)(*[s.h for s in shards])
    pp = (C.c_void_p * n)(*payload_ptrs)
    gp = (C.c_void_p * n)(*gathered_ptrs)
    _check(LIB.rslqr_cuda_shard_exchange(hs, n, pp, gp), "shard_exchange")
