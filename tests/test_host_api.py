"""CPU tests of the C host layer and of the C-ABI surface (no compute calls: no GPU here).
Mirrors test/lqrdata_test.c, test/utils_test.c, test/nddata_test.c, test/binarytree_test.c,
test/matrix_test.c, test/solver_test.c (placement part) of the reference."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import rslqr_b200 as R
from rslqr_b200 import lib as L
from helpers import GOLDEN, golden_json, jmat, problem_from_json

LIB = R.LIB
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_every_declared_symbol_is_exported():
    """Each function declared in include/*.h must be exported by librslqr.so."""
    missing = []
    for hdr in sorted(os.listdir(os.path.join(ROOT, "include"))):
        text = open(os.path.join(ROOT, "include", hdr)).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        text = re.sub(r"static inline[^{]*\{[^}]*\}", "", text)
        for name in re.findall(r"\b([A-Za-z_][A-Za-z0-9_]*)\s*\([^;{]*\)\s*;", text):
            if name in ("defined", "sizeof"):
                continue
            if not hasattr(LIB, name):
                missing.append(f"{hdr}:{name}")
    assert not missing, missing


def test_cabi_reports_no_device_without_fallback():
    if LIB.rslqr_cuda_device_count() > 0:
        pytest.skip("a GPU is present")
    h = C.c_void_p()
    rc = LIB.rslqr_cuda_create(6, 3, 8, 1, 0, 0, C.byref(h))
    assert rc == -2 and not h.value
    assert b"no CPU fallback" in LIB.rslqr_cuda_last_error()
    with pytest.raises(R.RslqrError):
        R.CudaSolver(6, 3, 8)


def test_cabi_argument_errors():
    h = C.c_void_p()
    assert LIB.rslqr_cuda_create(6, 3, 7, 1, 0, 0, C.byref(h)) == -1      # not a power of two
    assert LIB.rslqr_cuda_create(0, 3, 8, 1, 0, 0, C.byref(h)) == -1
    assert LIB.rslqr_cuda_create(6, 3, 8, 1, 0, 0, None) == -1
    assert LIB.rslqr_cuda_destroy(None) == -1
    assert LIB.rslqr_cuda_solve(None, 1) == -1
    assert LIB.rslqr_cuda_knot_stride(6, 3) == 79
    assert LIB.rslqr_cuda_knot_stride(12, 4) == 237


def test_json_problem_loader():
    # test/lqrdata_test.c:58-98: values of lqr_prob.json incl. x0 and the terminal knot
    n, m, N, knots, x0 = R.load_problem_json(os.path.join(GOLDEN, "lqr_prob.json"))
    pn, pm, pN, pknots, px0, d = problem_from_json("lqr_prob.json")
    assert (n, m, N) == (pn, pm, pN) == (6, 3, 8)
    assert np.array_equal(knots, pknots) and np.array_equal(x0, px0)
    assert knots[N - 1, 0] == 10.0          # terminal Q
    assert np.array_equal(x0, [1, -1, 2, -2, 3, -3])


def test_json_problem_loader_256():
    n, m, N, knots, x0 = R.load_problem_json(os.path.join(GOLDEN, "lqr_prob_256.json"))
    pn, pm, pN, pknots, px0, _ = problem_from_json("lqr_prob_256.json")
    assert N == 256 and np.array_equal(knots, pknots) and np.array_equal(x0, px0)


def test_json_lqrdata_and_matrix_loader():
    # test/lqrdata_test.c:16-38 and the "test" matrix == 1..12 column-major
    p = LIB.ndlqr_ReadLQRDataJSONFile(os.path.join(GOLDEN, "lqr_data.json").encode())
    assert p
    d = golden_json("lqr_data.json")
    kd = p.contents
    assert kd.nstates == 6 and kd.ninputs == 3
    assert [kd.Q[i] for i in range(6)] == d["Q"]
    assert kd.c[0] == d["c"]
    A = np.ctypeslib.as_array(kd.A, shape=(36,)).reshape(6, 6, order="F")
    assert np.array_equal(A, jmat(d["A"]))
    LIB.ndlqr_FreeLQRData(p)
    mat = LIB.ReadMatrixJSONFile(os.path.join(GOLDEN, "sample_problem.json").encode(), b"test")
    assert (mat.rows, mat.cols) == (3, 4)
    assert [mat.data[i] for i in range(12)] == list(range(1, 13))
    LIB.FreeMatrix(C.byref(mat))


def test_json_loader_errors(tmp_path):
    assert not LIB.ndlqr_ReadLQRProblemJSONFile(b"/nonexistent/file.json")
    bad = tmp_path / "bad.json"
    bad.write_text('{"nhorizon": 8, "lqrdata": [ {"nstates": 6 ')
    assert not LIB.ndlqr_ReadLQRProblemJSONFile(str(bad).encode())
    mat = LIB.ReadMatrixJSONFile(os.path.join(GOLDEN, "sample_problem.json").encode(), b"nope")
    assert mat.rows == 0 and mat.cols == 0 and not mat.data
    short = tmp_path / "short.json"
    short.write_text('{"nstates": 2, "ninputs": 1, "c": 0, "Q": [1], "R": [1], "q": [0,0], "r": [0],'
                     ' "d": [0,0], "A": [[1,0],[0,1]], "B": [[0,1]]}')
    assert not LIB.ndlqr_ReadLQRDataJSONFile(str(short).encode())   # Q too short


def test_binary_problem_container(tmp_path):
    # flat binary container; round trips, range reads, LQRProblem <-> file, error paths
    n, m, N = 12, 4, 16
    kn, x0 = R.synth_problems(n, m, N, 5, 0, 6)
    path = str(tmp_path / "batch.rslqr")
    L.write_problems_binary(path, n, m, N, kn, x0)
    assert os.path.getsize(path) == 64 + 8 * 6 * (n + N * R.knot_stride(n, m))
    rn, rm, rN, rk, rx = L.read_problems_binary(path)
    assert (rn, rm, rN) == (n, m, N) and np.array_equal(rk, kn) and np.array_equal(rx, x0)
    _, _, _, rk, rx = L.read_problems_binary(path, first=2, count=3)
    assert np.array_equal(rk, kn[2:5]) and np.array_equal(rx, x0[2:5])
    with pytest.raises(R.RslqrError):
        L.read_problems_binary(path, first=4, count=3)            # past the end
    # one record as an LQRProblem of the reference's struct layout, and back through the LQRProblem writer
    pp = LIB.rslqr_ReadLQRProblemBinary(path.encode(), 3)
    k3, x3 = L.problem_to_flat(pp, n, m)
    assert pp.contents.nhorizon == N and np.array_equal(k3, kn[3]) and np.array_equal(x3, x0[3])
    arr = (C.POINTER(L.LQRProblem) * 2)(pp, pp)
    path2 = str(tmp_path / "two.rslqr")
    assert LIB.rslqr_WriteLQRProblemsBinary(path2.encode(), arr, 2) == 0
    _, _, _, rk, rx = L.read_problems_binary(path2)
    assert np.array_equal(rk[0], kn[3]) and np.array_equal(rk[1], kn[3]) and np.array_equal(rx[1], x0[3])
    LIB.ndlqr_FreeLQRProblem(pp)
    # the JSON problem of the reference survives JSON -> binary -> LQRProblem
    jp = LIB.ndlqr_ReadLQRProblemJSONFile(os.path.join(GOLDEN, "lqr_prob.json").encode())
    arr1 = (C.POINTER(L.LQRProblem) * 1)(jp)
    path3 = str(tmp_path / "json.rslqr")
    assert LIB.rslqr_WriteLQRProblemsBinary(path3.encode(), arr1, 1) == 0
    bp = LIB.rslqr_ReadLQRProblemBinary(path3.encode(), 0)
    ka, xa = L.problem_to_flat(jp, 6, 3)
    kb, xb = L.problem_to_flat(bp, 6, 3)
    assert np.array_equal(ka, kb) and np.array_equal(xa, xb)
    LIB.ndlqr_FreeLQRProblem(jp)
    LIB.ndlqr_FreeLQRProblem(bp)
    # error paths: missing file, wrong magic, truncated payload, bad arguments
    assert not LIB.rslqr_ReadLQRProblemBinary(b"/nonexistent/file.rslqr", 0)
    bad = tmp_path / "bad.rslqr"
    bad.write_bytes(b"NOTRSLQR" + bytes(56))
    assert LIB.rslqr_ReadProblemsBinaryHeader(str(bad).encode(), None, None, None, None) == -1
    trunc = tmp_path / "trunc.rslqr"
    trunc.write_bytes(open(path, "rb").read()[:-100])
    with pytest.raises(R.RslqrError):
        L.read_problems_binary(str(trunc))
    assert not LIB.rslqr_ReadLQRProblemBinary(path.encode(), 6)     # index == count
    assert LIB.rslqr_WriteProblemsBinary(None, n, m, N, 1, None, None) == -1


def test_nddata_layout():
    # test/nddata_test.c:15-28,61-126
    assert not LIB.ndlqr_NewNdData(6, 3, 7, 6)
    assert not LIB.ndlqr_NewNdData(0, 3, 8, 6)
    nd = LIB.ndlqr_NewNdData(6, 3, 8, 6)
    assert nd.contents.depth == 3 and nd.contents.nsegments == 7
    f = C.POINTER(L.NdFactor)()
    assert LIB.ndlqr_GetNdFactor(nd, 2, 1, C.byref(f)) == 0
    base = C.addressof(nd.contents.data.contents)
    fs = (2 * 6 + 3) * 6
    assert C.addressof(f.contents.lambda_.data.contents) - base == (2 + 8 * 1) * fs * 8
    assert C.addressof(f.contents.state.data.contents) - base == ((2 + 8) * fs + 36) * 8
    assert C.addressof(f.contents.input.data.contents) - base == ((2 + 8) * fs + 72) * 8
    assert (f.contents.input.rows, f.contents.input.cols) == (3, 6)
    assert LIB.ndlqr_GetNdFactor(nd, 8, 0, C.byref(f)) == -1
    assert LIB.ndlqr_GetNdFactor(nd, 0, 3, C.byref(f)) == -1
    LIB.ndlqr_FreeNdData(nd)
    rhs = LIB.ndlqr_NewNdData(6, 3, 8, 1)
    assert rhs.contents.depth == 1
    LIB.ndlqr_GetNdFactor(rhs, 3, 0, C.byref(f))
    assert C.addressof(f.contents.state.data.contents) - C.addressof(rhs.contents.data.contents) == (3 * 15 + 6) * 8
    LIB.ndlqr_FreeNdData(rhs)


def test_tree_matches_reference_shape():
    # test/binarytree_test.c:4-61
    t = LIB.ndlqr_BuildTree(8)
    assert t.num_elements == 8 and t.depth == 3
    root = t.root.contents
    assert (root.idx, root.level) == (3, 2)
    assert (root.left_inds.start, root.left_inds.stop, root.right_inds.start, root.right_inds.stop) == (0, 3, 4, 7)
    assert root.left_child.contents.idx == 1 and root.left_child.contents.level == 1
    assert root.left_child.contents.left_child.contents.idx == 0
    assert root.right_child.contents.idx == 5
    assert root.right_child.contents.right_child.contents.idx == 6
    assert [LIB.ndlqr_GetIndexLevel(C.byref(t), k) for k in range(7)] == [0, 1, 0, 2, 0, 1, 0]
    assert LIB.ndlqr_GetIndexAtLevel(C.byref(t), 5, 0) == 4
    assert LIB.ndlqr_GetIndexAtLevel(C.byref(t), 2, 2) == 3
    assert LIB.ndlqr_GetIndexAtLevel(C.byref(t), 7, 0) == 6
    assert LIB.ndlqr_GetIndexFromLeaf(C.byref(t), 1, 1) == 5
    # ShouldCalcLambda == closed form used on the device
    for lvl in range(3):
        for k in range(8):
            idx = LIB.ndlqr_GetIndexAtLevel(C.byref(t), k, lvl)
            a = (k >> (lvl + 1)) << (lvl + 1)
            want = (k == 0) or not (k == a or k == idx + 1)
            assert bool(LIB.ndlqr_ShouldCalcLambda(C.byref(t), idx, k)) == want
    LIB.ndlqr_FreeTree(C.byref(t))


def test_cholesky_factor_slots():
    cf = LIB.ndlqr_NewCholeskyFactors(3, 8)
    assert cf.contents.numfacts == 16 + 7
    info = C.POINTER(L.CholeskyInfo)()
    base = C.addressof(cf.contents.cholinfo.contents)
    sz = C.sizeof(L.CholeskyInfo)
    for (fn, args, slot) in [(LIB.ndlqr_GetQFactorizon, (3,), 6), (LIB.ndlqr_GetRFactorizon, (3,), 7),
                             (LIB.ndlqr_GetSFactorization, (1, 1), 16 + 4 + 1), (LIB.ndlqr_GetSFactorization, (0, 2), 22)]:
        assert fn(cf, *args, C.byref(info)) == 0
        assert (C.addressof(info.contents) - base) // sz == slot
    assert LIB.ndlqr_GetRFactorizon(cf, 7, C.byref(info)) == -1
    assert LIB.ndlqr_GetSFactorization(cf, 4, 0, C.byref(info)) == -1
    LIB.ndlqr_FreeCholeskyFactors(cf)


def test_solver_init_placement_without_gpu(capfd):
    # test/solver_test.c:23,49-143. Works without a GPU: Initialize only fills mirrors; Solve must fail loudly.
    pp = LIB.ndlqr_ReadLQRProblemJSONFile(os.path.join(GOLDEN, "lqr_prob.json").encode())
    s = R.NdLqrSolver(6, 3, 8)
    assert s.c.nvars == 117 and s.c.depth == 3
    has_gpu = LIB.rslqr_cuda_device_count() > 0
    assert s.initialize(pp) == (0 if has_gpu or not s.c.device else 0)
    d = golden_json("lqr_prob.json")
    data = s.nddata_array("data")
    n, m = 6, 3
    X = data[0, 0][36:72].reshape(6, 6, order="F")
    assert np.array_equal(X, jmat(d["lqrdata"][0]["A"]).T)
    assert np.array_equal(data[0, 1][36:72].reshape(6, 6, order="F"), -np.eye(6))
    assert np.array_equal(data[2, 4][36:72].reshape(6, 6, order="F"), -np.eye(6))
    z = s.soln_array().reshape(8, 15)
    assert np.array_equal(z[0, :6], -np.array(d["x0"]))
    assert np.array_equal(z[3, :6], -np.array(d["lqrdata"][2]["d"]))
    assert np.array_equal(z[7, 6:12], -np.array(d["lqrdata"][7]["q"]))
    Q7 = np.ctypeslib.as_array(s.c.diagonals[14].data, shape=(36,)).reshape(6, 6)
    assert np.array_equal(np.diag(Q7), d["lqrdata"][7]["Q"])
    if not has_gpu:
        assert s.solve() == -1
        assert "no CPU fallback" in capfd.readouterr().err
    # horizon mismatch -> -1 (src/solver.c:125)
    s2 = R.NdLqrSolver(6, 3, 16)
    assert s2.initialize(pp) == -1
    LIB.ndlqr_FreeLQRProblem(pp)


def test_synth_generator_is_reproducible():
    k1, x1 = R.synth_problems(12, 4, 16, 2024, 5, 2)
    k2, x2 = R.synth_problems(12, 4, 16, 2024, 6, 1)
    assert np.array_equal(k1[1], k2[0]) and np.array_equal(x1[1], x2[0])
    assert not np.array_equal(k1[0], k1[1])
    st = R.knot_stride(12, 4)
    kn = k1[0]
    assert (kn[:-1, :12] > 0).all() and (kn[:-1, 12:16] > 0).all()      # Q, R > 0
    A = kn[0, 33:33 + 144].reshape(12, 12, order="F")
    assert np.array_equal(np.diag(A), np.ones(12)) and A[0, 6] == 0.02
    assert not kn[-1, 33:33 + 144 + 48].any()                            # terminal A, B = 0
    pp = LIB.rslqr_synth_problem(12, 4, 16, 2024, 5)
    kk, xx = L.problem_to_flat(pp, 12, 4)
    assert np.array_equal(kk, k1[0]) and np.array_equal(xx, x1[0])
    LIB.ndlqr_FreeLQRProblem(pp)


def _compile_consumer(src, out, extra=()):
    import subprocess
    libdir = os.path.join(ROOT, "rslqr_b200")
    cmd = ["/usr/bin/gcc", "-std=c11", "-O1", "-I" + os.path.join(ROOT, "include"), *extra, src, "-L" + libdir, "-lrslqr",
           "-Wl,-rpath," + libdir, "-lm", "-o", out]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return out


def test_reference_example_compiles_and_links(tmp_path):
    # drop-in at link level: a consumer written against the reference's headers builds
    # against include/ + librslqr.so; when the reference tree is mounted, its own example is built UNMODIFIED
    exe = _compile_consumer(os.path.join(ROOT, "tests", "c", "consumer_flow.c"), str(tmp_path / "consumer"))
    assert os.path.exists(exe)
    ref_example = "/root/reference/examples/installexample/main.c"
    if os.path.exists(ref_example):
        prob = os.path.join(GOLDEN, "lqr_prob.json")
        _compile_consumer(ref_example, str(tmp_path / "ref_example"), extra=[f'-DLQRPROBFILE="{prob}"'])
