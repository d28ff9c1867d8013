"""GPU parity tests (run on the B200 box): the CUDA path, called through the C-ABI and through the
reference-named C host API, against the oracle (oracle/ndlqr_port.c), the reference's golden vectors
and stored outputs of the unmodified reference. Tolerance (BASELINE.json north_star): relative
2-norm error <= 1e-10 per variable class (lambda, x, u); KKT residual no worse than the reference's."""
import ctypes as C
import os

import numpy as np
import pytest

import oracle_lib as O
import rslqr_b200 as R
from helpers import GOLDEN, blocks, golden_json, jmat, problem_from_json, rel_err_classes

pytestmark = pytest.mark.gpu
TOL = 1e-10


def _modes(solver):
    out = [R.MODE_STAGED]
    for mode in (R.MODE_FUSED, R.MODE_SCAN):
        try:
            solver.set_mode(mode)
            out.append(mode)
        except R.RslqrError:
            pass
    return out


def _solve_all_modes(n, m, N, knots, x0):
    """yield (mode, solutions[count, N*zs], solver) for every mode the shape supports"""
    count = x0.reshape(-1, n).shape[0]
    s = R.CudaSolver(n, m, N, batch=count, flags=R.KEEP_FACT)
    s.upload(knots, x0)
    for mode in _modes(s):
        s.set_mode(mode)
        s.reset()
        s.solve()
        assert not s.flags().any()
        yield mode, s.solution(), s
    s.close()


def test_stage_by_stage_vs_oracle():
    # the device state after every phase equals the oracle's (reference test/nested_dissection_test.c:138-215)
    n, m, N, knots, x0, _ = problem_from_json("lqr_prob.json")
    o = O.PortSolver(n, m, N)
    o.init(knots, x0)
    s = R.CudaSolver(n, m, N, flags=R.KEEP_FACT)
    s.upload(knots, x0)
    s.stage("PACK_RHS")
    assert np.array_equal(s.solution()[0], o.soln.reshape(-1))     # pure copies/negation: bit-exact
    s.stage("LEAVES"); o.leaves()
    assert np.allclose(s.fact(), o.fact, rtol=1e-13, atol=1e-15)
    assert np.allclose(s.solution()[0], o.soln.reshape(-1), rtol=1e-13, atol=1e-15)
    for lv in range(s.K):
        for name, fn in (("PRODUCTS", o.products), ("CHOLESKY", o.cholesky), ("SOLVES", o.solves), ("SCHUR", o.schur)):
            if lv == s.K - 1 and name in ("SOLVES", "SCHUR"):
                continue
            s.stage(name, level=lv); fn(lv)
            f = s.fact()
            assert np.allclose(f, o.fact, rtol=1e-11, atol=1e-12), (name, lv, np.abs(f - o.fact).max())
    for lv in range(s.K):
        s.stage("RHS_PRODUCTS", level=lv); s.stage("RHS_SOLVES", level=lv); s.stage("RHS_SCHUR", level=lv)
        o.rhs_level(lv)
        assert np.allclose(s.solution()[0], o.soln.reshape(-1), rtol=1e-10, atol=1e-11), lv
    s.close()


def test_stage_golden_fixtures():
    # sample_problem.json: "b" after the leaves, E{k}{p}{y,x,u} after the level-0 Schur update
    n, m, N, knots, x0, _ = problem_from_json("lqr_prob.json")
    sp = golden_json("sample_problem.json")
    s = R.CudaSolver(n, m, N, flags=R.KEEP_FACT)
    s.upload(knots, x0)
    s.stage("PACK_RHS"); s.stage("LEAVES")
    assert np.linalg.norm(s.solution()[0] - np.array(sp["b"]).reshape(-1)) < 1e-6
    for name in ("PRODUCTS", "CHOLESKY", "SOLVES", "SCHUR"):
        s.stage(name, level=0)
    f = s.fact()
    for k in (0, 1):
        for p in (1, 2):
            L, X, U = blocks(f[p, k], n, m)
            assert np.linalg.norm(L - jmat(sp[f"E{k}{p}y"])) < 1e-6
            assert np.linalg.norm(X - jmat(sp[f"E{k}{p}x"])) < 1e-6
            assert np.linalg.norm(U - jmat(sp[f"E{k}{p}u"])) < 1e-6
    s.close()


@pytest.mark.parametrize("name", ["lqr_prob.json", "lqr_prob_256.json"])
def test_full_solve_vs_golden_and_oracle(name):
    # BASELINE configs[0] and configs[1]
    n, m, N, knots, x0, d = problem_from_json(name)
    want = np.array(d["soln"], dtype=np.float64).reshape(-1)
    nvars = want.size
    osoln, ofact = O.port_solve(n, m, N, knots, x0)
    okkt = O.kkt_residual(n, m, N, knots, x0, osoln)
    for mode, z, s in _solve_all_modes(n, m, N, knots, x0):
        assert np.linalg.norm(z[0][:nvars] - want) < 1e-6                       # the reference's own test bar
        assert max(rel_err_classes(z[0], osoln, n, m, N)) < TOL, mode
        assert O.kkt_residual(n, m, N, knots, x0, z[0]) <= 2 * okkt + 1e-13 * np.linalg.norm(want)
        if mode == R.MODE_STAGED:
            assert np.allclose(s.fact(), ofact, rtol=1e-9, atol=1e-11)
            if name == "lqr_prob.json":
                sp = golden_json("sample_problem.json")
                f = s.fact()
                for k in range(N):
                    L, X, U = blocks(f[2, k], n, m)
                    assert np.linalg.norm(L - jmat(sp[f"F{k}2y"])) < 1e-9
                    assert np.linalg.norm(X - jmat(sp[f"F{k}2x"])) < 1e-9
                    assert np.linalg.norm(U - jmat(sp[f"F{k}2u"])) < 1e-9


def test_vs_stored_reference_outputs():
    g = np.load(os.path.join(GOLDEN, "ref_synth.npz"))
    for name, n, m, N, ids in [("quad_n12_m4_N256", 12, 4, 256, [0, 1, 4095]), ("quad_n12_m4_N32", 12, 4, 32, [0, 7]),
                               ("gen_n6_m3_N64", 6, 3, 64, [3]), ("gen_n16_m5_N16", 16, 5, 16, [2]),
                               ("gen_n64_m16_N8", 64, 16, 8, [0])]:
        kn = np.concatenate([R.synth_problems(n, m, N, 2024, pid, 1)[0] for pid in ids])
        x0 = np.concatenate([R.synth_problems(n, m, N, 2024, pid, 1)[1] for pid in ids])
        for mode, z, s in _solve_all_modes(n, m, N, kn, x0):
            for j, pid in enumerate(ids):
                key = f"{name}_id{pid}"
                assert max(rel_err_classes(z[j], g[key + "_soln"], n, m, N)) < TOL, (key, mode)
                kkt = O.kkt_residual(n, m, N, kn[j], x0[j], z[j])
                assert kkt <= 2 * float(g[key + "_kkt"]) + 1e-13 * np.linalg.norm(z[j]), (key, mode, kkt)
                if mode == R.MODE_STAGED:
                    f = s.fact(j)
                    assert np.allclose(f[-1], g[key + "_fact_top"], rtol=1e-8, atol=1e-11)
                    assert np.allclose(f[0, 1], g[key + "_fact_l0_k1"], rtol=1e-8, atol=1e-11)


@pytest.mark.skipif(O.REF is None, reason="oracle/_ref not built")
def test_vs_compiled_reference_live():
    # the unmodified reference, run on this host, on fresh random shapes
    for (n, m, N, seed) in [(12, 4, 128, 5), (6, 3, 32, 6), (3, 1, 4, 7), (20, 7, 16, 8)]:
        kn, x0 = R.synth_problems(n, m, N, seed, 0, 2)
        for mode, z, s in _solve_all_modes(n, m, N, kn, x0):
            for p in range(2):
                rsoln, rfact, _ = O.ref_solve(n, m, N, kn[p], x0[p], threads=2)
                assert max(rel_err_classes(z[p], rsoln, n, m, N)) < TOL, (n, m, N, mode)
                if mode == R.MODE_STAGED:
                    assert np.allclose(s.fact(p), rfact, rtol=1e-8, atol=1e-11)


@pytest.mark.parametrize("shape", [(12, 4, 2), (1, 1, 2), (1, 1, 8), (2, 3, 4), (7, 2, 64), (12, 4, 1024), (33, 9, 8), (64, 16, 16), (16, 16, 32), (64, 16, 128)])
def test_edge_shapes(shape):
    # minimum horizon, scalar systems, m > n, non-multiple-of-warp sizes, long horizon, large state
    n, m, N = shape
    kn, x0 = R.synth_problems(n, m, N, 11, 0, 3)
    for mode, z, s in _solve_all_modes(n, m, N, kn, x0):
        for p in range(3):
            want, _ = O.port_solve(n, m, N, kn[p], x0[p], threads=2)
            assert max(rel_err_classes(z[p], want, n, m, N)) < TOL, (shape, mode, p)


@pytest.mark.parametrize("g0", [8, 16])
@pytest.mark.parametrize("dfma", [0, 1])
@pytest.mark.parametrize("shape", [(12, 4, 8), (12, 4, 16), (12, 4, 32), (12, 4, 128), (12, 4, 256), (12, 4, 4096),
                                   (6, 3, 8), (6, 3, 64), (6, 3, 512)])
def test_tiered_path_variants(shape, g0, dfma, monkeypatch):
    # the tiered (fused) path for both group sizes and both Schur-update implementations (DMMA / DFMA),
    # from one group (N == G) to four tiers (N = 4096), against the oracle
    n, m, N = shape
    monkeypatch.setenv("RSLQR_G0", str(g0))
    monkeypatch.setenv("RSLQR_FORCE_DFMA", str(dfma))
    B = 5
    kn, x0 = R.synth_problems(n, m, N, 77, 0, B)
    s = R.CudaSolver(n, m, N, batch=B, flags=0)
    s.set_mode(R.MODE_FUSED)
    s.upload(kn, x0)
    s.solve()
    z = s.solution()
    assert not s.flags().any()
    for p in range(B):
        want, _ = O.port_solve(n, m, N, kn[p], x0[p], threads=2)
        assert max(rel_err_classes(z[p], want, n, m, N)) < TOL, (shape, g0, dfma, p)
    s.close()


@pytest.mark.parametrize("N", [4, 8, 16, 32, 64, 128, 256, 512, 4096])
def test_scan_path_all_horizons(N):
    # the scan path (one warp per 4-knot subtree, boundary entries at every level) from a single subtree (N = 4)
    # to four tiers above the leaves, against the oracle; batch 7 also exercises partially filled CTAs
    n, m, B = 12, 4, 7
    kn, x0 = R.synth_problems(n, m, N, 78, 0, B)
    s = R.CudaSolver(n, m, N, batch=B, flags=0)
    assert s.mode == R.MODE_SCAN        # the default for the covered shape
    s.upload(kn, x0)
    s.solve()
    z = s.solution()
    assert not s.flags().any()
    for p in range(B):
        want, _ = O.port_solve(n, m, N, kn[p], x0[p], threads=2)
        assert max(rel_err_classes(z[p], want, n, m, N)) < TOL, (N, p)
        assert O.kkt_residual(n, m, N, kn[p], x0[p], z[p]) < 1e-10 * np.linalg.norm(z[p])
    s.close()


@pytest.mark.parametrize("N", [4, 8, 64, 256, 2048])
@pytest.mark.parametrize("shape", [(6, 3), (12, 6), (8, 4), (16, 8), (4, 2), (12, 4)])
def test_dual_form_scan_path_shapes(shape, N, monkeypatch):
    # the scan path in dual form (kernels_dual.cuh: a subtree = its Schur complement onto the two multipliers bounding
    # it) for every covered shape, against the oracle; n = 12, m = 4 is routed through it with RSLQR_DUAL=1
    n, m = shape
    B = 5
    monkeypatch.setenv("RSLQR_DUAL", "1")
    kn, x0 = R.synth_problems(n, m, N, 41, 0, B)
    s = R.CudaSolver(n, m, N, batch=B, flags=0)
    assert s.mode == R.MODE_SCAN
    s.upload(kn, x0)
    s.solve()
    z = s.solution().copy()
    assert not s.flags().any()
    s.solve()
    assert np.array_equal(s.solution(), z)          # re-solve on the same handle: bit-identical
    for p in range(B):
        want, _ = O.port_solve(n, m, N, kn[p], x0[p], threads=2)
        assert max(rel_err_classes(z[p], want, n, m, N)) < TOL, (shape, N, p)
        assert O.kkt_residual(n, m, N, kn[p], x0[p], z[p]) < 1e-10 * max(1.0, np.linalg.norm(z[p]))
    # a caller-supplied right-hand side (pack_rhs = 0) takes the same path
    rhs = np.zeros((B, N, 2 * n + m))
    for p in range(B):
        ps = O.PortSolver(n, m, N)
        ps.init(kn[p], x0[p])
        rhs[p] = ps.soln
        ps.close()
    s.upload_rhs(rhs)
    s.solve(pack_rhs=False)
    assert np.linalg.norm(s.solution() - z) <= 1e-12 * np.linalg.norm(z)
    # a non-positive Q entry raises the status word of that leaf factor
    bad = kn.copy()
    bad[2, N // 2, 1] = -1.0
    s.upload(bad, x0)
    s.solve()
    assert list(s.flags()) == [0, 0, 1, 0, 0] and s.status(2)[2 * (N // 2)] == -1
    s.close()


@pytest.mark.parametrize("shape", [(6, 2), (8, 2), (8, 3), (10, 4), (10, 5), (12, 2), (12, 3), (12, 8), (14, 4), (14, 6), (16, 4), (16, 6),
                                   (3, 1), (5, 2), (7, 3), (9, 3), (11, 4), (13, 4), (15, 5),
                                   (2, 1), (4, 1), (4, 4), (6, 6), (7, 7), (8, 8), (14, 7), (16, 16)])
def test_dual_form_scan_path_more_shapes(shape):
    # the further (n <= 16, m <= 16; odd n too) instances of the dual-form kernels (RSLQR_DUAL_SHAPES, kernels_dual.cuh): default
    # mode is the scan path; two horizons (one tier / several tiers) against the oracle, plus the no-allocation entry
    # point's shape list (it dispatches through the same table)
    n, m = shape
    for N in (8, 512):
        B = 3
        kn, x0 = R.synth_problems(n, m, N, 43, 0, B)
        s = R.CudaSolver(n, m, N, batch=B, flags=0)
        assert s.mode == R.MODE_SCAN
        s.upload(kn, x0)
        s.solve()
        z = s.solution().copy()
        assert not s.flags().any()
        s.close()
        for p in range(B):
            want, _ = O.port_solve(n, m, N, kn[p], x0[p], threads=2)
            assert max(rel_err_classes(z[p], want, n, m, N)) < TOL, (shape, N, p)
            assert O.kkt_residual(n, m, N, kn[p], x0[p], z[p]) < 1e-10 * max(1.0, np.linalg.norm(z[p]))


@pytest.mark.parametrize("shape", [(9, 4), (11, 3), (17, 1), (20, 7), (32, 8), (33, 9), (40, 10), (48, 16), (64, 8), (5, 16),
                                   (24, 20), (40, 17), (64, 32)])
@pytest.mark.parametrize("N", [2, 16, 128])
def test_large_block_path_pads_any_shape(shape, N):
    # every (n <= 64, m <= 32) without an instance of the warp kernels runs on the large-block dual path (kernels_big.cuh)
    # with n padded on chip to blocks of 32 or 64: default mode is the scan mode; against the oracle; re-solve bit-identical;
    # a caller-supplied right-hand side; a non-positive Q entry raises that problem's flag and the leaf's status word
    n, m = shape
    B = 3
    kn, x0 = R.synth_problems(n, m, N, 47, 0, B)
    s = R.CudaSolver(n, m, N, batch=B, flags=0)
    assert s.mode == R.MODE_SCAN
    s.upload(kn, x0)
    s.solve()
    z = s.solution().copy()
    assert not s.flags().any() and np.isfinite(z).all()
    s.solve()
    assert np.array_equal(s.solution(), z)
    for p in range(B):
        want, _ = O.port_solve(n, m, N, kn[p], x0[p], threads=2)
        assert max(rel_err_classes(z[p], want, n, m, N)) < TOL, (shape, N, p)
        assert O.kkt_residual(n, m, N, kn[p], x0[p], z[p]) < 1e-10 * max(1.0, np.linalg.norm(z[p]))
    rhs = np.zeros((B, N, 2 * n + m))
    for p in range(B):
        ps = O.PortSolver(n, m, N)
        ps.init(kn[p], x0[p])
        rhs[p] = ps.soln
        ps.close()
    s.upload_rhs(rhs)
    s.solve(pack_rhs=False)
    assert np.linalg.norm(s.solution() - z) <= 1e-12 * np.linalg.norm(z)
    bad = kn.copy()
    bad[1, N // 2, 0] = -1.0
    s.upload(bad, x0)
    s.solve()
    assert list(s.flags()) == [0, 1, 0] and s.status(1)[2 * (N // 2)] == -1
    s.close()


@pytest.mark.parametrize("shape", [(20, 7, 64), (33, 9, 32), (9, 4, 128), (13, 4, 256)])
def test_host_api_takes_the_fast_paths_on_any_shape(shape):
    # ndlqr_NewNdLqrSolver + Initialize + Solve on shapes outside the reference's fixtures: the default solve runs the
    # scan mode (large-block path with on-chip padding, or a warp-kernel instance), solution layout as src/solve.h:47-53
    n, m, N = shape
    pp = R.LIB.rslqr_synth_problem(n, m, N, 77, 3)
    s = R.NdLqrSolver(n, m, N)
    assert s.initialize(pp) == 0 and s.solve() == 0
    assert R.LIB.rslqr_cuda_get_mode(s.c.device) == R.MODE_SCAN
    z = s.solution().copy()
    kn, x0 = R.synth_problems(n, m, N, 77, 3, 1)
    want, _ = O.port_solve(n, m, N, kn[0], x0[0])
    nv = (2 * n + m) * N - m
    assert np.linalg.norm(z[:nv] - want[:nv]) <= TOL * np.linalg.norm(want[:nv])


def test_shipped_n6_problem_runs_in_scan_mode():
    # lqr_prob_256.json (n=6, m=3: the only shape the reference ships) takes the scan path by default
    n, m, N, kn, x0 = R.load_problem_json(os.path.join(GOLDEN, "lqr_prob_256.json"))
    want = np.array(golden_json("lqr_prob_256.json")["soln"]).reshape(-1)
    s = R.CudaSolver(n, m, N, batch=1, flags=0)
    assert s.mode == R.MODE_SCAN
    s.upload(kn[None], x0[None])
    s.solve()
    z = s.solution()[0]
    assert np.linalg.norm(z[:want.size] - want) < 1e-6
    ref, _ = O.port_solve(n, m, N, kn, x0)
    assert max(rel_err_classes(z, ref, n, m, N)) < TOL
    s.close()


def test_latency_mode_dependent_launch_is_bit_identical(monkeypatch):
    # batch < 64: kernels are chained by programmatic dependent launch (pdl_sync / launch_k, kernels_staged.cuh). A kernel
    # that touched global memory before its griddepcontrol.wait would race with its predecessor: 30 back-to-back solves
    # with the chain on must reproduce the solve with the chain off bit for bit (RSLQR_PDL is read when a handle is made).
    for (n, m, N) in ((12, 4, 2048), (6, 3, 256), (12, 6, 64)):
        kn, x0 = R.synth_problems(n, m, N, 21, 0, 2)
        monkeypatch.setenv("RSLQR_PDL", "0")
        s0 = R.CudaSolver(n, m, N, batch=2, flags=0)
        s0.upload(kn, x0)
        s0.solve()
        want = s0.solution().copy()
        s0.close()
        monkeypatch.setenv("RSLQR_PDL", "1")
        s1 = R.CudaSolver(n, m, N, batch=2, flags=0)
        s1.upload(kn, x0)
        for _ in range(30):
            s1.solve_async(True)
        s1.sync()
        assert np.array_equal(s1.solution(), want)
        assert not s1.flags().any()
        s1.close()


def test_scan_path_is_bit_reproducible():
    # warp-synchronous shared-memory code with TMA copies: a data race would show up as run-to-run differences.
    # 40 repeated solves of the same resident batch (more warps than one wave of the GPU) must be bit-identical,
    # and so must the same problems solved alone (different CTA / warp placement).
    n, m, N, B = 12, 4, 64, 1500
    kn, x0 = R.synth_problems(n, m, N, 11, 0, B)
    s = R.CudaSolver(n, m, N, batch=B, flags=0)
    s.upload(kn, x0)
    s.solve()
    first = s.solution().copy()
    for _ in range(40):
        s.solve()
        assert np.array_equal(s.solution(), first)
    s.close()
    one = R.CudaSolver(n, m, N, batch=1, flags=0)
    for p in (0, 777, 1499):
        one.upload(kn[p:p + 1], x0[p:p + 1])
        one.solve()
        assert np.array_equal(one.solution()[0], first[p])
    one.close()


@pytest.mark.parametrize("mode", [R.MODE_FUSED, R.MODE_SCAN])
def test_tiered_path_reports_cholesky_failure(mode):
    n, m, N = 12, 4, 64
    kn, x0 = R.synth_problems(n, m, N, 1, 0, 3)
    kn[1, 3, 2] = -1.0      # Q[2] of knot 3 of problem 1: leaf factorization fails
    s = R.CudaSolver(n, m, N, batch=3, flags=0)
    s.set_mode(mode)
    s.upload(kn, x0)
    s.solve()
    fl = s.flags()
    assert fl[0] == 0 and fl[1] == 1
    assert s.status(1)[2 * 3] == -1 and s.status(0).sum() == 0
    s.close()


def test_batch_quadrotor_properties():
    # BASELINE configs[2] shape (n=12,m=4,N=256) at a batch the CPU checks in seconds: sample vs oracle,
    # KKT residual of EVERY problem, and independence of problems from their batch position.
    n, m, N, B = 12, 4, 256, 512
    kn, x0 = R.synth_problems(n, m, N, 2024, 0, B)
    for mode, z, s in _solve_all_modes(n, m, N, kn, x0):
        for p in (0, 1, 255, 511):
            want, _ = O.port_solve(n, m, N, kn[p], x0[p], threads=2)
            assert max(rel_err_classes(z[p], want, n, m, N)) < TOL, (mode, p)
        res = np.array([O.kkt_residual(n, m, N, kn[p], x0[p], z[p]) for p in range(B)])
        assert res.max() < 1e-10 * np.linalg.norm(z, axis=1).max(), res.max()
        # same problem at another batch slot gives the bit-identical answer
        s2 = R.CudaSolver(n, m, N, batch=2, flags=R.KEEP_FACT)
        s2.set_mode(mode)
        s2.upload(kn[[300, 7]], x0[[300, 7]])
        s2.solve()
        z2 = s2.solution()
        assert np.array_equal(z2[0], z[300]) and np.array_equal(z2[1], z[7])
        s2.close()


def test_linearity_in_rhs():
    # size-independent property: the solve is linear in the right-hand side (fixed K)
    n, m, N = 12, 4, 256
    kn, x0 = R.synth_problems(n, m, N, 3, 0, 1)
    rng = np.random.default_rng(0)
    s = R.CudaSolver(n, m, N, batch=3, flags=R.KEEP_FACT)
    for mode in _modes(s):
        s.set_mode(mode)
        s.reset()
        s.upload(np.repeat(kn, 3, axis=0), np.repeat(x0, 3, axis=0))
        b1 = rng.standard_normal(N * s.zs); b2 = rng.standard_normal(N * s.zs)
        b1[-m:] = 0; b2[-m:] = 0
        s.upload_rhs(np.stack([b1, b2, 2.0 * b1 - 0.5 * b2]))
        s.solve(pack_rhs=False)
        z = s.solution()
        comb = 2.0 * z[0] - 0.5 * z[1]
        assert np.linalg.norm(z[2] - comb) / np.linalg.norm(comb) < 1e-11, mode
    s.close()


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(12, 4, 256), (6, 3, 8), (64, 16, 32)])
def test_factor_once_solve_many(shape):
    # the stored staged factorization applied to new right-hand sides must give exactly what a full
    # solve of the same right-hand side gives (same rhs-phase kernels, same leaf scaling arithmetic)
    n, m, N = shape
    kn, x0 = R.synth_problems(n, m, N, 11, 0, 2)
    rng = np.random.default_rng(5)
    s = R.CudaSolver(n, m, N, batch=2, flags=R.KEEP_FACT)
    s.set_mode(R.MODE_STAGED)
    s.upload(kn, x0)
    with pytest.raises(R.RslqrError):
        s.solve_rhs()                      # nothing factored yet
    # (1) factor, then the packed rhs of the problems themselves == the full solve
    s.factor()
    s.solve_rhs(pack_rhs=True)
    z_split = s.solution().copy()
    s.reset()
    s.solve(pack_rhs=True)
    z_full = s.solution().copy()
    assert np.array_equal(z_split, z_full)
    want, _ = O.port_solve(n, m, N, kn[0], x0[0], threads=2)
    assert max(rel_err_classes(z_full[0], want, n, m, N)) < TOL
    # (2) the factorization left behind by a full solve serves any number of new right-hand sides
    fact_before = s.fact(0).copy()
    for _ in range(3):
        b = rng.standard_normal((2, N * s.zs))
        b[:, -m:] = 0
        s.upload_rhs(b)
        s.solve_rhs()
        z_many = s.solution().copy()
        t = R.CudaSolver(n, m, N, batch=2, flags=R.KEEP_FACT)
        t.set_mode(R.MODE_STAGED)
        t.upload(kn, x0)
        t.upload_rhs(b)
        t.solve(pack_rhs=False)
        assert np.array_equal(z_many, t.solution())
        t.close()
    assert np.array_equal(fact_before, s.fact(0))   # the rhs phase does not touch `fact`
    # (3) reset drops the factorization; the scan / tiered kernels keep none
    s.reset()
    with pytest.raises(R.RslqrError):
        s.solve_rhs()
    if (n, m) == (12, 4):
        s.set_mode(R.MODE_SCAN)
        with pytest.raises(R.RslqrError):
            s.factor()
    s.close()


@pytest.mark.gpu
def test_host_api_solve_rhs_only():
    # ndlqr_Solve, then new right-hand sides through ndlqr_SolveRhsOnly (include/solve.h) vs fresh full solves
    LIB = R.LIB
    pp = LIB.ndlqr_ReadLQRProblemJSONFile(os.path.join(GOLDEN, "lqr_prob_256.json").encode())
    s = R.NdLqrSolver(6, 3, 256)
    t = R.NdLqrSolver(6, 3, 256)
    t.c.mirror_fact = 1                                 # fresh full solves on the staged kernels (bitwise comparison)
    assert LIB.ndlqr_SolveRhsOnly(s.p) == -1            # nothing initialized yet
    assert s.initialize(pp) == 0
    rhs0 = s.soln_array().copy()
    assert s.solve() == 0
    z0 = s.solution().copy()
    rng = np.random.default_rng(9)
    for scale in (1.0, -0.5, 3.0):
        rhs = rhs0 * scale + 0.01 * rng.standard_normal(rhs0.shape)
        rhs[-3:] = 0
        s.soln_array()[:] = rhs
        assert LIB.ndlqr_SolveRhsOnly(s.p) == 0
        t.reset()
        assert t.initialize(pp) == 0
        t.soln_array()[:] = rhs
        assert t.solve() == 0
        assert np.array_equal(s.solution(), t.solution())
    # the original rhs again: agrees with the first solve, which took the default (on-chip) path -- different kernels,
    # same answer to rounding
    s.soln_array()[:] = rhs0
    assert LIB.ndlqr_SolveRhsOnly(s.p) == 0
    assert np.linalg.norm(s.solution() - z0) <= TOL * np.linalg.norm(z0)
    LIB.ndlqr_FreeLQRProblem(pp)


@pytest.mark.parametrize("N", [2, 4, 16, 256, 1024])
def test_riccati_comparison_solver_matches_nested_dissection(N):
    # the reference's own baseline recursion (src/riccati_solve.c) on the device solves the same KKT
    # system; its [lambda, x, u] vector must agree with the oracle's nested-dissection solve
    n, m = 12, 4
    kn, x0 = R.synth_problems(n, m, N, 21, 0, 5)
    s = R.CudaSolver(n, m, N, batch=5, flags=R.KEEP_FACT)
    s.upload(kn, x0)
    s.riccati()
    z = s.solution().copy()
    assert not s.flags().any()
    for p in range(5):
        want, _ = O.port_solve(n, m, N, kn[p], x0[p], threads=2)
        assert max(rel_err_classes(z[p], want, n, m, N)) < 1e-9, (N, p)
        assert O.kkt_residual(n, m, N, kn[p], x0[p], z[p]) < 1e-9 * max(1.0, np.linalg.norm(z[p]))
    s.solve()
    assert max(rel_err_classes(z[0], s.solution()[0], n, m, N)) < 1e-9
    s.close()


def test_cholesky_failure_is_reported():
    # a non-positive Q entry must raise the status word (reference stores cholinfo->success, src/linalg.c:84)
    n, m, N = 6, 3, 8
    kn, x0 = R.synth_problems(n, m, N, 1, 0, 2)
    kn[1, 3, 2] = -1.0      # Q[2] of knot 3 of problem 1
    for mode, flags_expected in ((R.MODE_STAGED, [0, 1]),):
        s = R.CudaSolver(n, m, N, batch=2, flags=R.KEEP_FACT)
        s.set_mode(mode)
        s.upload(kn, x0)
        s.solve()
        assert list(s.flags()) == flags_expected
        st = s.status(1)
        assert st[2 * 3] == -1 and s.status(0).sum() == 0
        s.close()


# ---------------------------------------------------------------------------------------------
# the reference-named C host API (drop-in flow of examples/installexample/main.c)
# ---------------------------------------------------------------------------------------------
def test_host_api_solve_flow_and_solve_twice():
    # test/nested_dissection_test.c:239-313 (RunSolve, SolveTwice) through ndlqr_* entry points
    LIB = R.LIB
    pp = LIB.ndlqr_ReadLQRProblemJSONFile(os.path.join(GOLDEN, "lqr_prob.json").encode())
    want = np.array(golden_json("lqr_prob.json")["soln"]).reshape(-1)
    s = R.NdLqrSolver(6, 3, 8)
    assert s.c.device
    assert s.initialize(pp) == 0
    assert s.solve() == 0
    z = s.solution()
    assert z.size == 117 and np.linalg.norm(z - want) < 1e-6
    out = np.zeros(117)
    assert LIB.ndlqr_CopySolution(s.p, out.ctypes.data_as(C.POINTER(C.c_double))) == 117
    assert np.array_equal(out, z)
    # the default solve moves the solution only: the factor mirror is untouched until it is asked for
    assert not s.c.fact_dirty and not s.nddata_array("fact").any()
    assert s.sync_fact() == 0
    # fact mirror holds the reference's F{k}2* blocks
    sp = golden_json("sample_problem.json")
    f = s.nddata_array("fact")
    for k in range(8):
        L_, X, U = blocks(f[2, k], 6, 3)
        assert np.linalg.norm(X - jmat(sp[f"F{k}2x"])) < 1e-9
    # diagonals hold the Cholesky factors sqrt(Q) (mirrored together with fact)
    Q0 = np.ctypeslib.as_array(s.c.diagonals[0].data, shape=(36,)).reshape(6, 6)
    assert np.allclose(np.diag(Q0), np.sqrt(golden_json("lqr_prob.json")["lqrdata"][0]["Q"]))
    assert s.c.cholfacts.contents.cholinfo[0].success == 0
    assert s.c.solve_time_ms > 0
    # solve twice: reset + initialize + solve reproduces the answer bit for bit
    s.reset()
    assert s.initialize(pp) == 0 and s.solve() == 0
    assert np.array_equal(s.solution(), z)
    # solver->mirror_fact = 1: every solve runs the reference's phases one launch each and leaves `fact` and the
    # per-phase profile behind (reference src/solve.c:15-25)
    s.c.mirror_fact = 1
    s.reset()
    assert s.initialize(pp) == 0 and s.solve() == 0
    assert np.linalg.norm(s.solution() - want) < 1e-6 and np.linalg.norm(s.solution() - z) / np.linalg.norm(z) < TOL
    prof = LIB.ndlqr_GetProfile(s.p)
    assert prof.t_shur_ms > 0 and prof.t_leaves_ms > 0
    f2 = s.nddata_array("fact").copy()
    assert np.array_equal(f2, f)
    s.c.mirror_fact = 0
    # the convenience entry named by the task statement
    assert LIB.ndlqr_SolveLQRProblem(pp, s.p) == 0
    assert np.array_equal(s.solution(), z)
    LIB.ndlqr_FreeLQRProblem(pp)


@pytest.mark.parametrize("mirror", [0, 1])
def test_host_api_reference_solve_twice_sequences(mirror):
    # the reference's documented re-solve flow (src/solve.h:28-32, test/nested_dissection_test.c:289-313 SolveTwice,
    # test/sample_problem_test.c:144): ndlqr_ResetNdData(solver->fact); Initialize; Solve -- and a plain
    # Initialize + Solve repeated without any reset. Both on the default path and on the staged path (mirror_fact = 1),
    # whose kernels accumulate into the device factor array and therefore need it cleared between solves.
    LIB = R.LIB
    pp = LIB.ndlqr_ReadLQRProblemJSONFile(os.path.join(GOLDEN, "lqr_prob.json").encode())
    want = np.array(golden_json("lqr_prob.json")["soln"]).reshape(-1)
    s = R.NdLqrSolver(6, 3, 8)
    s.c.mirror_fact = mirror
    assert s.initialize(pp) == 0 and s.solve() == 0
    z = s.solution()
    assert np.linalg.norm(z - want) < 1e-6
    f = s.nddata_array("fact").copy()
    LIB.ndlqr_ResetNdData(s.c.fact)                     # reference SolveTwice
    assert not s.nddata_array("fact").any()
    assert s.initialize(pp) == 0 and s.solve() == 0
    assert np.array_equal(s.solution(), z)
    assert np.array_equal(s.nddata_array("fact"), f)
    for _ in range(2):                                  # no reset at all
        assert s.initialize(pp) == 0 and s.solve() == 0
        assert np.array_equal(s.solution(), z)
    if mirror:
        assert f.any() and np.array_equal(s.nddata_array("fact"), f)
    # a failing solve must not stick: break Q, solve (status raised), restore, solve again (status clear)
    q00 = pp.contents.lqrdata[2].contents.Q[1]
    pp.contents.lqrdata[2].contents.Q[1] = -1.0
    assert s.initialize(pp) == 0 and s.solve() == 0
    assert s.c.cholfacts.contents.cholinfo[2 * 2].success != 0
    pp.contents.lqrdata[2].contents.Q[1] = q00
    assert s.initialize(pp) == 0 and s.solve() == 0
    assert s.c.cholfacts.contents.cholinfo[2 * 2].success == 0
    assert np.array_equal(s.solution(), z)
    LIB.ndlqr_FreeLQRProblem(pp)


def test_cuda_abi_resolve_without_reset_and_sticky_status():
    # C-ABI: every solve starts from cleared status words / flags and (staged) a zeroed factor array
    n, m, N = 6, 3, 16
    kn, x0 = R.synth_problems(n, m, N, 5, 0, 3)
    bad = kn.copy()
    bad[1, 4, 0] = -2.0
    for mode in (R.MODE_STAGED, R.MODE_AUTO):
        s = R.CudaSolver(n, m, N, batch=3, flags=0)          # no KEEP_FACT: the factor array appears on first need
        s.set_mode(mode)
        s.upload(kn, x0)
        s.solve()
        z = s.solution().copy()
        s.solve()
        assert np.array_equal(s.solution(), z), mode
        s.upload(bad, x0)
        s.solve()
        assert list(s.flags()) == [0, 1, 0]
        s.upload(kn, x0)
        s.solve()
        assert not s.flags().any() and not s.status(1).any()
        assert np.array_equal(s.solution(), z), mode
        s.close()


def test_host_api_stage_calls():
    # test/nested_dissection_test.c:11-237 with the reference's stage signatures
    LIB = R.LIB
    n, m, N = 6, 3, 8
    pp = LIB.ndlqr_ReadLQRProblemJSONFile(os.path.join(GOLDEN, "lqr_prob.json").encode())
    sp = golden_json("sample_problem.json")
    s = R.NdLqrSolver(n, m, N)
    s.initialize(pp)
    assert LIB.ndlqr_SolveLeaves(s.p) == 0
    assert np.linalg.norm(s.soln_array() - np.array(sp["b"]).reshape(-1)) < 1e-6
    c = s.c
    for leaf in range(4):
        idx = 2 * leaf
        for p in range(3):
            assert LIB.ndlqr_FactorInnerProduct(c.data, c.fact, idx, 0, p) == 0
    for leaf in range(4):
        assert LIB.ndlqr_SolverFactorCholesky(s.p, leaf, 0) == 0
        for p in (1, 2):
            assert LIB.ndlqr_SolveCholeskyFactor(c.fact, None, 2 * leaf, 0, p) == 0
    for p in (1, 2):
        assert LIB.ndlqr_ComputeShurCompliment(s.p, 0, 0, p) == 0
    f = s.nddata_array("fact")
    for k in (0, 1):
        for p in (1, 2):
            L_, X, U = blocks(f[p, k], n, m)
            assert np.linalg.norm(L_ - jmat(sp[f"E{k}{p}y"])) < 1e-6
            assert np.linalg.norm(X - jmat(sp[f"E{k}{p}x"])) < 1e-6
            assert np.linalg.norm(U - jmat(sp[f"E{k}{p}u"])) < 1e-6
    LIB.ndlqr_FreeLQRProblem(pp)


def test_batch_host_api():
    LIB = R.LIB
    LIB.ndlqr_NewBatchSolver.restype = C.c_void_p
    LIB.ndlqr_NewBatchSolver.argtypes = [C.c_int] * 5
    for fn in ("ndlqr_FreeBatchSolver", "ndlqr_BatchSolve", "ndlqr_BatchGetNumVars"):
        getattr(LIB, fn).argtypes = [C.c_void_p]
    LIB.ndlqr_BatchInitialize.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
    LIB.ndlqr_BatchCopySolution.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
    n, m, N, B = 12, 4, 64, 5
    probs = [LIB.rslqr_synth_problem(n, m, N, 99, i) for i in range(B)]
    arr = (C.c_void_p * B)(*[C.cast(p, C.c_void_p) for p in probs])
    bs = LIB.ndlqr_NewBatchSolver(n, m, N, B, 0)
    assert bs
    assert LIB.ndlqr_BatchInitialize(bs, arr, B) == 0
    assert LIB.ndlqr_BatchSolve(bs) == 0
    nv = LIB.ndlqr_BatchGetNumVars(bs)
    assert nv == 28 * N - 4
    kn, x0 = R.synth_problems(n, m, N, 99, 0, B)
    for p in range(B):
        z = np.zeros(nv)
        assert LIB.ndlqr_BatchCopySolution(bs, p, z.ctypes.data_as(C.c_void_p)) == nv
        want, _ = O.port_solve(n, m, N, kn[p], x0[p])
        assert np.linalg.norm(z - want[:nv]) / np.linalg.norm(want) < TOL
    LIB.ndlqr_FreeBatchSolver(bs)
    for p in probs:
        LIB.ndlqr_FreeLQRProblem(p)


def test_c_consumer_flow_runs(tmp_path):
    # the compiled C consumer (tests/c/consumer_flow.c) solves the shipped problems on the GPU and matches "soln"
    import subprocess
    from test_host_api import _compile_consumer
    exe = _compile_consumer(os.path.join(os.path.dirname(GOLDEN), "c", "consumer_flow.c"), str(tmp_path / "consumer"))
    for name in ("lqr_prob.json", "lqr_prob_256.json"):
        r = subprocess.run([exe, os.path.join(GOLDEN, name)], capture_output=True, text=True)
        assert r.returncode == 0, (name, r.returncode, r.stdout, r.stderr)


@pytest.mark.parametrize("mode", [R.MODE_SCAN, R.MODE_FUSED, R.MODE_STAGED])
def test_solve_host_pipeline_matches_resident_solve(mode):
    # the chunked host-to-host call (copies overlapped with kernels) returns bit-identical solutions
    n, m, N, B = 12, 4, 64, 37
    kn, x0 = R.synth_problems(n, m, N, 5, 0, B)
    s = R.CudaSolver(n, m, N, batch=B, flags=R.KEEP_FACT)
    s.set_mode(mode)
    s.upload(kn, x0)
    s.reset()
    s.solve()
    want = s.solution().copy()
    for chunk in (0, 1, 8, 37, 100):
        out = np.zeros((B, N * s.zs))
        s.upload(np.zeros_like(kn), np.zeros_like(x0))   # make sure the call really uploads
        s.solve_host(kn, x0, out, chunk=chunk)
        s.sync()
        assert np.array_equal(out, want), (mode, chunk)
    # a partial count leaves the rest untouched
    out = np.full((B, N * s.zs), 7.0)
    s.solve_host(kn, x0, out, count=5, chunk=2)
    s.sync()
    assert np.array_equal(out[:5], want[:5]) and (out[5:] == 7.0).all()
    s.close()


@pytest.mark.parametrize("shape", [(64, 16, 16), (6, 3, 32), (16, 8, 16)])
def test_solve_host_chunks_on_the_dual_form_paths(shape):
    # chunks of the host-to-host call address the per-separator buffers of the dual-form paths (elements, records,
    # factors of the large-block path) at an offset: every chunking must reproduce the resident solve bit for bit, twice
    n, m, N = shape
    B = 7
    kn, x0 = R.synth_problems(n, m, N, 13, 0, B)
    s = R.CudaSolver(n, m, N, batch=B, flags=0)
    assert s.mode == R.MODE_SCAN
    s.upload(kn, x0)
    s.solve()
    want = s.solution().copy()
    ref, _ = O.port_solve(n, m, N, kn[B - 1], x0[B - 1])
    assert max(rel_err_classes(want[B - 1], ref, n, m, N)) < TOL
    for chunk in (3, 1, 0, 3):
        out = np.zeros((B, N * s.zs))
        s.upload(np.zeros_like(kn), np.zeros_like(x0))
        s.solve_host(kn, x0, out, chunk=chunk)
        s.sync()
        assert np.array_equal(out, want), (shape, chunk)
    s.close()


@pytest.mark.parametrize("shape,nranks", [((12, 4, 256), 2), ((12, 4, 256), 8), ((12, 4, 2048), 4), ((6, 3, 128), 8),
                                          ((12, 4, 64), 1)])
def test_long_horizon_shards_emulated_on_one_gpu(shape, nranks):
    # BASELINE configs[3] path: one problem split into `nranks` subtrees; ranks are emulated as shard handles on ONE
    # device (B200_PROFILING.md: never run waiting ranks as separate launches); the exchange is the peer-copy all-gather
    import torch
    n, m, N = shape
    kn, x0 = R.synth_problems(n, m, N, 9, 0, 1)
    want, _ = O.port_solve(n, m, N, kn[0], x0[0], threads=2)
    shards = [R.ShardSolver(n, m, N, r, nranks, device=0) for r in range(nranks)]
    pay = shards[0].payload_doubles
    d_pay = [torch.zeros(pay, dtype=torch.float64, device="cuda:0") for _ in range(nranks)]
    d_gat = [torch.zeros(pay * nranks, dtype=torch.float64, device="cuda:0") for _ in range(nranks)]
    for s in shards:
        s.upload_shard(kn[0], x0[0])
    for s, p in zip(shards, d_pay):
        s.reduce(p.data_ptr())
    R.shard_exchange(shards, [p.data_ptr() for p in d_pay], [g.data_ptr() for g in d_gat])
    for s, g in zip(shards, d_gat):
        s.finish(g.data_ptr())
    z = []
    for s in shards:
        s.sync()
        assert not s.flags().any()
        z.append(s.solution()[0])
    z = np.concatenate(z)
    assert max(rel_err_classes(z, want, n, m, N)) < TOL, (shape, nranks)
    # every rank computed the same cross-GPU records: the single-handle solve agrees to rounding as well
    one = R.CudaSolver(n, m, N, batch=1, flags=0)
    one.upload(kn, x0)
    one.solve()
    assert np.linalg.norm(one.solution()[0] - z) <= 1e-12 * np.linalg.norm(z)
    one.close()
    for s in shards:
        s.close()


@pytest.mark.parametrize("N,nranks", [(256, 2), (1024, 8)])
def test_long_horizon_peer_store_exchange_emulated(N, nranks):
    # the exchange without a collective: every shard stores its payload into every rank's gathered buffer and raises a
    # flag; the cross-GPU tier kernel waits on the flags. Ranks are handles on ONE device here, their "peer mappings"
    # plain device pointers; two epochs exercise the double buffering.
    import torch
    n, m = 12, 4
    kn, x0 = R.synth_problems(n, m, N, 21, 0, 1)
    want, _ = O.port_solve(n, m, N, kn[0], x0[0], threads=2)
    shards = [R.ShardSolver(n, m, N, r, nranks, device=0) for r in range(nranks)]
    pay = shards[0].payload_doubles
    bufs = [torch.zeros(2 * nranks * pay + nranks, dtype=torch.float64, device="cuda:0") for _ in range(nranks)]
    base = [b.data_ptr() for b in bufs]
    flag_off = 2 * nranks * pay * 8
    for s in shards:
        s.upload_shard(kn[0], x0[0])
    for epoch in (1, 2):
        half = (epoch & 1) * nranks * pay * 8
        for r, s in enumerate(shards):      # all pushes are queued before any finish: nothing can spin forever here
            own = base[r] + half + r * pay * 8
            s.reduce(own)
            s.push(own, [b + half for b in base], [b + flag_off for b in base], epoch)
        for r, s in enumerate(shards):
            s.finish_wait(base[r] + half, base[r] + flag_off, epoch)
        z = []
        for s in shards:
            s.sync()
            assert not s.flags().any()
            z.append(s.solution()[0])
        z = np.concatenate(z)
        assert max(rel_err_classes(z, want, n, m, N)) < TOL, (N, nranks, epoch)
    for b in bufs:   # every flag word carries the last epoch
        assert (b[-nranks:].view(torch.int64) == 2).all()
    for s in shards:
        s.close()


def test_long_horizon_16384_eight_shards_kkt():
    # BASELINE configs[3] at full size: n=12, m=4, N=16384 as 8 subtrees of 2048 knots (ranks emulated on one device).
    # Size-independent checks: KKT residual and agreement with the single-handle tiered solve.
    import torch
    n, m, N, G = 12, 4, 16384, 8
    kn, x0 = R.synth_problems(n, m, N, 2024, 0, 1)
    shards = [R.ShardSolver(n, m, N, r, G, device=0) for r in range(G)]
    pay = shards[0].payload_doubles
    d_pay = [torch.zeros(pay, dtype=torch.float64, device="cuda:0") for _ in range(G)]
    d_gat = [torch.zeros(pay * G, dtype=torch.float64, device="cuda:0") for _ in range(G)]
    for s in shards:
        s.upload_shard(kn[0], x0[0])
    for s, p in zip(shards, d_pay):
        s.reduce(p.data_ptr())
    R.shard_exchange(shards, [p.data_ptr() for p in d_pay], [g.data_ptr() for g in d_gat])
    for s, g in zip(shards, d_gat):
        s.finish(g.data_ptr())
    z = np.concatenate([(s.sync(), s.solution()[0])[1] for s in shards])
    res = O.kkt_residual(n, m, N, kn[0], x0[0], z)
    assert res < 1e-10 * np.linalg.norm(z), res
    one = R.CudaSolver(n, m, N, batch=1, flags=0)
    one.upload(kn, x0)
    one.solve()
    assert np.linalg.norm(one.solution()[0] - z) <= 1e-11 * np.linalg.norm(z)
    one.close()
    for s in shards:
        s.close()


# ---------------------------------------------------------------------------------------------
# reference parity AT SIZE for BASELINE configs[3] and configs[4]: outputs of the unmodified reference stored by
# tests/golden/make_golden_large.py (same spirit as test/sample_problem_test.c:150-157: solve vs stored solution)
# ---------------------------------------------------------------------------------------------
def _large_golden():
    return np.load(os.path.join(GOLDEN, "ref_synth_large.npz"))


def test_long_horizon_16384_vs_reference():
    # configs[3]: n=12, m=4, N=16384. (a) one handle, (b) the 8-shard split (ranks emulated on one device), both against
    # the reference's own ndlqr_Solve output: <= 1e-10 relative per variable class, KKT residual no worse than 2x its.
    import torch
    n, m, N, G = 12, 4, 16384, 8
    g = _large_golden()
    ref, ref_kkt = g["quad_n12_m4_N16384_id0_soln"], float(g["quad_n12_m4_N16384_id0_kkt"])
    kn, x0 = R.synth_problems(n, m, N, 2024, 0, 1)
    one = R.CudaSolver(n, m, N, batch=1, flags=0)
    assert one.mode == R.MODE_SCAN
    one.upload(kn, x0)
    one.solve()
    z1 = one.solution()[0].copy()
    assert not one.flags().any()
    one.close()
    shards = [R.ShardSolver(n, m, N, r, G, device=0) for r in range(G)]
    pay = shards[0].payload_doubles
    d_pay = [torch.zeros(pay, dtype=torch.float64, device="cuda:0") for _ in range(G)]
    d_gat = [torch.zeros(pay * G, dtype=torch.float64, device="cuda:0") for _ in range(G)]
    for s in shards:
        s.upload_shard(kn[0], x0[0])
    for s, p in zip(shards, d_pay):
        s.reduce(p.data_ptr())
    R.shard_exchange(shards, [p.data_ptr() for p in d_pay], [gg.data_ptr() for gg in d_gat])
    for s, gg in zip(shards, d_gat):
        s.finish(gg.data_ptr())
    z8 = np.concatenate([(s.sync(), s.solution()[0])[1] for s in shards])
    for s in shards:
        assert not s.flags().any()
        s.close()
    for what, z in (("single handle", z1), ("8 shards", z8)):
        errs = rel_err_classes(z, ref, n, m, N)
        assert max(errs) < TOL, (what, errs)
        kkt = O.kkt_residual(n, m, N, kn[0], x0[0], z)
        assert kkt <= 2 * ref_kkt + 1e-13 * np.linalg.norm(z), (what, kkt, ref_kkt)


def test_full_baseline_batch_4096_properties():
    # BASELINE configs[2] at its own size (the bench workload: n=12, m=4, N=256, seed 2024, problems 0..4095): no failure
    # flag; samples against the oracle; KKT residual of a spread of problems no worse than 2x the oracle's; and a
    # size-independent property: a problem solved inside the full batch is bit-identical to the same problem solved in a
    # batch of 8 (other CTA / warp placement, other tier arity)
    n, m, N, B = 12, 4, 256, 4096
    kn, x0 = R.synth_problems(n, m, N, 2024, 0, B)
    s = R.CudaSolver(n, m, N, batch=B, flags=0)
    assert s.mode == R.MODE_SCAN
    s.upload(kn, x0)
    s.solve()
    assert not s.flags().any()
    z = s.solution().copy()
    s.close()
    assert np.isfinite(z).all()
    for p in (0, 1, 2047, 4095):
        ref, _ = O.port_solve(n, m, N, kn[p], x0[p])
        assert max(rel_err_classes(z[p], ref, n, m, N)) < TOL, p
        kkt_ref = O.kkt_residual(n, m, N, kn[p], x0[p], ref)
        assert O.kkt_residual(n, m, N, kn[p], x0[p], z[p]) <= 2 * kkt_ref + 1e-13 * np.linalg.norm(ref), p
    worst = max(O.kkt_residual(n, m, N, kn[p], x0[p], z[p]) / np.linalg.norm(z[p]) for p in range(0, B, 61))
    assert worst < 1e-14 * N, worst
    ids = [5, 700, 701, 1999, 2500, 3333, 4000, 4095]
    small = R.CudaSolver(n, m, N, batch=len(ids), flags=0)
    small.upload(np.ascontiguousarray(kn[ids]), np.ascontiguousarray(x0[ids]))
    small.solve()
    assert np.array_equal(small.solution(), z[ids])
    small.close()


def test_large_state_n64_N1024_vs_reference():
    # configs[4]: n=64, m=16, N=1024 (755 MB dense factor array per problem) against the reference's stored solution
    n, m, N = 64, 16, 1024
    g = _large_golden()
    ref, ref_kkt = g["gen_n64_m16_N1024_id0_soln"], float(g["gen_n64_m16_N1024_id0_kkt"])
    kn, x0 = R.synth_problems(n, m, N, 2024, 0, 1)
    s = R.CudaSolver(n, m, N, batch=1, flags=0)
    s.upload(kn, x0)
    s.solve()
    z = s.solution()[0].copy()
    assert not s.flags().any()
    s.close()
    errs = rel_err_classes(z, ref, n, m, N)
    assert max(errs) < TOL, errs
    kkt = O.kkt_residual(n, m, N, kn[0], x0[0], z)
    assert kkt <= 2 * ref_kkt + 1e-13 * np.linalg.norm(z), (kkt, ref_kkt)


def test_default_drop_in_solve_takes_the_fast_path_at_16384():
    # ndlqr_NewNdLqrSolver(12, 4, 16384) + Initialize + Solve: the on-chip (scan) kernels, no 617 MB factor array on
    # either side of the bus
    LIB = R.LIB
    n, m, N = 12, 4, 16384
    pp = LIB.rslqr_synth_problem(n, m, N, 2024, 0)
    g = _large_golden()
    ref = g["quad_n12_m4_N16384_id0_soln"]
    s = R.NdLqrSolver(n, m, N)
    assert s.initialize(pp) == 0 and s.solve() == 0
    assert LIB.rslqr_cuda_get_mode(s.c.device) == R.MODE_SCAN
    assert not s.c.pinned_fact and not s.c.fact_dirty
    ptrs = [C.c_void_p() for _ in range(4)]
    LIB.rslqr_cuda_device_pointers(s.c.device, *[C.byref(p) for p in ptrs])
    assert not ptrs[3].value                                  # no device factor array was ever allocated
    z = np.zeros(N * 28)
    z[:s.c.nvars] = s.solution()
    assert max(rel_err_classes(z, ref, n, m, N)) < TOL
    for _ in range(3):
        assert s.initialize(pp) == 0 and s.solve() == 0
    assert s.c.solve_time_ms < 5.0                            # (0.2-0.3 ms on a B200; the staged path takes ~40 ms)
    s.close()
    LIB.ndlqr_FreeLQRProblem(pp)


# ---------------------------------------------------------------------------------------------
# more than one device (skipped on a one-GPU box; run with `gpurun --gpus 2`)
# ---------------------------------------------------------------------------------------------
def _ndev():
    return R.LIB.rslqr_cuda_device_count()


@pytest.mark.skipif(_ndev() < 2, reason="needs two GPUs")
def test_two_devices_in_one_process():
    # the > 48 KB shared-memory opt-in is per device (cudaFuncSetAttribute): handles on devices 0 and 1 of ONE process
    # must both launch the scan / staged kernels
    LIB = R.LIB
    n, m, N, B = 12, 4, 64, 7
    kn, x0 = R.synth_problems(n, m, N, 33, 0, B)
    want = [O.port_solve(n, m, N, kn[p], x0[p])[0] for p in range(B)]
    hs = [R.CudaSolver(n, m, N, batch=B, device=d, flags=0) for d in (0, 1)]
    for mode in (R.MODE_SCAN, R.MODE_FUSED, R.MODE_STAGED):
        for h in hs:
            h.set_mode(mode)
            h.upload(kn, x0)
            h.solve_async(True)
        for d, h in enumerate(hs):
            h.sync()
            assert not h.flags().any()
            z = h.solution()
            for p in range(B):
                assert max(rel_err_classes(z[p], want[p], n, m, N)) < TOL, (mode, d, p)
    for h in hs:
        h.riccati()
        h.close()
    # n=64 staged kernels with > 48 KB of dynamic shared memory on the second device
    kn2, x2 = R.synth_problems(64, 16, 16, 3, 0, 1)
    w2, _ = O.port_solve(64, 16, 16, kn2[0], x2[0])
    for d in (1, 0):
        h = R.CudaSolver(64, 16, 16, batch=1, device=d, flags=0)
        h.upload(kn2, x2)
        h.solve()
        assert max(rel_err_classes(h.solution()[0], w2, 64, 16, 16)) < TOL
        h.close()
    # the batch host API on device 1
    LIB.ndlqr_NewBatchSolver.restype = C.c_void_p
    LIB.ndlqr_NewBatchSolver.argtypes = [C.c_int] * 5
    for fn in ("ndlqr_FreeBatchSolver", "ndlqr_BatchSolve", "ndlqr_BatchGetNumVars"):
        getattr(LIB, fn).argtypes = [C.c_void_p]
    LIB.ndlqr_BatchInitialize.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
    LIB.ndlqr_BatchCopySolution.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
    probs = [LIB.rslqr_synth_problem(n, m, N, 33, i) for i in range(B)]
    arr = (C.c_void_p * B)(*[C.cast(p, C.c_void_p) for p in probs])
    bs = LIB.ndlqr_NewBatchSolver(n, m, N, B, 1)
    assert bs and LIB.ndlqr_BatchInitialize(bs, arr, B) == 0 and LIB.ndlqr_BatchSolve(bs) == 0
    nv = LIB.ndlqr_BatchGetNumVars(bs)
    for p in range(B):
        z = np.zeros(nv)
        assert LIB.ndlqr_BatchCopySolution(bs, p, z.ctypes.data_as(C.c_void_p)) == nv
        assert np.linalg.norm(z - want[p][:nv]) / np.linalg.norm(want[p]) < TOL
    LIB.ndlqr_FreeBatchSolver(bs)
    for p in probs:
        LIB.ndlqr_FreeLQRProblem(p)


@pytest.mark.skipif(_ndev() < 2, reason="needs two GPUs")
def test_long_horizon_split_over_real_gpus():
    # SURVEY 8e row 2 on real devices: one rank per GPU under torchrun, both transports (peer stores + flags, NCCL),
    # KKT residual and agreement with the one-GPU solve (tests/dist_long_check.py asserts them on rank 0)
    import subprocess
    import sys
    world = 2 if _ndev() < 4 else (4 if _ndev() < 8 else 8)
    here = os.path.dirname(os.path.abspath(__file__))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", "29731", os.path.join(here, "dist_long_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.stdout[-2000:], r.stderr[-2000:])
    assert "exchange=p2p" in r.stdout or "exchange=nccl" in r.stdout
