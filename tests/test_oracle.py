"""CPU tests that PIN the oracle: the C port (oracle/ndlqr_port.c) against the reference's own golden
vectors (tests/golden/*.json = the reference's data files), against outputs of the unmodified
reference stored in tests/golden/ref_synth.npz, and - when oracle/_ref is built - against the
compiled reference directly. Reference tests mirrored: test/nested_dissection_test.c,
test/sample_problem_test.c, test/binarytree_test.c, test/solver_test.c."""
import os

import numpy as np
import pytest

import oracle_lib as O
from helpers import GOLDEN, blocks, golden_json, jmat, problem_from_json, rel_err_classes

import rslqr_b200 as R


def test_tree_closed_forms():
    # test/binarytree_test.c:23-61
    assert [O.PORT.port_index_level(k) for k in range(7)] == [0, 1, 0, 2, 0, 1, 0]
    assert O.PORT.port_index_at_level(5, 0) == 4
    assert O.PORT.port_index_at_level(3, 0) == 2
    assert O.PORT.port_index_at_level(2, 2) == 3
    assert O.PORT.port_index_at_level(7, 2) == 3
    assert O.PORT.port_index_at_level(7, 0) == 6
    assert O.PORT.port_index_from_leaf(1, 1) == 5


def test_init_placement():
    # test/solver_test.c:49-143 (N=8, n=6, m=3): A^T/B^T/-I placement and negated rhs
    n, m, N, knots, x0, d = problem_from_json("lqr_prob.json")
    s = O.PortSolver(n, m, N)
    s.init(knots, x0)
    A0 = jmat(d["lqrdata"][0]["A"])
    B0 = jmat(d["lqrdata"][0]["B"])
    L, X, U = blocks(s.data[0, 0], n, m)
    assert np.array_equal(X, A0.T) and np.array_equal(U, B0.T) and not L.any()
    L, X, U = blocks(s.data[0, 1], n, m)
    assert np.array_equal(X, -np.eye(n)) and not U.any()
    L, X, U = blocks(s.data[2, 3], n, m)  # knot 3 has level 2
    assert np.array_equal(X, jmat(d["lqrdata"][3]["A"]).T)
    assert np.array_equal(s.soln[0, :n], -x0)
    assert np.array_equal(s.soln[1, :n], -np.array(d["lqrdata"][0]["d"]))
    assert np.array_equal(s.soln[0, n:2 * n], -np.array(d["lqrdata"][0]["q"]))
    assert np.array_equal(s.soln[N - 1, 2 * n:], np.zeros(m))
    assert s.diag[N - 1, 0] == d["lqrdata"][N - 1]["Q"][0]


def test_leaf_stage_golden():
    # test/nested_dissection_test.c:11-114: rhs after the leaves == sample_problem.json "b"
    n, m, N, knots, x0, d = problem_from_json("lqr_prob.json")
    sp = golden_json("sample_problem.json")
    s = O.PortSolver(n, m, N)
    s.init(knots, x0)
    s.leaves()
    b = np.array(sp["b"], dtype=np.float64).reshape(-1)
    assert np.linalg.norm(s.soln.reshape(-1) - b) < 1e-6
    # F(0,0) = [-A0^T ; 0 ; R \ B0^T]
    A0, B0 = jmat(d["lqrdata"][0]["A"]), jmat(d["lqrdata"][0]["B"])
    R0 = np.array(d["lqrdata"][0]["R"])
    L, X, U = blocks(s.fact[0, 0], n, m)
    assert np.allclose(L, -A0.T, atol=1e-14) and not X.any()
    assert np.allclose(U, B0.T / R0[:, None], atol=1e-13)


def test_schur_stage_golden():
    # test/nested_dissection_test.c:138-237: level-0 Schur results E{k}{p}{y,x,u}
    n, m, N, knots, x0, _ = problem_from_json("lqr_prob.json")
    sp = golden_json("sample_problem.json")
    s = O.PortSolver(n, m, N)
    s.init(knots, x0)
    s.leaves()
    s.products(0)
    s.cholesky(0)
    s.solves(0)
    s.schur(0)
    for k in (0, 1):
        for p in (1, 2):
            L, X, U = blocks(s.fact[p, k], n, m)
            assert np.linalg.norm(L - jmat(sp[f"E{k}{p}y"])) < 1e-6
            assert np.linalg.norm(X - jmat(sp[f"E{k}{p}x"])) < 1e-6
            assert np.linalg.norm(U - jmat(sp[f"E{k}{p}u"])) < 1e-6


@pytest.mark.parametrize("name", ["lqr_prob.json", "lqr_prob_256.json"])
def test_full_solve_golden(name):
    # test/nested_dissection_test.c:239-283, test/sample_problem_test.c: solve vs stored `soln`
    n, m, N, knots, x0, d = problem_from_json(name)
    soln, fact = O.port_solve(n, m, N, knots, x0, threads=2)
    want = np.array(d["soln"], dtype=np.float64).reshape(-1)
    nvars = (2 * n + m) * N - m
    assert np.linalg.norm(soln[:nvars] - want) < 1e-6
    assert np.linalg.norm(soln[:nvars] - want) / np.linalg.norm(want) < 1e-13
    assert O.kkt_residual(n, m, N, knots, x0, soln) < 1e-9
    if name == "lqr_prob.json":  # F{k}2{y,x,u}: top-level column of the finished factorization
        sp = golden_json("sample_problem.json")
        for k in range(N):
            L, X, U = blocks(fact[2, k], n, m)
            assert np.linalg.norm(L - jmat(sp[f"F{k}2y"])) < 1e-9
            assert np.linalg.norm(X - jmat(sp[f"F{k}2x"])) < 1e-9
            assert np.linalg.norm(U - jmat(sp[f"F{k}2u"])) < 1e-9


def test_port_vs_stored_reference_outputs():
    # outputs of the unmodified reference on synthetic problems (tests/golden/make_golden.py)
    g = np.load(os.path.join(GOLDEN, "ref_synth.npz"))
    for name, n, m, N, ids in [("quad_n12_m4_N256", 12, 4, 256, [0, 4095]), ("quad_n12_m4_N32", 12, 4, 32, [0, 7]),
                               ("gen_n6_m3_N64", 6, 3, 64, [3]), ("gen_n16_m5_N16", 16, 5, 16, [2]),
                               ("gen_n64_m16_N8", 64, 16, 8, [0])]:
        for pid in ids:
            kn, x0 = R.synth_problems(n, m, N, 2024, pid, 1)
            soln, fact = O.port_solve(n, m, N, kn[0], x0[0])
            key = f"{name}_id{pid}"
            assert max(rel_err_classes(soln, g[key + "_soln"], n, m, N)) < 1e-11, key
            assert np.allclose(fact[-1], g[key + "_fact_top"], rtol=1e-9, atol=1e-12), key
            assert O.kkt_residual(n, m, N, kn[0], x0[0], soln) <= 2 * float(g[key + "_kkt"]) + 1e-13


@pytest.mark.skipif(O.REF is None, reason="oracle/_ref not built")
def test_port_vs_compiled_reference_stagewise():
    # every intermediate `fact` block and the rhs agree with the compiled reference
    for (n, m, N, pid) in [(6, 3, 8, 1), (12, 4, 64, 5), (5, 2, 16, 9)]:
        kn, x0 = R.synth_problems(n, m, N, 77, pid, 1)
        soln, fact = O.port_solve(n, m, N, kn[0], x0[0], threads=3)
        rsoln, rfact, _ = O.ref_solve(n, m, N, kn[0], x0[0], threads=2)
        assert np.allclose(fact, rfact, rtol=1e-12, atol=1e-13)
        assert np.allclose(soln, rsoln, rtol=1e-12, atol=1e-12)


def test_riccati_agrees():
    # test/riccati_solver_test.c:339-364: Riccati vs stored solution (1e-10 at N=8)
    n, m, N, knots, x0, d = problem_from_json("lqr_prob.json")
    z = O.riccati(n, m, N, knots, x0)
    want = np.array(d["soln"], dtype=np.float64).reshape(-1)
    assert np.linalg.norm(z[:want.size] - want) < 1e-8


def test_port_rejects_bad_horizon():
    # test/nddata_test.c:15-22
    assert not O.PORT.port_new(6, 3, 7)
    assert not O.PORT.port_new(0, 3, 8)
