"""Pins the CPU oracle (oracle/sysml_oracle.cpp).

The reference holds no golden vectors for numeric ops (its tests regenerate inputs with java.util.Random
and compare against R), so the oracle is pinned by (a) known answers of
java.util.Random, (b) exact rational arithmetic on small cases (tests/golden/*.json, made by
tests/golden/make_golden.py), (c) extended-precision numpy at the reference tests' own shapes, seeds and
tolerances, (d) the reference's documented edge-case semantics.
"""
import json
import os

import numpy as np
import pytest

from oracle import oracle as O

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_java_random_known_answers():
    # new java.util.Random(42).nextDouble() and new Random(0).nextDouble(): JDK-documented LCG
    assert O.java_next_doubles(42, 1)[0] == 0.7275636800328681
    assert O.java_next_doubles(0, 1)[0] == 0.730967787376657


def test_generate_test_matrix_sparsity():
    X = O.java_test_matrix(3471, 567, 0, 1, 0.7, 7)
    assert abs((X != 0).mean() - 0.7) < 0.01 and X.min() >= 0 and X.max() < 1
    assert np.array_equal(X, O.java_test_matrix(3471, 567, 0, 1, 0.7, 7))


def test_balanced_blocks_rule():
    # UtilFunctions.getBalancedBlockSizesDefault (UtilFunctions.java:156-180)
    assert O.balanced_blocks(100, 8, True) == [13, 13, 13, 13, 12, 12, 12, 12]
    b = O.balanced_blocks(1000, 8, False)  # nk = roundToNext(min(64, 31), 8) = 32
    assert len(b) == 32 and sum(b) == 1000 and max(b) - min(b) <= 1
    assert O.balanced_blocks(3, 8, True) == [1, 1, 1]


@pytest.mark.parametrize("name", ["mmchain_small", "tsmm_small", "matmult_small", "agg_small"])
def test_golden_exact_rational(name):
    g = json.load(open(os.path.join(GOLD, name + ".json")))
    X = np.array(g["X"], dtype=np.float64)
    if name == "mmchain_small":
        v, w = np.array(g["v"]), np.array(g["w"])
        for ct, key in ((0, "XtXv"), (1, "XtwXv"), (2, "XtXvy")):
            for k in (1, 4):
                got = O.mmchain_dense(X, v, w if ct else None, ct, k)
                assert np.allclose(got, np.array(g[key]), rtol=1e-14, atol=0)
    elif name == "tsmm_small":
        T, nnz = O.tsmm_dense(X, True, 1)
        assert np.allclose(T, np.array(g["left"]), rtol=1e-14, atol=0) and nnz == g["left_nnz"]
        T, nnz = O.tsmm_dense(X, False, 1)
        assert np.allclose(T, np.array(g["right"]), rtol=1e-14, atol=0)
    elif name == "matmult_small":
        B = np.array(g["B"])
        Cm, nnz = O.matmult_dense(X, B, 1)
        assert np.allclose(Cm, np.array(g["C"]), rtol=1e-14, atol=0)
        rp, ci, va = O.dense_to_csr(X)
        Cs, _ = O.matmult_csr_dense(rp, ci, va, X.shape, B, 1)
        assert np.allclose(Cs, np.array(g["C"]), rtol=1e-14, atol=0)
    else:
        assert np.allclose(O.agg_unary_dense(X, O.SUM, O.ALL), g["sum"], rtol=1e-15)
        assert np.allclose(O.agg_unary_dense(X, O.SUM, O.ROW), g["rowsums"], rtol=1e-15)
        assert np.allclose(O.agg_unary_dense(X, O.SUM, O.COL), g["colsums"], rtol=1e-15)
        assert np.allclose(O.agg_unary_dense(X, O.SUMSQ, O.ALL), g["sumsq"], rtol=1e-15)
        assert np.array_equal(O.agg_unary_dense(X, O.MAXINDEX, O.ROW), g["rowindexmax"])
        assert np.array_equal(O.agg_unary_dense(X, O.MININDEX, O.ROW), g["rowindexmin"])


@pytest.mark.parametrize("sparsity", [0.7, 0.1])
def test_mmchain_at_reference_test_shape(sparsity):
    """MapMultChainTest: 3471 x 567, seeds 7/3/10, eps 1e-8 vs R (here: numpy longdouble)."""
    X = O.java_test_matrix(3471, 567, 0, 1, sparsity, 7)
    v = O.java_test_matrix(567, 1, 0, 1, 0.7, 3).ravel()
    w = O.java_test_matrix(3471, 1, 0, 1, sparsity, 10).ravel()
    Xl, vl, wl = X.astype(np.longdouble), v.astype(np.longdouble), w.astype(np.longdouble)
    refs = {0: Xl.T @ (Xl @ vl), 1: Xl.T @ (wl * (Xl @ vl)), 2: Xl.T @ ((Xl @ vl) - wl)}
    rp, ci, va = O.dense_to_csr(X)
    for ct, ref in refs.items():
        for k in (1, 8):
            d = O.mmchain_dense(X, v, w if ct else None, ct, k)
            s = O.mmchain_csr(rp, ci, va, X.shape, v, w if ct else None, ct, k)
            assert np.max(np.abs(d - ref)) < 1e-8 and np.max(np.abs(s - ref)) < 1e-8
            assert np.max(np.abs((d - ref) / ref)) < 1e-13


def test_tsmm_at_reference_test_shape():
    """FullMatrixMultiplicationTransposeSelfTest: 3500 x 1500 LEFT (scaled to 3500 x 300 for CPU time), eps 1e-10."""
    X = O.java_test_matrix(3500, 300, 0, 1, 0.7, 7)
    Xl = X.astype(np.longdouble)
    T, nnz = O.tsmm_dense(X, True, 8)
    ref = Xl.T @ Xl
    assert np.max(np.abs(T - ref)) < 1e-10 and np.array_equal(T, T.T) and nnz == np.count_nonzero(T)
    rp, ci, va = O.dense_to_csr(X)
    S, nnz2 = O.tsmm_csr_left(rp, ci, va, X.shape, 8)
    assert np.max(np.abs(S - ref)) < 1e-10 and nnz2 == nnz
    X2 = O.java_test_matrix(300, 2000, 0, 1, 0.7, 7)
    T2, _ = O.tsmm_dense(X2, False, 8)
    assert np.max(np.abs(T2 - X2.astype(np.longdouble) @ X2.astype(np.longdouble).T)) < 1e-10


@pytest.mark.parametrize("m,cd,n", [(1, 500, 1), (300, 1, 40), (300, 500, 1), (300, 3000, 1), (1, 500, 70), (9, 500, 70),
                                    (400, 300, 20), (200, 70, 90), (33, 2000, 3)])
def test_matmult_dispatch_cases(m, cd, n):
    """Every special case of matrixMultDenseDense / SparseDense (LibMatrixMult.java:880-1103, 1170-1395)."""
    A = O.java_test_matrix(m, cd, -1, 1, 0.6, 5)
    B = O.java_test_matrix(cd, n, -1, 1, 0.9, 6)
    ref = A.astype(np.longdouble) @ B.astype(np.longdouble)
    rp, ci, va = O.dense_to_csr(A)
    for k in (1, 8):
        Cd, nnz = O.matmult_dense(A, B, k)
        assert np.max(np.abs(Cd - ref)) < 1e-11 and nnz == np.count_nonzero(Cd)
        Cs, _ = O.matmult_csr_dense(rp, ci, va, A.shape, B, k)
        assert np.max(np.abs(Cs - ref)) < 1e-11


def test_kahan_sum_and_threaded_merge():
    rng = np.random.default_rng(0)
    X = rng.uniform(-1, 1, (20000, 64)) * 1e8 + 1.0
    ref = float(np.sum(X.astype(np.longdouble)))
    for k in (1, 8):
        s = O.agg_unary_dense(X, O.SUM, O.ALL, k)[0]
        assert abs(s - ref) <= 2 * np.spacing(abs(ref))
    cs = O.agg_unary_dense(X, O.SUM, O.COL, 8)
    assert np.max(np.abs(cs - np.sum(X.astype(np.longdouble), axis=0).astype(np.float64))) <= 4 * np.spacing(np.max(np.abs(cs)))


def test_aggregate_inf_nan_semantics():
    """AggregateInfTest / AggregateNaNTest: KahanPlus fast path keeps +-Inf; Math.max propagates NaN."""
    X = np.array([[1.0, np.inf, 2.0], [3.0, 4.0, 5.0]])
    assert O.agg_unary_dense(X, O.SUM, O.ALL)[0] == np.inf
    assert O.agg_unary_dense(-X, O.SUM, O.ALL)[0] == -np.inf
    Y = np.array([[1.0, np.nan, 2.0], [3.0, 4.0, 5.0]])
    assert np.isnan(O.agg_unary_dense(Y, O.MAX, O.ALL)[0]) and np.isnan(O.agg_unary_dense(Y, O.SUM, O.ALL)[0])
    assert O.agg_unary_dense(Y, O.MAX, O.ROW)[1] == 5.0


def test_rowindexmax_semantics():
    """indexmax uses >=: the LAST maximum wins, 1-based (LibMatrixAgg.java:3081-3089); sparse rows resolve
    implicit zeros to the last zero column only when 0 STRICTLY beats the best stored value (:2550-2557)."""
    X = np.array([[1.0, 3.0, 3.0, 2.0], [-1.0, -2.0, -1.0, -3.0], [0.0, 0.0, 0.0, 0.0], [-5.0, 0.0, -2.0, 0.0],
                  [np.nan, 1.0, np.nan, 1.0], [0.0, -1.0, 0.0, -2.0]])
    assert O.agg_unary_dense(X, O.MAXINDEX, O.ROW).tolist() == [3, 3, 4, 4, 4, 3]
    assert O.agg_unary_dense(X, O.MININDEX, O.ROW).tolist() == [1, 4, 4, 1, 4, 4]
    rp, ci, va = O.dense_to_csr(X)
    assert O.agg_unary_csr(rp, ci, va, X.shape, O.MAXINDEX, O.ROW).tolist() == [3, 3, 4, 4, 4, 3]
    assert O.agg_unary_csr(rp, ci, va, X.shape, O.MININDEX, O.ROW).tolist() == [1, 4, 4, 1, 4, 4]


def test_cellwise_value_functions():
    nan, inf = float("nan"), float("inf")
    B = O.binary_scalar
    assert B(7.5, 2.0, 17) == 1.5 and B(-7.5, 2.0, 17) == 0.5 and np.isnan(B(1.0, 0.0, 17))  # %% = x - floor(x/y)*y
    assert B(7.5, 2.0, 18) == 3.0 and B(-7.5, 2.0, 18) == -4.0 and B(1.0, 0.0, 18) == inf      # %/%
    assert B(0.5, 4.0, 15) == -1.0 and B(0.0, 3.0, 16) == 0.0 and B(2.0, 3.0, 16) == -1.0      # 1-*  and -nz
    assert B(2.0, 0.0, 13) == 0.0 and B(2.0, -1.0, 13) == 1.0 and B(0.0, 0.0, 14) == 0.0       # && ||
    assert np.isnan(B(nan, 1.0, 12)) and np.isnan(B(1.0, nan, 11))                              # Math.max/min
    assert np.signbit(B(-0.0, 0.0, 11)) and not np.signbit(B(-0.0, 0.0, 12))
    assert B(2.0, 10.0, 4) == 1024.0 and B(1.0, 1.0, 9) == 1.0 and B(nan, nan, 9) == 0.0 and B(nan, nan, 10) == 1.0


def test_rand_generator_is_shardable():
    full = O.rand_uniform(1000, seed=9)
    assert np.array_equal(full[300:700], O.rand_uniform(400, seed=9, offset=300))
    assert full.min() >= -1 and full.max() < 1 and abs(full.mean()) < 0.1


def test_variance_matches_two_pass_longdouble():
    X = O.java_test_matrix(500, 40, -1, 1, 0.9, 42) + 3.0
    Xl = X.astype(np.longdouble)
    assert abs(O.agg_unary_dense(X, O.VAR, O.ALL)[0] - float(Xl.var(ddof=1))) < 1e-13
    assert np.max(np.abs(O.agg_unary_dense(X, O.VAR, O.ROW) - Xl.var(axis=1, ddof=1).astype(np.float64))) < 1e-13
    assert np.max(np.abs(O.agg_unary_dense(X, O.VAR, O.COL) - Xl.var(axis=0, ddof=1).astype(np.float64))) < 1e-13
    assert O.agg_unary_dense(np.array([[5.0]]), O.VAR, O.ALL)[0] == 0.0  # w == 1 -> 0 (CM_COV_Object.java:170)
