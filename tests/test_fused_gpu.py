"""Fused cellwise chains (sysml_b200_fused_cellwise) against the UNFUSED reference evaluation of the same chain
(oracle.fused_cellwise: one LibMatrixBincell op per operator, then LibMatrixAgg) and against this library's own
unfused kernels.  The chains are the inner-loop expressions of the reference's algorithm scripts:
  l2-svm.dml        out = 1 - Y * Xw;  sv = out > 0;  g = sum(out * sv * Y * Xd);  h = sum(Xd * sv * Xd)
  GLM.dml (poisson) w = exp(eta); z = eta + (y - w) / w
  MultiLogReg.dml   P = exp(LT - rowMaxs) ; rowSums(P)
  Kmeans.dml        D = -2 * XC + t(rowSums(C^2)); P = (D <= minD)
"""
import numpy as np
import pytest

from oracle import oracle as O
from tests.util import assert_equal_matrices

pytestmark = pytest.mark.gpu


def _run(gctx, arrays, expr, consts=(), agg=None, rows=None, cols=None):
    from systemml_b200 import LibMatrixB200 as L
    objs = [gctx.dense(a) for a in arrays]
    got = L.fusedCellwise(gctx, "spoof", objs, expr, consts, agg, rows, cols)
    if not isinstance(got, float):
        h = got.to_host()
        got.free()
        got = h
    for o in objs:
        o.free()
    return got


def _close(exp, got, what):
    """Chains that contain exp(): numpy's and CUDA's exp differ by an ulp and later terms cancel, so the bound is
    absolute, a few ulp of the largest intermediate (operands here are O(1..10))."""
    assert np.array_equal(np.isnan(exp), np.isnan(got)), what
    d = np.nanmax(np.abs(exp - got)) if exp.size else 0.0
    assert d <= 64 * np.finfo(np.float64).eps * max(1.0, float(np.nanmax(np.abs(exp)))), "%s: max abs diff %.3e" % (what, d)


def _oracle(arrays, expr, consts, agg, rows, cols):
    from systemml_b200.lib_matrix import FUSED_AGG
    return O.fused_cellwise(arrays, consts, expr.tokens, rows, cols, FUSED_AGG[agg])


@pytest.mark.parametrize("n", [1, 7, 4096, 100_001, 1_000_000])
def test_l2svm_chain_vectors(gctx, n):
    from systemml_b200.lib_matrix import Expr as E
    rng = np.random.default_rng(n)
    Y = np.where(rng.uniform(size=(n, 1)) > 0.5, 1.0, -1.0)
    Xw, Xd = rng.normal(size=(n, 1)), rng.normal(size=(n, 1))
    one, zero = E.const(0), E.const(1)
    out = one - E.input(0) * E.input(1)              # 1 - Y * Xw
    sv = out.binary(zero, ">")                        # out > 0
    g = out * sv * E.input(0) * E.input(2)            # out * sv * Y * Xd
    for expr, agg in ((out, None), (sv, None), (g, "sum"), (E.input(2) * sv * E.input(2), "sum")):
        exp = _oracle([Y, Xw, Xd], expr, (1.0, 0.0), agg, n, 1)
        got = _run(gctx, [Y, Xw, Xd], expr, (1.0, 0.0), agg)
        if agg is None:
            assert np.array_equal(exp, got), "a fused chain without an aggregate is bit-identical to the unfused ops"
        else:
            assert abs(exp[0] - got) <= 1e-9 * max(abs(exp[0]), 1e-300) + 1e-12 * np.sqrt(n)


@pytest.mark.parametrize("rows,cols", [(1, 1), (64, 130), (1024, 21), (2049, 64), (777, 1025), (5, 5000), (100_000, 50), (3, 1)])
def test_broadcast_chains_all_aggregates(gctx, rows, cols):
    """matrix (op) column vector (op) row vector (op) scalar, with every aggregate."""
    from systemml_b200.lib_matrix import Expr as E
    rng = np.random.default_rng(rows * 31 + cols)
    A, B = rng.uniform(-1, 1, (rows, cols)), rng.uniform(0.5, 2, (rows, cols))
    cv, rv = rng.uniform(-1, 1, (rows, 1)), rng.uniform(1, 2, (1, cols))
    # (A * cv - B / rv) * k0 + exp(A)            -- Kmeans / MultiLogReg style mixes
    expr = (E.input(0) * E.input(2) - E.input(1) / E.input(3)) * E.const(0) + E.input(0).unary("exp")
    for agg in (None, "sum", "rowSums", "colSums"):
        exp = _oracle([A, B, cv, rv], expr, (-2.0,), agg, rows, cols)
        got = _run(gctx, [A, B, cv, rv], expr, (-2.0,), agg)
        if agg is None:
            _close(exp, got, "fused broadcast chain")
        elif agg == "sum":
            assert abs(exp[0] - got) <= 1e-9 * abs(exp[0]) + 1e-10
        else:
            assert_equal_matrices(exp.reshape(-1), got.reshape(-1), rel=1e-9, what="fused %s" % agg)


def test_matches_unfused_kernels_bitwise(gctx):
    """Same chain through this library's own one-op-per-launch kernels: bit-identical (same value functions)."""
    from systemml_b200 import LibMatrixB200 as L
    from systemml_b200.lib_matrix import Expr as E
    rng = np.random.default_rng(9)
    rows, cols = 5000, 37
    A, B = rng.uniform(-3, 3, (rows, cols)), rng.uniform(-3, 3, (rows, cols))
    cv = rng.uniform(0.1, 2, (rows, 1))
    dA, dB, dc = gctx.dense(A), gctx.dense(B), gctx.dense(cv)
    # ((A %% cv) ^ 2 min B) 1-* A ; sigmoid(.)
    t1 = L.matrixMatrixArithmetic(gctx, "%%", dA, dc, "%%")
    t2 = L.matrixScalarArithmetic(gctx, "^", t1, 2.0, "^")
    t3 = L.matrixMatrixArithmetic(gctx, "min", t2, dB, "min")
    t4 = L.matrixMatrixArithmetic(gctx, "1-*", t3, dA, "1-*")
    t5 = L.unaryOp(gctx, "sigmoid", t4, "sigmoid")
    expr = E.input(0).binary(E.input(2), "%%").binary(E.const(0), "^").binary(E.input(1), "min").binary(E.input(0), "1-*").unary("sigmoid")
    got = L.fusedCellwise(gctx, "spoof", [dA, dB, dc], expr, (2.0,))
    assert np.array_equal(t5.to_host(), got.to_host())
    # and the cached kernel is reused for new data / new constants
    got2 = L.fusedCellwise(gctx, "spoof", [dB, dA, dc], expr, (3.0,))
    u1 = L.matrixMatrixArithmetic(gctx, "%%", dB, dc, "%%")
    u2 = L.matrixScalarArithmetic(gctx, "^", u1, 3.0, "^")
    u3 = L.matrixMatrixArithmetic(gctx, "min", u2, dA, "min")
    u4 = L.matrixMatrixArithmetic(gctx, "1-*", u3, dB, "1-*")
    u5 = L.unaryOp(gctx, "sigmoid", u4, "sigmoid")
    assert np.array_equal(u5.to_host(), got2.to_host())


def test_glm_poisson_chain(gctx):
    from systemml_b200.lib_matrix import Expr as E
    rng = np.random.default_rng(4)
    n = 200_003
    eta, y = rng.uniform(-2, 2, (n, 1)), rng.poisson(2.0, (n, 1)).astype(np.float64)
    w = E.input(0).unary("exp")
    z = E.input(0) + (E.input(1) - w) / w
    exp = _oracle([eta, y], z, (), None, n, 1)
    got = _run(gctx, [eta, y], z)
    _close(exp, got, "GLM working response")
    dev = _run(gctx, [eta, y], (E.input(1) - w) * (E.input(1) - w) / w, (), "sum")
    e = _oracle([eta, y], (E.input(1) - w) * (E.input(1) - w) / w, (), "sum", n, 1)[0]
    assert abs(dev - e) <= 1e-9 * abs(e)


def test_special_values_and_errors(gctx):
    from systemml_b200 import LibMatrixB200 as L, native
    from systemml_b200.lib_matrix import Expr as E
    A = np.array([[np.nan, 0.0, -0.0, np.inf], [1.0, -np.inf, 5.0, -7.5]])
    B = np.array([[1.0, -0.0, 0.0, 2.0], [0.0, 3.0, np.nan, 2.0]])
    for op in ("max", "min", "%%", "%/%", "/", "!=", "-nz"):
        expr = E.input(0).binary(E.input(1), op)
        exp = _oracle([A, B], expr, (), None, 2, 4)
        got = _run(gctx, [A, B], expr)
        assert np.array_equal(exp.view(np.int64), got.view(np.int64)) or np.array_equal(np.isnan(exp), np.isnan(got)) and np.array_equal(
            np.nan_to_num(exp, nan=0.0), np.nan_to_num(got, nan=0.0)), op
    dA = gctx.dense(A)
    with pytest.raises(native.NativeError):  # stack underflow
        L.fusedCellwise(gctx, "spoof", [dA], E([(2 << 24) | 0]))
    with pytest.raises(native.NativeError):  # two values left
        L.fusedCellwise(gctx, "spoof", [dA], E([0, 0]))
    with pytest.raises(native.NativeError):  # input index out of range
        L.fusedCellwise(gctx, "spoof", [dA], E([3]))
    with pytest.raises(native.NativeError):  # shape that does not broadcast
        L.fusedCellwise(gctx, "spoof", [dA, gctx.dense(np.zeros((3, 4)))], E.input(0) + E.input(1))
    # empty operand: sums are zero
    assert L.fusedCellwise(gctx, "spoof", [gctx.dense(np.zeros((0, 5)))], E.input(0), (), "sum", rows=0, cols=5) == 0.0


@pytest.mark.parametrize("rows,cols", [(1, 1), (64, 130), (1024, 21), (777, 1025), (5, 5000), (100_000, 50), (4097, 1)])
def test_min_max_endings(gctx, rows, cols):
    """Chains ending in min / max (SpoofCellwise.AggOp MIN, MAX), all three directions: exact -- no rounding in a
    min or max -- including NaN propagation and -0.0 < +0.0 (Kmeans: rowMins(-2*XC + t(c2)) without materialising D)."""
    from systemml_b200.lib_matrix import Expr as E
    rng = np.random.default_rng(rows + 7 * cols)
    XC = rng.uniform(-1, 1, (rows, cols))
    c2 = rng.uniform(0, 2, (1, cols))
    if rows * cols > 50:
        XC.flat[3], XC.flat[11], XC.flat[20] = np.nan, 0.0, -0.0
    expr = E.const(0) * E.input(0) + E.input(1)
    for agg in ("min", "rowMins", "colMins", "max", "rowMaxs", "colMaxs"):
        exp = _oracle([XC, c2], expr, (-2.0,), agg, rows, cols).reshape(-1)
        got = _run(gctx, [XC, c2], expr, (-2.0,), agg)
        got = np.array([got]) if isinstance(got, float) else got.reshape(-1)
        assert np.array_equal(np.isnan(exp), np.isnan(got)), agg
        ok = ~np.isnan(exp)
        assert np.array_equal(exp[ok].view(np.int64), got[ok].view(np.int64)), agg


@pytest.mark.parametrize("n,cols", [(100_003, 1), (65_536, 1), (7, 1), (3001, 17)])
def test_multi_aggregate_equals_separate_sums(gctx, n, cols):
    """sysml_b200_fused_multi_agg (the code generator's multi-aggregate template, SpoofMultiAggregate.java): l2-svm.dml's two
    line-search sums over the same three vectors in one pass, bit for bit what one fused sum per expression returns, and
    within the usual bar of the unfused oracle evaluation; also with a broadcast operand and four results."""
    from systemml_b200 import LibMatrixB200 as L, native
    E = L.Expr
    Y = np.where(O.rand_uniform(n * cols, seed=1) > 0, 1.0, -1.0).reshape(n, cols)
    Xw = O.rand_uniform(n * cols, seed=2).reshape(n, cols)
    Xd = O.rand_uniform(n * cols, seed=3).reshape(n, cols)
    objs = [gctx.dense(a) for a in (Y, Xw, Xd)]
    consts = (1.0, 0.0, 0.37)
    out_e = E.const(0) - E.input(0) * (E.input(1) + E.const(2) * E.input(2))
    g_e = E.let(out_e, lambda o: o * o.binary(E.const(1), ">") * E.input(0) * E.input(2))
    h_e = E.input(2) * out_e.binary(E.const(1), ">") * E.input(2)
    both = L.fusedMultiAggregate(gctx, "spoof", objs, [g_e, h_e], consts)
    one = [L.fusedCellwise(gctx, "spoof", objs, e, consts, "sum") for e in (g_e, h_e)]
    assert both == one, "same accumulation, same merge order"
    out = 1 - Y * (Xw + 0.37 * Xd)
    sv = (out > 0).astype(np.float64)
    for got, exp in zip(both, (np.sum(out * sv * Y * Xd), np.sum(Xd * sv * Xd))):
        assert abs(got - exp) <= 1e-9 * max(abs(exp), 1.0)
    # a row-vector operand broadcast over the rows, four results (the limit)
    if cols > 1:
        rv = gctx.dense(O.rand_uniform(cols, seed=4).reshape(1, cols))
        four = L.fusedMultiAggregate(gctx, "spoof", objs + [rv], [E.input(0) * E.input(3), E.input(1) * E.input(3), E.input(2), E.input(3)])
        sep = [L.fusedCellwise(gctx, "spoof", objs + [rv], e, (), "sum") for e in (E.input(0) * E.input(3), E.input(1) * E.input(3), E.input(2), E.input(3))]
        assert four == sep
    with pytest.raises(native.NativeError):
        L.fusedMultiAggregate(gctx, "spoof", objs, [E.input(0)] * 5)
    # a program that leaves fewer values than results asked for is malformed
    import ctypes as C
    ptrs = (C.c_void_p * 3)(*[o.ptr for o in objs])
    kd = (C.c_int * 3)(0, 0, 0)
    pg = (C.c_int32 * 1)(0)
    outb = gctx.empty(4, 1)
    rc = gctx._lib.sysml_b200_fused_multi_agg(gctx.handle, ptrs, kd, 3, None, 0, pg, 1, n, cols, 2, C.c_void_p(outb.ptr))
    assert rc == native.E_INVALID
