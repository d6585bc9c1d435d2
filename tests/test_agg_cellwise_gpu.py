"""Aggregate-unary and cellwise parity vs the CPU oracle.

Grids follow gpu/AggregateUnaryOpTests + UnaryOpTestsBase (shapes x sparsities, seed 42, rel 1e-9) and
gpu/MatrixMatrixElementWiseOpTests (sizes {1,64,1024,2049} x {64,130,1024,2049}; matrix-matrix,
matrix-colvector, matrix-rowvector; + - * / ^ %/% %% 1-* < <= > >= == != min max, axpy)."""
import numpy as np
import pytest

from oracle import oracle as O
from tests.util import assert_equal_matrices

pytestmark = pytest.mark.gpu

SHAPES = [(1, 1), (1, 2049), (2049, 1), (64, 130), (1024, 21), (1000, 50), (777, 1025), (5, 5000), (100_000, 50), (3, 3)]


def _mat(rows, cols, sparsity, seed=42):
    return O.java_test_matrix(rows, cols, -1, 1, sparsity, seed)


@pytest.mark.parametrize("rows,cols", SHAPES)
@pytest.mark.parametrize("sparsity", [0.0003, 0.3, 0.9])
def test_sums(gctx, rows, cols, sparsity):
    from systemml_b200 import LibMatrixB200 as L
    X = _mat(rows, cols, sparsity)
    if not X.any():
        X[0, 0] = 0.5
    dX = gctx.dense(X)
    k = O.num_threads()
    for opc, op, d in (("uak+", O.SUM, O.ALL), ("uark+", O.SUM, O.ROW), ("uack+", O.SUM, O.COL),
                       ("uasqk+", O.SUMSQ, O.ALL), ("uarsqk+", O.SUMSQ, O.ROW), ("uacsqk+", O.SUMSQ, O.COL)):
        exp = O.agg_unary_dense(X, op, d, k)
        got = L.unaryAggregate(gctx, opc, dX, opc)
        got = np.array([got]) if d == O.ALL else got.to_host().reshape(-1)
        # compensated on both sides: results agree to a few ulp of the LARGEST partial sum
        tol = 8 * np.spacing(np.max(np.abs(X)) * max(rows, cols)) + 1e-300
        assert np.max(np.abs(exp - got)) <= max(tol, 1e-12 * np.max(np.abs(exp))), opc
    dX.free()


@pytest.mark.parametrize("rows,cols", SHAPES)
def test_min_max_index(gctx, rows, cols):
    from systemml_b200 import LibMatrixB200 as L
    X = _mat(rows, cols, 0.5)
    X[X > 0.8] = 0.8  # force ties so that the last-index rule is exercised
    X[X < -0.8] = -0.8
    dX = gctx.dense(X)
    for opc, op, d in (("uamax", O.MAX, O.ALL), ("uarmax", O.MAX, O.ROW), ("uacmax", O.MAX, O.COL),
                       ("uamin", O.MIN, O.ALL), ("uarmin", O.MIN, O.ROW), ("uacmin", O.MIN, O.COL),
                       ("uarimax", O.MAXINDEX, O.ROW), ("uarimin", O.MININDEX, O.ROW)):
        exp = O.agg_unary_dense(X, op, d, 1)
        got = L.unaryAggregate(gctx, opc, dX, opc)
        got = np.array([got]) if d == O.ALL else got.to_host().reshape(-1)
        assert np.array_equal(exp, got), "%s must be bit-exact" % opc
    dX.free()


def test_index_edge_cases(gctx):
    from systemml_b200 import LibMatrixB200 as L
    X = np.array([[1.0, 3.0, 3.0, 2.0], [-1.0, -2.0, -1.0, -3.0], [0.0, 0.0, 0.0, 0.0], [-5.0, 0.0, -2.0, 0.0],
                  [np.nan, 1.0, np.nan, 1.0], [0.0, -1.0, 0.0, -2.0], [np.inf, 1.0, np.inf, -np.inf]])
    dX = gctx.dense(X)
    assert L.unaryAggregate(gctx, "uarimax", dX, "uarimax").to_host().reshape(-1).tolist() == O.agg_unary_dense(X, O.MAXINDEX, O.ROW).tolist()
    assert L.unaryAggregate(gctx, "uarimin", dX, "uarimin").to_host().reshape(-1).tolist() == O.agg_unary_dense(X, O.MININDEX, O.ROW).tolist()
    # wide row: ties across lanes
    W = np.zeros((3, 1000))
    W[0, [10, 500, 999]] = 7.0
    W[1, :] = -1.0
    W[2, 123] = -3.0
    dW = gctx.dense(W)
    assert L.unaryAggregate(gctx, "uarimax", dW, "uarimax").to_host().reshape(-1).tolist() == [1000, 1000, 1000]
    assert L.unaryAggregate(gctx, "uarimin", dW, "uarimin").to_host().reshape(-1).tolist() == [999, 1000, 124]


def test_inf_nan_aggregates(gctx):
    """AggregateInfTest / AggregateNaNTest."""
    from systemml_b200 import LibMatrixB200 as L
    X = _mat(300, 70, 0.9)
    X[17, 3] = np.inf
    dX = gctx.dense(X)
    assert L.unaryAggregate(gctx, "uak+", dX, "uak+") == np.inf
    rs = L.unaryAggregate(gctx, "uark+", dX, "uark+").to_host().reshape(-1)
    assert rs[17] == np.inf and np.isfinite(np.delete(rs, 17)).all()
    cs = L.unaryAggregate(gctx, "uack+", dX, "uack+").to_host().reshape(-1)
    assert cs[3] == np.inf
    X[17, 3] = np.nan
    dY = gctx.dense(X)
    assert np.isnan(L.unaryAggregate(gctx, "uamax", dY, "uamax")) and np.isnan(L.unaryAggregate(gctx, "uak+", dY, "uak+"))
    mx = L.unaryAggregate(gctx, "uarmax", dY, "uarmax").to_host().reshape(-1)
    assert np.isnan(mx[17]) and mx[16] == np.max(X[16])


def test_mean_prod(gctx):
    from systemml_b200 import LibMatrixB200 as L
    X = O.rand_uniform(300 * 40, seed=5, lo=0.9, hi=1.1).reshape(300, 40)
    dX = gctx.dense(X)
    for opc, op, d in (("uamean", O.MEAN, O.ALL), ("uarmean", O.MEAN, O.ROW), ("uacmean", O.MEAN, O.COL),
                       ("ua*", O.PROD, O.ALL), ("uar*", O.PROD, O.ROW), ("uac*", O.PROD, O.COL)):
        exp = O.agg_unary_dense(X, op, d, 1)
        got = L.unaryAggregate(gctx, opc, dX, opc)
        got = np.array([got]) if d == O.ALL else got.to_host().reshape(-1)
        assert_equal_matrices(exp, got, what=opc)


OPS = ["+", "-", "*", "/", "^", "<", "<=", ">", ">=", "==", "!=", "min", "max", "&&", "||", "1-*", "-nz", "%%", "%/%"]


@pytest.mark.parametrize("rows,cols", [(1, 64), (64, 130), (1024, 1024), (2049, 64), (2049, 2049), (100_003, 7)])
@pytest.mark.parametrize("op", OPS)
def test_binary_matrix_matrix_and_vectors(gctx, rows, cols, op):
    from systemml_b200 import LibMatrixB200 as L
    A = _mat(rows, cols, 0.9, 42)
    dA = gctx.dense(A)
    for shape, seed in (((rows, cols), 43), ((rows, 1), 44), ((1, cols), 45)):
        B = _mat(shape[0], shape[1], 0.9, seed)
        if op == "^":
            A2 = np.abs(A) + 0.1  # keep pow() in its well-conditioned domain
            dA2 = gctx.dense(A2)
            exp = O.binary_mm(A2, B, L.BINARY_OPS[op])
            got = L.matrixMatrixArithmetic(gctx, op, dA2, gctx.dense(B), op).to_host()
            dA2.free()
        else:
            exp = O.binary_mm(A, B, L.BINARY_OPS[op])
            got = L.matrixMatrixArithmetic(gctx, op, dA, gctx.dense(B), op).to_host()
        if op in ("^",):
            assert_equal_matrices(exp, got, rel=1e-13, what="pow")
        else:
            assert np.array_equal(exp, got, equal_nan=True), "%s %s must be bit-exact" % (op, shape)
    dA.free()


@pytest.mark.parametrize("op", OPS)
@pytest.mark.parametrize("left", [False, True])
def test_binary_scalar(gctx, op, left):
    from systemml_b200 import LibMatrixB200 as L
    A = _mat(1031, 77, 0.8, 7)
    if op == "^":
        A = np.abs(A) + 0.1
    for s in (2.0, -0.5, 0.0):
        exp = O.binary_ms(A, s, L.BINARY_OPS[op], left)
        got = L.matrixScalarArithmetic(gctx, op, gctx.dense(A), s, op, left).to_host()
        if op == "^":
            assert_equal_matrices(exp, got, rel=1e-13)
        else:
            assert np.array_equal(exp, got, equal_nan=True), "%s scalar=%s left=%s" % (op, s, left)


def test_axpy_and_unary_and_transpose(gctx):
    from systemml_b200 import LibMatrixB200 as L
    A, B = _mat(513, 129, 0.9, 1), _mat(513, 129, 0.9, 2)
    dA = gctx.dense(A)
    for Bm in (B, B[:, :1].copy(), B[:1, :].copy()):
        for alpha in (2.5, -1.0):
            assert np.array_equal(O.axpy(A, Bm, alpha), L.axpy(gctx, "+*", dA, gctx.dense(Bm), alpha).to_host())
    assert np.array_equal(L.transpose(gctx, "r'", dA).to_host(), A.T)
    P = np.abs(A) + 0.5
    dP = gctx.dense(P)
    for name, fn in (("exp", np.exp), ("log", np.log), ("sqrt", np.sqrt), ("abs", np.abs), ("floor", np.floor), ("ceil", np.ceil),
                     ("sign", np.sign), ("sin", np.sin), ("cos", np.cos), ("tanh", np.tanh)):
        src, d = (A, dA) if name in ("abs", "floor", "ceil", "sign", "exp", "sin", "cos", "tanh") else (P, dP)
        assert_equal_matrices(fn(src), L.unaryOp(gctx, name, d, name).to_host(), rel=1e-14, what=name)
    assert_equal_matrices(1 / (1 + np.exp(-A)), L.unaryOp(gctx, "sigmoid", dA, "sigmoid").to_host(), rel=1e-14)
    assert np.array_equal(O.java_round(A * 10), L.unaryOp(gctx, "round", gctx.dense(A * 10), "round").to_host())
    # round = (double) Math.round(x) and sign = Math.signum(x) on the cases where a shortcut goes wrong (known answers in
    # tests/test_known_answers_cpu.py): halves, 0.49999999999999994, odd integers above 2^52, NaN -> 0, saturation, signed zeros
    S = np.array([[2.5, -2.5, 0.5, -0.5, 0.49999999999999994, -0.5000000000000001, -0.2, -0.0, 0.0, 4503599627370497.0, -4503599627370497.0,
                   np.nan, np.inf, -np.inf, 1e300, -1e300, 2.4999999999999996, 3.5, -3.5, 1e15 + 0.5]])
    for name, ref in (("round", O.java_round), ("sign", O.java_signum)):
        got, exp = L.unaryOp(gctx, name, gctx.dense(S), name).to_host(), ref(S)
        assert np.array_equal(got, exp, equal_nan=True) and np.array_equal(np.signbit(got), np.signbit(exp)), (name, got, exp)


def test_nnz_and_conversions(gctx):
    from systemml_b200 import LibMatrixB200 as L
    X = _mat(1500, 333, 0.05, 9)
    X[3, 4] = -0.0
    X[5, 6] = np.nan
    dX = gctx.dense(X)
    assert dX.getNnz() == O.count_nnz(X)  # -0.0 is zero, NaN is non-zero
    S = L.denseToSparse(gctx, dX)
    rp, ci, va = S.to_host()
    erp, eci, eva = O.dense_to_csr(X)
    assert np.array_equal(rp, erp.astype(np.int32)) and np.array_equal(ci, eci) and np.array_equal(va, eva, equal_nan=True)
    back = L.sparseToDense(gctx, S).to_host()
    assert np.array_equal(np.nan_to_num(back, nan=123.0), np.nan_to_num(np.where(X == 0, 0.0, X), nan=123.0))


def test_fused_cg_vector_update(gctx):
    """sysml_b200_cg_vector_update against the same block written out in numpy (LinearRegCG.dml:183-195)."""
    import ctypes as C
    m, lam = 1000, 0.01
    rng = np.random.default_rng(5)
    p, q, r, beta = (rng.uniform(-1, 1, (m, 1)) for _ in range(4))
    nr2 = np.array([[float((r * r).sum())]])
    dp, dq, dr, db, dn = (gctx.dense(x) for x in (p, q, r, beta, nr2))
    gctx.call("sysml_b200_cg_vector_update", *(C.c_void_p(o.ptr) for o in (dp, dq, dr, db, dn)), m, lam)
    q2 = q + lam * p
    a = nr2[0, 0] / float((p * q2).sum())
    beta2, r2 = beta + a * p, r + a * q2
    nr2n = float((r2 * r2).sum())
    p2 = -r2 + (nr2n / nr2[0, 0]) * p
    assert_equal_matrices(q2, dq.to_host(), rel=1e-14)
    assert_equal_matrices(beta2, db.to_host(), rel=1e-12)
    assert_equal_matrices(r2, dr.to_host(), rel=1e-12)
    assert_equal_matrices(p2, dp.to_host(), rel=1e-11)
    assert abs(dn.to_host()[0, 0] - nr2n) <= 1e-12 * nr2n


@pytest.mark.parametrize("rows,cols", [(1, 1), (64, 130), (2049, 21), (1000, 1), (1, 777), (50_000, 50)])
def test_variance(gctx, rows, cols):
    """uavar / uarvar / uacvar (AggregateUnaryOpTests.var, rel 1e-9) vs the restated CM update."""
    from systemml_b200 import LibMatrixB200 as L
    X = O.java_test_matrix(rows, cols, -1, 1, 0.9, 42) + 3.0
    dX = gctx.dense(X)
    for opc, d in (("uavar", O.ALL), ("uarvar", O.ROW), ("uacvar", O.COL)):
        exp = O.agg_unary_dense(X, O.VAR, d, 1)
        got = L.unaryAggregate(gctx, opc, dX, opc)
        got = np.array([got]) if d == O.ALL else got.to_host().reshape(-1)
        assert_equal_matrices(exp, got, rel=1e-9, what=opc)
    dX.free()


@pytest.mark.parametrize("n,m", [(1000, 37), (3, 5000), (257, 129)])
def test_outer_vector_vector_cellwise(gctx, n, m):
    """n x 1 op 1 x m -> n x m (BinaryAccessType.OUTER_VECTOR_VECTOR, LibMatrixBincell.java:150-163; safeBinaryVVGeneric
    :565-589; the sorted-row-vector comparison shortcut performBinOuterOperation :814-860 gives the same cells):
    bit-exact for the arithmetic and relational value functions."""
    from systemml_b200 import LibMatrixB200 as L
    a = O.java_test_matrix(n, 1, -2, 2, 0.9, 5)
    b = np.sort(O.java_test_matrix(1, m, -2, 2, 0.9, 6), axis=1)      # sorted: the reference would take its shortcut
    da, db = gctx.dense(a), gctx.dense(b)
    for opn, code in (("+", 0), ("-", 1), ("*", 2), ("/", 3), ("<", 5), ("<=", 6), (">", 7), (">=", 8), ("==", 9), ("!=", 10),
                      ("min", 11), ("max", 12), ("1-*", 15), ("%%", 17)):
        out = L.matrixMatrixArithmetic(gctx, opn, da, db, opn)
        got = out.to_host()
        assert got.shape == (n, m)
        assert np.array_equal(O.binary_mm(a, b, code), got, equal_nan=True), opn
        out.free()
    exp_lt = (a < b).astype(np.float64)                               # independent statement of one of them
    assert np.array_equal(L.matrixMatrixArithmetic(gctx, "<", da, db, "<").to_host(), exp_lt)
    for o in (da, db):
        o.free()
