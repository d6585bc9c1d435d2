"""tsmm and dense matmult parity (DMMA kernels) vs the CPU oracle.

Shapes from the reference tests: FullMatrixMultiplicationTransposeSelfTest (3500x1500 / 7000x750, seed 7,
sparsity 0.7/0.1, eps 1e-10), gpu/MatrixMultiplicationOpTest (dims {1,128,1024}, {1,121,123,122}, 5000x50,
all four transpose combinations, rel 1e-5 there; 1e-9 here)."""
import numpy as np
import pytest

from oracle import oracle as O
from tests.util import assert_equal_matrices, assert_parity

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("rows,cols,sparsity", [(3500, 1500, 0.7), (7000, 750, 0.1), (1000, 128, 0.7), (517, 1, 0.7),
                                                (64, 129, 1.0), (2000, 1000, 1.0), (33, 7, 0.5)])
def test_tsmm_left(gctx, rows, cols, sparsity):
    from systemml_b200 import LibMatrixB200 as L
    X = O.java_test_matrix(rows, cols, 0, 1, sparsity, 7)
    exp, nnz = O.tsmm_dense(X, True, O.num_threads())
    dX = gctx.dense(X)
    out, gnnz = L.matmultTSMM(gctx, "tsmm", dX, True, with_nnz=True)
    got = out.to_host()
    assert_equal_matrices(exp, got, what="tsmm LEFT %dx%d" % (rows, cols))
    assert np.max(np.abs(exp - got)) < 1e-10 * max(1.0, rows / 3500)  # the reference test's absolute bar
    assert np.array_equal(got, got.T), "result must be exactly symmetric (upper copied to lower)"
    assert gnnz == nnz, "nnz must be bit-exact (LibMatrixMult.java:3651-3668)"
    dX.free()
    out.free()


@pytest.mark.parametrize("rows,cols", [(750, 7000), (128, 1024), (1, 300), (130, 17)])
def test_tsmm_right(gctx, rows, cols):
    from systemml_b200 import LibMatrixB200 as L
    X = O.java_test_matrix(rows, cols, 0, 1, 0.7, 7)
    exp, nnz = O.tsmm_dense(X, False, O.num_threads())
    dX = gctx.dense(X)
    out, gnnz = L.matmultTSMM(gctx, "tsmm", dX, False, with_nnz=True)
    assert_equal_matrices(exp, out.to_host(), what="tsmm RIGHT")
    assert gnnz == nnz
    dX.free()
    out.free()


def test_tsmm_zero_columns_nnz(gctx):
    from systemml_b200 import LibMatrixB200 as L
    X = O.java_test_matrix(400, 60, -1, 1, 1.0, 3)
    X[:, [5, 17, 40]] = 0.0
    exp, nnz = O.tsmm_dense(X, True, 1)
    out, gnnz = L.matmultTSMM(gctx, "tsmm", gctx.dense(X), True, with_nnz=True)
    assert gnnz == nnz == 57 * 57
    assert_equal_matrices(exp, out.to_host())


DIMS = [1, 121, 128]


@pytest.mark.parametrize("m", DIMS)
@pytest.mark.parametrize("k", [1, 123, 1024])
@pytest.mark.parametrize("n", [1, 122, 128])
@pytest.mark.parametrize("ta,tb", [(False, False), (True, False), (False, True), (True, True)])
def test_matmult_all_transposes(gctx, m, k, n, ta, tb):
    from systemml_b200 import LibMatrixB200 as L
    A = O.rand_uniform(m * k, seed=42).reshape((k, m) if ta else (m, k))
    B = O.rand_uniform(k * n, seed=43).reshape((n, k) if tb else (k, n))
    Ao = np.ascontiguousarray(A.T) if ta else A
    Bo = np.ascontiguousarray(B.T) if tb else B
    exp, _ = O.matmult_dense(Ao, Bo, 4)
    dA, dB = gctx.dense(A), gctx.dense(B)
    out = L.matmult(gctx, "ba+*", dA, dB, ta, tb)
    got = out.to_host()
    # 1e-9 element-wise (GPUTests.assertEqualMatrices); cancelling cells of these zero-mean inputs take the
    # rounding-bound exemption of tests/util.py
    assert_parity(exp, got, np.abs(Ao) @ np.abs(Bo), k, what="matmult %dx%dx%d ta=%s tb=%s" % (m, k, n, ta, tb))
    for o in (dA, dB, out):
        o.free()


@pytest.mark.parametrize("m,k,n", [(5000, 50, 20), (5000, 50, 1), (20000, 500, 20), (30000, 100, 50), (257, 1000, 300),
                                   (1000, 33, 7), (64, 4100, 65)])
def test_matmult_ml_shapes(gctx, m, k, n):
    """'common ML' shapes of MatrixMultiplicationOpTest plus the skinny right-hand sides of C4 / C5."""
    from systemml_b200 import LibMatrixB200 as L
    A = O.rand_uniform(m * k, seed=1, lo=0.0, hi=1.0).reshape(m, k)
    B = O.rand_uniform(k * n, seed=2, lo=0.0, hi=1.0).reshape(k, n)
    exp, _ = O.matmult_dense(A, B, O.num_threads())
    out = L.matmult(gctx, "ba+*", gctx.dense(A), gctx.dense(B))
    assert_equal_matrices(exp, out.to_host(), what="matmult %dx%dx%d" % (m, k, n))
    # t(A) %*% Y with a tall contraction (split-K path)
    Y = O.rand_uniform(m * n, seed=3, lo=0.0, hi=1.0).reshape(m, n)
    exp2, _ = O.matmult_dense(np.ascontiguousarray(A.T), Y, O.num_threads())
    out2 = L.matmult(gctx, "ba+*", gctx.dense(A), gctx.dense(Y), True, False)
    assert_equal_matrices(exp2, out2.to_host(), what="t(A)Y %dx%dx%d" % (m, k, n))


def test_matmult_dimension_error(gctx):
    from systemml_b200 import LibMatrixB200 as L, native
    with pytest.raises(native.NativeError) as ei:
        L.matmult(gctx, "ba+*", gctx.dense(np.ones((3, 4))), gctx.dense(np.ones((5, 2))))
    assert ei.value.code == native.E_DIM


def test_tsmm_tall_split_k_determinism(gctx):
    from systemml_b200 import LibMatrixB200 as L
    X = gctx.rand(400_000, 256, seed=8)
    a = L.matmultTSMM(gctx, "tsmm", X, True).to_host()
    b = L.matmultTSMM(gctx, "tsmm", X, True).to_host()
    assert np.array_equal(a, b)
    # oracle on a row subsample is not comparable; check against the trace identity sum(diag) = sum(X^2)
    ssq = L.unaryAggregate(gctx, "uasqk+", X, "uasqk+")
    assert abs(np.trace(a) - ssq) <= 1e-11 * abs(ssq)
    X.free()
