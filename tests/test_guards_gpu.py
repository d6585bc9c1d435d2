"""Out-of-bounds WRITE detection with guard zones (compute-sanitizer is closed on this GPU pool): every
output buffer is embedded in a larger allocation whose borders hold a sentinel; after the call the borders
must be untouched and every output cell must have been written (no sentinel left inside)."""
import ctypes as C

import numpy as np
import pytest

from oracle import oracle as O

pytestmark = pytest.mark.gpu

GUARD = 64
SENT = -7.25e300


class Guarded:
    def __init__(self, gctx, n):
        self.gctx, self.n = gctx, n
        self.buf = gctx.dense(np.full(n + 2 * GUARD, SENT))
        self.ptr = self.buf.ptr + 8 * GUARD

    def check(self, what):
        h = self.buf.to_host().reshape(-1)
        assert np.all(h[:GUARD] == SENT) and np.all(h[GUARD + self.n:] == SENT), "%s wrote outside its output" % what
        assert not np.any(h[GUARD:GUARD + self.n] == SENT), "%s left output cells unwritten" % what
        self.buf.free()
        return h[GUARD:GUARD + self.n]


def p(x):
    return C.c_void_p(x)


@pytest.mark.parametrize("n,m", [(1, 1), (3, 7), (101, 63), (257, 129), (999, 1000), (130, 1025), (65, 4096), (9, 5003)])
def test_mmchain_mv_vm_guards(gctx, n, m):
    X, v, w = gctx.rand(n, m, seed=1), gctx.rand(m, 1, seed=2), gctx.rand(n, 1, seed=3)
    for ct in (0, 1, 2):
        o = Guarded(gctx, m)
        gctx.call("sysml_b200_mmchain_dense", p(X.ptr), p(v.ptr), p(w.ptr) if ct else None, p(o.ptr), n, m, ct)
        got = o.check("mmchain")
        exp = O.mmchain_dense(X.to_host(), v.to_host(), w.to_host() if ct else None, ct)
        assert np.max(np.abs(got - exp)) <= 1e-9 * max(np.max(np.abs(exp)), 1e-300)
    o = Guarded(gctx, n)
    gctx.call("sysml_b200_matmult_dense", p(X.ptr), p(v.ptr), p(o.ptr), n, m, m, 1, 0, 0)
    o.check("MV")
    o = Guarded(gctx, m)
    gctx.call("sysml_b200_matmult_dense", p(X.ptr), p(w.ptr), p(o.ptr), n, m, n, 1, 1, 0)
    o.check("VM")


@pytest.mark.parametrize("rows,cols", [(1, 1), (5, 3), (100, 50), (2049, 21), (33, 1000), (70000, 50), (1000, 129), (64, 2049)])
def test_aggregate_guards(gctx, rows, cols):
    X = gctx.rand(rows, cols, seed=4)
    for op in (0, 1, 2, 3, 6, 7, 8):
        for d, n in ((0, 1), (1, rows), (2, cols)):
            o = Guarded(gctx, n)
            gctx.call("sysml_b200_agg_unary_dense", p(X.ptr), rows, cols, op, d, p(o.ptr))
            o.check("agg op=%d dir=%d" % (op, d))
    for op in (4, 5):
        o = Guarded(gctx, rows)
        gctx.call("sysml_b200_agg_unary_dense", p(X.ptr), rows, cols, op, 1, p(o.ptr))
        o.check("rowIndex")


@pytest.mark.parametrize("M,K,N", [(1, 5, 3), (130, 17, 9), (257, 100, 50), (64, 300, 130), (129, 33, 65), (1000, 64, 20), (50, 2000, 100)])
def test_gemm_and_tsmm_guards(gctx, M, K, N):
    A, B = gctx.rand(M, K, seed=5), gctx.rand(K, N, seed=6)
    At, Bt = gctx.rand(K, M, seed=7), gctx.rand(N, K, seed=8)
    for (a, b, ta, tb, ar, ac, br, bc) in ((A, B, 0, 0, M, K, K, N), (At, B, 1, 0, K, M, K, N), (A, Bt, 0, 1, M, K, N, K), (At, Bt, 1, 1, K, M, N, K)):
        o = Guarded(gctx, M * N)
        gctx.call("sysml_b200_matmult_dense", p(a.ptr), p(b.ptr), p(o.ptr), ar, ac, br, bc, ta, tb)
        o.check("matmult ta=%d tb=%d" % (ta, tb))
    for left, dim in ((1, K), (0, M)):
        o = Guarded(gctx, dim * dim)
        gctx.call("sysml_b200_tsmm_dense", p(A.ptr), p(o.ptr), M, K, left, None)
        got = o.check("tsmm left=%d" % left).reshape(dim, dim)
        assert np.array_equal(got, got.T)


@pytest.mark.parametrize("rows,cols", [(1, 1), (7, 3), (1025, 33), (33, 1025)])
def test_cellwise_guards(gctx, rows, cols):
    A, B = gctx.rand(rows, cols, seed=9), gctx.rand(rows, cols, seed=10)
    cv, rv = gctx.rand(rows, 1, seed=11), gctx.rand(1, cols, seed=12)
    for bc, b in ((0, B), (1, cv), (2, rv)):
        for op in (0, 3, 6, 17):
            o = Guarded(gctx, rows * cols)
            gctx.call("sysml_b200_binary_mm", p(A.ptr), p(b.ptr), p(o.ptr), rows, cols, bc, op)
            o.check("binary_mm")
        o = Guarded(gctx, rows * cols)
        gctx.call("sysml_b200_axpy", p(A.ptr), p(b.ptr), 0.5, p(o.ptr), rows, cols, bc)
        o.check("axpy")
    for fn, args in (("sysml_b200_binary_ms", (-2.0,)),):
        o = Guarded(gctx, rows * cols)
        gctx.call("sysml_b200_binary_ms", p(A.ptr), -2.0, p(o.ptr), rows * cols, 2, 0)
        o.check("binary_ms")
    o = Guarded(gctx, rows * cols)
    gctx.call("sysml_b200_unary", p(A.ptr), p(o.ptr), rows * cols, 0)
    o.check("unary")
    o = Guarded(gctx, rows * cols)
    gctx.call("sysml_b200_transpose", p(A.ptr), p(o.ptr), rows, cols)
    o.check("transpose")


def test_sparse_guards(gctx):
    X = O.java_test_matrix(700, 130, -1, 1, 0.07, 3)
    rp, ci, va = O.dense_to_csr(X)
    S = gctx.csr(rp, ci, va, X.shape)
    v, y = gctx.rand(130, 1, seed=1), gctx.rand(700, 1, seed=2)
    B = gctx.rand(130, 33, seed=3)
    o = Guarded(gctx, 700)
    gctx.call("sysml_b200_matmult_csr_dense", p(S.csr.rowPtr), p(S.csr.colInd), p(S.csr.val), p(v.ptr), p(o.ptr), 700, 130, 130, 1, 0)
    o.check("spmv")
    o = Guarded(gctx, 130)
    gctx.call("sysml_b200_matmult_csr_dense", p(S.csr.rowPtr), p(S.csr.colInd), p(S.csr.val), p(y.ptr), p(o.ptr), 700, 130, 700, 1, 1)
    o.check("t(csr) y")
    o = Guarded(gctx, 700 * 33)
    gctx.call("sysml_b200_matmult_csr_dense", p(S.csr.rowPtr), p(S.csr.colInd), p(S.csr.val), p(B.ptr), p(o.ptr), 700, 130, 130, 33, 0)
    o.check("spmm")
    o = Guarded(gctx, 130)
    gctx.call("sysml_b200_mmchain_csr", p(S.csr.rowPtr), p(S.csr.colInd), p(S.csr.val), p(v.ptr), p(y.ptr), p(o.ptr), 700, 130, 1)
    o.check("csr mmchain")
    for op, d, n in ((0, 0, 1), (0, 1, 700), (0, 2, 130), (2, 1, 700), (3, 2, 130), (4, 1, 700)):
        o = Guarded(gctx, n)
        gctx.call("sysml_b200_agg_unary_csr", p(S.csr.rowPtr), p(S.csr.colInd), p(S.csr.val), 700, 130, op, d, p(o.ptr))
        o.check("csr agg")
