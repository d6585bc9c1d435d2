"""CSR kernels vs the CPU oracle: SpMV / SpMM / t(X)%*%y, sparse mmchain, sparse tsmm, sparse aggregates."""
import numpy as np
import pytest

from oracle import oracle as O
from tests.util import abs_terms_mmchain, assert_equal_matrices, assert_parity

pytestmark = pytest.mark.gpu


def _csr(gctx, X):
    rp, ci, va = O.dense_to_csr(X)
    return gctx.csr(rp, ci, va, X.shape), (rp, ci, va)


@pytest.mark.parametrize("rows,cols,sp", [(3471, 567, 0.1), (3471, 567, 0.7), (20000, 3000, 0.002), (1000, 10000, 0.001),
                                          (50, 40, 0.5), (1, 100, 0.3), (5000, 200, 0.0)])
@pytest.mark.parametrize("ct", [0, 1, 2])
def test_sparse_mmchain(gctx, rows, cols, sp, ct):
    from systemml_b200 import LibMatrixB200 as L
    X = O.java_test_matrix(rows, cols, 0, 1, sp, 7)
    v = O.java_test_matrix(cols, 1, 0, 1, 0.7, 3)
    w = O.java_test_matrix(rows, 1, 0, 1, max(sp, 0.1), 10)
    dX, (rp, ci, va) = _csr(gctx, X)
    exp = O.mmchain_csr(rp, ci, va, X.shape, v, w if ct else None, ct, 4)
    out = L.mmchain(gctx, "mmchain", dX, gctx.dense(v), gctx.dense(w) if ct else None, ct)
    got = out.to_host().reshape(-1)
    scale = max(np.max(np.abs(exp)), 1e-300)
    assert np.max(np.abs(exp - got)) <= 1e-9 * scale
    if ct != 2:
        assert_equal_matrices(exp, got, what="sparse mmchain")


@pytest.mark.parametrize("rows,cols,sp,n", [(3471, 567, 0.1, 1), (3471, 567, 0.7, 1), (20000, 3000, 0.002, 1), (20000, 3000, 0.002, 20),
                                            (999, 10000, 0.001, 1), (2000, 500, 0.05, 64), (2000, 500, 0.05, 130), (300, 80, 0.3, 7)])
def test_spmv_spmm(gctx, rows, cols, sp, n):
    from systemml_b200 import LibMatrixB200 as L
    X = O.java_test_matrix(rows, cols, 0, 1, sp, 7)
    B = O.java_test_matrix(cols, n, 0, 1, 0.9, 5)
    dX, (rp, ci, va) = _csr(gctx, X)
    exp, _ = O.matmult_csr_dense(rp, ci, va, X.shape, B, 4)
    got = L.matmult(gctx, "ba+*", dX, gctx.dense(B)).to_host()
    assert_equal_matrices(exp, got, what="CSR x dense n=%d" % n)
    # t(X) %*% Y (scatter form)
    Y = O.java_test_matrix(rows, n, 0, 1, 0.9, 6)
    exp2 = (X.astype(np.longdouble).T @ Y.astype(np.longdouble)).astype(np.float64)
    got2 = L.matmult(gctx, "ba+*", dX, gctx.dense(Y), isLeftTransposed=True).to_host()
    assert_equal_matrices(exp2, got2, rel=1e-9, what="t(CSR) x dense")


def test_sparse_tsmm(gctx):
    from systemml_b200 import LibMatrixB200 as L
    X = O.java_test_matrix(3500, 300, 0, 1, 0.1, 7)
    dX, (rp, ci, va) = _csr(gctx, X)
    exp, nnz = O.tsmm_csr_left(rp, ci, va, X.shape, 4)
    out, gnnz = L.matmultTSMM(gctx, "tsmm", dX, True, with_nnz=True)
    assert_equal_matrices(exp, out.to_host(), what="sparse tsmm")
    assert gnnz == nnz


@pytest.mark.parametrize("rows,cols,sp", [(1000, 50, 0.3), (2049, 130, 0.03), (500, 1000, 0.9), (64, 64, 0.0003)])
def test_sparse_aggregates(gctx, rows, cols, sp):
    from systemml_b200 import LibMatrixB200 as L
    X = O.java_test_matrix(rows, cols, -1, 1, sp, 42)
    if not X.any():
        X[1, 2] = -0.25
    dX, (rp, ci, va) = _csr(gctx, X)
    for opc, op, d, exact in (("uak+", O.SUM, O.ALL, False), ("uark+", O.SUM, O.ROW, False), ("uack+", O.SUM, O.COL, False),
                              ("uasqk+", O.SUMSQ, O.ALL, False), ("uamax", O.MAX, O.ALL, True), ("uarmax", O.MAX, O.ROW, True),
                              ("uacmax", O.MAX, O.COL, True), ("uamin", O.MIN, O.ALL, True), ("uarmin", O.MIN, O.ROW, True),
                              ("uacmin", O.MIN, O.COL, True), ("uarimax", O.MAXINDEX, O.ROW, True), ("uarimin", O.MININDEX, O.ROW, True)):
        exp = O.agg_unary_csr(rp, ci, va, X.shape, op, d, 1)
        got = L.unaryAggregate(gctx, opc, dX, opc)
        got = np.array([got]) if d == O.ALL else got.to_host().reshape(-1)
        if exact:
            assert np.array_equal(exp, got), opc
        else:
            assert np.max(np.abs(exp - got)) <= 1e-12 * max(1.0, np.max(np.abs(exp))), opc


def test_sparse_rowindex_edge_cases(gctx):
    from systemml_b200 import LibMatrixB200 as L
    X = np.array([[1.0, 3.0, 3.0, 2.0], [-1.0, -2.0, -1.0, -3.0], [0.0, 0.0, 0.0, 0.0], [-5.0, 0.0, -2.0, 0.0],
                  [0.0, -1.0, 0.0, -2.0], [0.0, 0.0, -1.0, -1.0], [2.0, 0.0, 0.0, 0.0]])
    dX, (rp, ci, va) = _csr(gctx, X)
    for opc, op in (("uarimax", O.MAXINDEX), ("uarimin", O.MININDEX)):
        exp = O.agg_unary_csr(rp, ci, va, X.shape, op, O.ROW)
        assert np.array_equal(exp, O.agg_unary_dense(X, op, O.ROW)), "oracle: sparse and dense must agree"
        assert L.unaryAggregate(gctx, opc, dX, opc).to_host().reshape(-1).tolist() == exp.tolist()


@pytest.mark.parametrize("m,k,n,spa,spb", [(300, 500, 200, 0.1, 0.05), (1, 400, 300, 1.0, 0.1), (700, 50, 1200, 0.02, 0.3), (64, 64, 64, 0.5, 0.0)])
def test_remaining_sparsity_combinations(gctx, m, k, n, spa, spb):
    """dense x sparse, sparse x sparse (FullMatrixMultiplicationTest's sparsity grid) and t(dense) x sparse."""
    from systemml_b200 import LibMatrixB200 as L
    A = O.java_test_matrix(m, k, -1, 1, spa, 11)
    B = O.java_test_matrix(k, n, -1, 1, spb, 12)
    exp = (A.astype(np.longdouble) @ B.astype(np.longdouble)).astype(np.float64)
    dB, _ = _csr(gctx, B)
    scale = max(np.max(np.abs(exp)), 1e-300)
    got = L.matmult(gctx, "ba+*", gctx.dense(A), dB).to_host()            # dense x CSR
    assert np.max(np.abs(exp - got)) <= 1e-12 * scale
    dA, _ = _csr(gctx, A)
    got = L.matmult(gctx, "ba+*", dA, dB).to_host()                        # CSR x CSR
    assert np.max(np.abs(exp - got)) <= 1e-12 * scale
    At = np.ascontiguousarray(A.T)
    got = L.matmult(gctx, "ba+*", gctx.dense(At), dB, isLeftTransposed=True).to_host()  # t(dense) x CSR
    assert np.max(np.abs(exp - got)) <= 1e-12 * scale


def test_sparse_tsmm_right(gctx):
    from systemml_b200 import LibMatrixB200 as L
    X = O.java_test_matrix(300, 2000, 0, 1, 0.1, 7)
    dX, _ = _csr(gctx, X)
    exp, nnz = O.tsmm_dense(X, False, 4)
    out, gnnz = L.matmultTSMM(gctx, "tsmm", dX, False, with_nnz=True)
    assert_equal_matrices(exp, out.to_host(), what="sparse tsmm RIGHT")
    assert gnnz == nnz


def _big_csr(rows, cols, mean, seed, full_rows=(), empty_every=0):
    """CSR arrays built directly (no dense detour): ~Poisson(mean) entries per row, sorted distinct columns."""
    rng = np.random.default_rng(seed)
    cnt = np.minimum(rng.poisson(mean, rows), cols).astype(np.int64)
    if empty_every:
        cnt[::empty_every] = 0
    for r in full_rows:
        cnt[r] = cols
    rp = np.zeros(rows + 1, dtype=np.int64)  # the oracle takes int64 row pointers; the device copy is int32
    rp[1:] = np.cumsum(cnt)
    ci = np.empty(rp[-1], dtype=np.int32)
    for r in range(rows):
        c = cnt[r]
        if c:
            ci[rp[r]:rp[r + 1]] = np.sort(rng.choice(cols, size=c, replace=False)) if c < cols else np.arange(cols)
    va = rng.uniform(-1, 1, rp[-1])
    va[va == 0] = 0.5
    return rp, ci, va


@pytest.mark.parametrize("rows,cols,mean,full_rows,empty_every", [
    (40001, 3000, 30, (), 0),                       # ragged last tile
    (40960, 3000, 30, (5, 20000, 40959), 7),        # tiles that overflow a stage + empty rows
    (150000, 10000, 10, (), 0),                     # the L2SVM shape class: 10 per row
    (400000, 500, 3, (), 0),                        # very short rows: one lane per row
])
def test_csr_streamed_rows(gctx, rows, cols, mean, full_rows, empty_every, monkeypatch):
    """The bulk-copy streamed CSR kernel (rows >= 128 x #SMs, nnz >= 2^20) for SpMV, t(X)%*%y and the three
    sparse mmchain types, against the oracle and against the row-at-a-time kernel."""
    from systemml_b200 import LibMatrixB200 as L
    rp, ci, va = _big_csr(rows, cols, mean, 11, full_rows, empty_every)
    assert rp[-1] >= (1 << 20) or rows * mean < (1 << 20)
    shape = (rows, cols)
    dX = gctx.csr(rp, ci, va, shape)
    rng = np.random.default_rng(5)
    v = rng.uniform(-1, 1, (cols, 1))
    y = rng.uniform(-1, 1, (rows, 1))
    y[::5] = 0.0  # zero weights are skipped (LibMatrixMult.java:1667)
    exp, _ = O.matmult_csr_dense(rp, ci, va, shape, v, 4)
    got = L.matmult(gctx, "ba+*", dX, gctx.dense(v)).to_host()
    assert_equal_matrices(exp, got, what="streamed SpMV")
    monkeypatch.setenv("SYSML_B200_CSR_NO_STREAM", "1")
    got_row = L.matmult(gctx, "ba+*", dX, gctx.dense(v)).to_host()
    monkeypatch.delenv("SYSML_B200_CSR_NO_STREAM")
    assert_equal_matrices(got_row, got, rel=1e-9, what="streamed vs row kernel")
    for ct in (0, 1, 2):
        e = O.mmchain_csr(rp, ci, va, shape, v, y if ct else None, ct, 4)
        g = L.mmchain(gctx, "mmchain", dX, gctx.dense(v), gctx.dense(y) if ct else None, ct).to_host().reshape(-1)
        assert np.max(np.abs(e - g)) <= 1e-9 * max(np.max(np.abs(e)), 1e-300), ct
    # row aggregates on the streamed kernel (colind is not even read): sums to a few ulp, max / min exact
    for opc, op, exact in (("uark+", O.SUM, False), ("uarsqk+", O.SUMSQ, False), ("uarmax", O.MAX, True), ("uarmin", O.MIN, True)):
        e = O.agg_unary_csr(rp, ci, va, shape, op, O.ROW, 1)
        gr = L.unaryAggregate(gctx, opc, dX, opc).to_host().reshape(-1)
        if exact:
            assert np.array_equal(e, gr), opc
        else:
            assert np.max(np.abs(e - gr)) <= 1e-12 * max(1.0, np.max(np.abs(e))), opc
    # t(X) %*% y: exact column sums in extended precision
    e2 = np.zeros(cols, dtype=np.longdouble)
    np.add.at(e2, ci, va.astype(np.longdouble) * np.repeat(y.reshape(-1), np.diff(rp)).astype(np.longdouble))
    g2 = L.matmult(gctx, "ba+*", dX, gctx.dense(y), isLeftTransposed=True).to_host().reshape(-1)
    assert np.max(np.abs(e2.astype(np.float64) - g2)) <= 1e-9 * max(float(np.max(np.abs(e2))), 1e-300)


def _csr_to_dense(rp, ci, va, shape):
    D = np.zeros(shape)
    for r in range(shape[0]):
        D[r, ci[rp[r]:rp[r + 1]]] = va[rp[r]:rp[r + 1]]
    return D


def _check_csr(got, exp_dense, exact=True):
    """CSR invariants (sorted distinct columns, no explicit zeros, nnz = count of non-zeros) + values."""
    rp, ci, va = got.to_host()
    rows, cols = exp_dense.shape
    assert rp[0] == 0 and rp[-1] == len(ci) == len(va) == got.csr.nnz
    assert np.all(np.diff(rp) >= 0)
    for r in range(rows):
        seg = ci[rp[r]:rp[r + 1]]
        assert np.all(np.diff(seg) > 0) and (len(seg) == 0 or (seg[0] >= 0 and seg[-1] < cols))
    assert np.all(va != 0.0)
    assert got.csr.nnz == int(np.count_nonzero(exp_dense))
    D = _csr_to_dense(rp, ci, va, exp_dense.shape)
    if exact:
        assert np.array_equal(D, exp_dense)
    else:
        assert_equal_matrices(exp_dense, D, rel=1e-12, what="sparse result")


@pytest.mark.parametrize("rows,cols,sp", [(300, 200, 0.1), (1, 500, 0.3), (700, 3, 0.5), (64, 64, 0.0), (2049, 1300, 0.01)])
def test_sparse_transpose_and_add(gctx, rows, cols, sp):
    """CSR results: t(X) (csr2csc) and alpha*A + beta*B (csrgeam), incl. cancellation to exact zeros."""
    from systemml_b200 import LibMatrixB200 as L
    A = O.java_test_matrix(rows, cols, -1, 1, sp, 21)
    B = O.java_test_matrix(rows, cols, -1, 1, sp, 22)
    if sp > 0:
        B[::3] = A[::3]  # A - B cancels exactly on these rows
    dA, _ = _csr(gctx, A)
    dB, _ = _csr(gctx, B)
    _check_csr(L.sparseTranspose(gctx, "r'", dA), np.ascontiguousarray(A.T))
    _check_csr(L.sparseSparseAdd(gctx, "+", dA, dB), A + B)
    _check_csr(L.sparseSparseAdd(gctx, "-", dA, dB, 1.0, -1.0), A - B)
    _check_csr(L.sparseSparseAdd(gctx, "+*", dA, dB, 1.0, 0.37), A + 0.37 * B)


def _ref_sparse_sparse(A, B):
    """LibMatrixMult.matrixMultSparseSparse (LibMatrixMult.java:1397-1500): per row of A, for k ascending,
    c[i, cols(B[k,:])] += a * vals(B[k,:]) -- products and sums separately rounded."""
    m, n = A.shape[0], B.shape[1]
    Cm = np.zeros((m, n))
    bnz = [np.nonzero(B[k])[0] for k in range(B.shape[0])]
    for i in range(m):
        for k in np.nonzero(A[i])[0]:
            j = bnz[k]
            if len(j):
                Cm[i, j] += A[i, k] * B[k, j]
    return Cm


@pytest.mark.parametrize("m,k,n,spa,spb", [(300, 500, 200, 0.1, 0.05), (1, 400, 300, 1.0, 0.1), (700, 50, 1200, 0.02, 0.3),
                                           (64, 64, 64, 0.5, 0.0), (1500, 900, 700, 0.004, 0.004)])
def test_sparse_sparse_matmult_sparse_output(gctx, m, k, n, spa, spb):
    from systemml_b200 import LibMatrixB200 as L
    A = O.java_test_matrix(m, k, -1, 1, spa, 11)
    B = O.java_test_matrix(k, n, -1, 1, spb, 12)
    dA, _ = _csr(gctx, A)
    dB, _ = _csr(gctx, B)
    got = L.sparseSparseMatmult(gctx, "ba+*", dA, dB)
    _check_csr(got, _ref_sparse_sparse(A, B))  # bit-exact: same products, same summation order
    # and it agrees with the dense-output entry point
    dense = L.matmult(gctx, "ba+*", dA, dB).to_host()
    rp, ci, va = got.to_host()
    assert np.max(np.abs(dense - _csr_to_dense(rp, ci, va, (m, n)))) <= 1e-12 * max(np.max(np.abs(dense)), 1e-300)  # other summation order
    # integer-valued operands: products that cancel exactly leave no entry behind
    Ai = np.round(O.java_test_matrix(40, 30, -2, 2, 0.5, 5))
    Bi = np.round(O.java_test_matrix(30, 50, -2, 2, 0.5, 6))
    dAi, _ = _csr(gctx, Ai)
    dBi, _ = _csr(gctx, Bi)
    _check_csr(L.sparseSparseMatmult(gctx, "ba+*", dAi, dBi), Ai @ Bi)


def test_sparse_output_edge_cases(gctx):
    """Empty operands, empty results and the product-count limit of the CSR-result operations."""
    from systemml_b200 import LibMatrixB200 as L, native
    Z = np.zeros((40, 30))
    A = O.java_test_matrix(40, 30, -1, 1, 0.2, 3)
    dZ, _ = _csr(gctx, Z)
    dA, _ = _csr(gctx, A)
    _check_csr(L.sparseTranspose(gctx, "r'", dZ), Z.T.copy())
    _check_csr(L.sparseSparseAdd(gctx, "+", dZ, dZ), Z)
    _check_csr(L.sparseSparseAdd(gctx, "+", dA, dZ), A)
    _check_csr(L.sparseSparseAdd(gctx, "-", dA, dA, 1.0, -1.0), Z)          # everything cancels: nnz 0, no arrays
    dAt = L.sparseTranspose(gctx, "r'", dA)
    _check_csr(L.sparseSparseMatmult(gctx, "ba+*", dZ, dAt), np.zeros((40, 40)))
    _check_csr(L.sparseSparseMatmult(gctx, "ba+*", dA, dAt), _ref_sparse_sparse(A, np.ascontiguousarray(A.T)))
    with pytest.raises(native.NativeError):                                     # inner dimensions
        L.sparseSparseMatmult(gctx, "ba+*", dA, dA)
    with pytest.raises(native.NativeError):
        L.sparseSparseAdd(gctx, "+", dA, dAt)
    # t(t(A)) == A exactly
    rp, ci, va = L.sparseTranspose(gctx, "r'", dAt).to_host()
    rp0, ci0, va0 = dA.to_host()
    assert np.array_equal(rp, rp0) and np.array_equal(ci, ci0) and np.array_equal(va, va0)


def test_device_csr_generator_matches_cpu_twin(gctx):
    """sysml_b200_rand_csr (BASELINE config C3's shape: ~Poisson(mean) entries per row) against its CPU twin, bit for bit
    -- structure (int32, exact) and values -- including a row shard taken with row_offset."""
    for rows, cols, mean, seed, off in ((20_000, 10_000, 10, 2, 0), (5_000, 10_000, 10, 2, 7_000), (3_000, 64, 3.5, 9, 0), (100, 5, 40, 1, 0)):
        o = gctx.rand_csr(rows, cols, mean, seed, row_offset=off)
        rp, ci, va = o.to_host()
        erp, eci, eva = O.rand_csr(rows, cols, mean, seed, row_offset=off)
        assert np.array_equal(rp.astype(np.int64), erp), (rows, cols, mean)
        assert np.array_equal(ci, eci) and np.array_equal(va, eva)
        assert o.csr.nnz == int(erp[-1])
        for r in range(0, rows, max(rows // 50, 1)):        # sorted, distinct, in range
            c = ci[rp[r]:rp[r + 1]]
            assert np.all(np.diff(c) > 0) and (c.size == 0 or (c[0] >= 0 and c[-1] < cols))
        o.free()


@pytest.mark.parametrize("n,m,sp,panel", [(5000, 300, 0.03, 512), (4097, 10_000, 0.001, 1024), (1000, 17, 0.4, 64), (70_000, 1000, 0.01, 8192),
                                          (300, 2000, 0.05, 8192), (2049, 64, 1.0, 256), (100, 33, 0.0, 32), (9000, 40, 0.3, 8192)])
def test_panelled_sliced_ell_vm_and_mmchain(gctx, n, m, sp, panel):
    """t(X) %*% y and the three chain types through the sliced-ELL column copy (csrc/psell.cu) against the oracle's
    matrixMultChainSparse / VM (LibMatrixMult.java:1655-1684, 1262-1286): 1e-9 element-wise, the same bits on every run.
    Shapes cover several panels, a ragged last panel, empty slices, nnz = 0, and columns far longer than a piece (a column
    of ones, fully dense blocks): cut columns share a group, whose adds are then serialised."""
    import ctypes as C
    from systemml_b200.gpu_context import PSELLPointer
    X = O.java_test_matrix(n, m, -1, 1, sp, 77) if sp > 0 else np.zeros((n, m))
    if sp >= 1.0:
        X[X == 0] = 0.5
    X[:, m // 2] = np.where(np.arange(n) % 3 == 0, 0.0, 1.5) if sp > 0 else 0.0     # one long column run
    if sp > 0:
        X[:, 0] = 1.0                                                                # an intercept column
    rp, ci, va = O.dense_to_csr(X)
    d = gctx.csr(rp, ci, va, X.shape)
    pc = PSELLPointer(gctx, d.csr, n, m, panel)
    assert pc.slots >= d.csr.nnz and pc.slots % 32 == 0
    y = O.rand_uniform(n, seed=5)
    v = O.rand_uniform(m, seed=6)
    w = O.rand_uniform(n, seed=7, lo=0.0, hi=1.0)
    dy, dv, dw = gctx.dense(y.reshape(n, 1)), gctx.dense(v.reshape(m, 1)), gctx.dense(w.reshape(n, 1))
    out, tmp = gctx.empty(m, 1), gctx.empty(n, 1)
    runs = []
    for _ in range(2):
        pc.vm(dy.ptr, out.ptr)
        runs.append(out.to_host().reshape(-1))
    assert np.array_equal(runs[0], runs[1]), "t(X) %*% y must be bit-identical run to run"
    exp = (X.astype(np.longdouble).T @ y.astype(np.longdouble)).astype(np.float64)
    assert_parity(exp, runs[0], np.abs(X).T @ np.abs(y), n, what="psell t(X)y %dx%d" % (n, m))
    for wop, yy in ((1, y * w), (2, y - w)):
        pc.vm(dy.ptr, out.ptr, dw.ptr, wop)
        e2 = (X.astype(np.longdouble).T @ yy.astype(np.longdouble)).astype(np.float64)
        assert_parity(e2, out.to_host().reshape(-1), np.abs(X).T @ np.abs(yy), n, what="psell t(X)(y op w) %dx%d" % (n, m))
    for ct in (0, 1, 2):
        pc.mmchain(d.csr, dv.ptr, dw.ptr if ct else None, tmp.ptr, out.ptr, ct)
        got = out.to_host().reshape(-1)
        e = O.mmchain_csr(rp, ci, va, X.shape, v, w if ct else None, ct, 1)
        assert_parity(e, got, abs_terms_mmchain(X, v, w, ct), n + m, what="psell mmchain %dx%d ct=%d" % (n, m, ct))
    # non-finite y in a row no column stores must not leak through the padded slots (they gather y_s[R] = 0)
    if sp > 0 and sp < 1.0:
        X2 = X.copy()
        X2[n // 2, :] = 0.0
        rp2, ci2, va2 = O.dense_to_csr(X2)
        d2 = gctx.csr(rp2, ci2, va2, X2.shape)
        pc2 = PSELLPointer(gctx, d2.csr, n, m, panel)
        y2 = y.copy()
        y2[n // 2] = np.inf
        dy2 = gctx.dense(y2.reshape(n, 1))
        pc2.vm(dy2.ptr, out.ptr)
        assert np.all(np.isfinite(out.to_host()))
        pc2.deallocate()
        d2.free()
        dy2.free()
    pc.deallocate()
    for o in (d, dy, dv, dw, out, tmp):
        o.free()


def test_gpuobject_builds_the_column_ordered_copy_on_second_use(gctx):
    """LibMatrixB200.matmult / mmchain: the first transposed use of a CSR operand scatters (one-shot operands never pay
    for the sort), the second builds the copy, later ones gather; all agree to 1e-9 and from then on bit for bit."""
    from systemml_b200 import LibMatrixB200 as L
    X = O.java_test_matrix(20_000, 500, -1, 1, 0.02, 3)
    rp, ci, va = O.dense_to_csr(X)
    d = gctx.csr(rp, ci, va, X.shape)
    y = gctx.dense(O.rand_uniform(20_000, seed=9).reshape(-1, 1))
    r = []
    for i in range(4):
        o = L.matmult(gctx, "ba+*", d, y, isLeftTransposed=True)
        r.append(o.to_host().reshape(-1))
        o.free()
        assert (d.colcopy is not None) == (i >= 1)
    assert_equal_matrices(r[0], r[1], what="scatter vs gather")
    assert np.array_equal(r[1], r[2]) and np.array_equal(r[2], r[3])
    v = gctx.dense(O.rand_uniform(500, seed=10).reshape(-1, 1))
    q = L.mmchain(gctx, "mmchain", d, v, None, "XtXv")
    assert_equal_matrices(O.mmchain_csr(rp, ci, va, X.shape, O.rand_uniform(500, seed=10), None, 0, 1), q.to_host().reshape(-1))
    d.free()
    assert d.colcopy is None


@pytest.mark.parametrize("m,k,n", [(5000, 700, 300), (40_000, 64, 33), (1, 50, 1000), (300, 300, 5)])
def test_ultra_sparse_left_matmult_sparse_result(gctx, m, k, n):
    """A selection / permutation-like left operand (a handful of cells) times a dense matrix, CSR result, against the oracle's
    LibMatrixMult.matrixMultUltraSparse (left branch, LibMatrixMult.java:1510-1566): structure and values bit for bit --
    single-cell rows select and scale a row of B, longer rows add their products in ascending k -- and the same cells as
    the general CSR x dense kernel."""
    from systemml_b200 import LibMatrixB200 as L, native
    rng = np.random.default_rng(5)
    A = np.zeros((m, k))
    rows = rng.choice(m, size=min(m, 12), replace=False)
    for i, r in enumerate(rows):
        cols = rng.choice(k, size=1 if i % 3 else min(k, 4), replace=False)
        A[r, cols] = np.where(i % 2, 1.0, rng.uniform(-2, 2, size=len(cols)))
    B = O.java_test_matrix(k, n, -1, 1, 0.6, 9)
    B[:, n // 2] = 0.0
    a = O.dense_to_csr(A)
    dA, dB = gctx.csr(a[0].astype(np.int32), a[1], a[2], A.shape), gctx.dense(B)
    got = L.ultraSparseMatmult(gctx, "ba+*", dA, dB)
    ref = O.matmult_ultra_sparse_left(a, B)
    rp, ci, va = got.to_host()
    assert got.csr.nnz == ref[0][-1] and np.array_equal(np.asarray(rp, dtype=np.int64), ref[0])
    assert np.array_equal(ci[:got.csr.nnz], ref[1]) and np.array_equal(va[:got.csr.nnz], ref[2])
    dense = L.matmult(gctx, "ba+*", dA, dB).to_host()
    assert_equal_matrices(dense, _csr_to_dense(rp, ci, va, (m, n)), what="ultra-sparse vs general kernel")
    # empty left operand, and a dimension error
    dZ = gctx.csr(np.zeros(m + 1, dtype=np.int32), np.zeros(0, dtype=np.int32), np.zeros(0), A.shape)
    assert L.ultraSparseMatmult(gctx, "ba+*", dZ, dB).csr.nnz == 0
    with pytest.raises(native.NativeError):
        L.ultraSparseMatmult(gctx, "ba+*", dA, gctx.dense(np.zeros((k + 1, n))))
