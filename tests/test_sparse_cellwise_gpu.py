"""Sparse-safe cellwise operators whose operand and result stay CSR (csrc/sparse_cellwise.cu) against the oracle's
restatement of LibMatrixBincell's sparse branches: safeBinaryScalar SPARSE <- SPARSE (LibMatrixBincell.java:952-995),
safeBinaryMVSparse (:338-440), safeBinaryMMSparseDenseSkip (:765-789), safeBinaryMMSparseSparse / mergeForSparseBinary
(:597-660, 1279-1340), MatrixBlock.sparseUnaryOperations (MatrixBlock.java:2640-2672).  Structure (row pointers, column
indices, nnz) and values are compared bit for bit -- the value functions are the ones the dense kernels use, which are
bit-exact against the oracle for everything but pow and the transcendentals (compared at 1e-13 there).  Every result
is ALSO compared with the dense form of the same operator (what the reference's GPU twin would compute after
sparseToDense), so the two product paths cannot drift apart."""
import numpy as np
import pytest

from oracle import oracle as O

pytestmark = pytest.mark.gpu

OPS = {"+": 0, "-": 1, "*": 2, "/": 3, "^": 4, "<": 5, "<=": 6, ">": 7, ">=": 8, "==": 9, "!=": 10, "min": 11, "max": 12,
       "&&": 13, "||": 14, "1-*": 15, "-nz": 16, "%%": 17, "%/%": 18}
UN = {"sqrt": 1, "round": 2, "abs": 3, "floor": 5, "ceil": 6, "sign": 7, "sin": 9, "tan": 11, "tanh": 12}
INEXACT = {"^", "sin", "tan", "tanh"}


def _csr(gctx, A):
    rp, ci, va = O.dense_to_csr(A)
    return gctx.csr(rp.astype(np.int32), ci, va, A.shape), (rp, ci, va)


def _same(got, ref, name, shape):
    rp, ci, va = got.to_host()
    rrp, rci, rva = ref
    assert got.csr.nnz == len(rci) == rrp[-1], name
    assert np.array_equal(np.asarray(rp, dtype=np.int64), rrp), name
    assert np.array_equal(ci[:len(rci)], rci), name
    if name in INEXACT:
        np.testing.assert_allclose(va[:len(rva)], rva, rtol=1e-13, atol=0, err_msg=name)
    else:
        assert np.array_equal(va[:len(rva)], rva, equal_nan=True), name
        assert np.array_equal(np.signbit(va[:len(rva)]), np.signbit(rva)), name
    assert np.all(va[:len(rva)] != 0.0), "no explicit zeros (SparseRow.append)"


def _dense_of(csr, shape):
    rp, ci, va = csr
    D = np.zeros(shape)
    for r in range(shape[0]):
        D[r, ci[rp[r]:rp[r + 1]]] = va[rp[r]:rp[r + 1]]
    return D


SHAPES = [(300, 200, 0.1), (1, 500, 0.3), (700, 3, 0.5), (33, 64, 1.0), (2049, 1300, 0.01), (5000, 40, 0.002)]


@pytest.mark.parametrize("rows,cols,sp", SHAPES)
def test_csr_scalar_and_unary(gctx, rows, cols, sp):
    from systemml_b200 import LibMatrixB200 as L, native
    A = O.java_test_matrix(rows, cols, -2, 2, sp, 31)
    A.flat[np.flatnonzero(A)[::7]] = 1.0
    dA, a = _csr(gctx, A)
    # (operator, constant, scalar on the left?) pairs the reference treats as sparse-safe (ScalarOperator.java:53-59,
    # RightScalarOperator.java:42-49, LeftScalarOperator.java:41-46) and that map 0 to 0
    cases = [("*", 0.37, False), ("*", -3.0, True), ("*", 0.0, False), ("/", 4.0, False), ("^", 2.0, False), ("^", 0.5, False),
             (">", 0.5, False), (">", 0.0, False), (">=", 0.25, False), ("<", -0.5, False), ("<=", -1e-9, False), ("<", 0.5, True),
             ("!=", 0.0, False), ("==", 1.0, False), ("-", 0.0, False), ("max", -0.5, False), ("min", 0.5, False), ("&&", 3.0, False),
             ("-nz", 0.5, False), ("%%", 0.75, False), ("%/%", 0.3, False), ("min", 0.25, True)]
    for name, s, left in cases:
        ref = O.csr_scalar_op(*a, rows, s, OPS[name], left)
        got = L.sparseScalarOp(gctx, name, dA, s, OPS[name], left)
        _same(got, ref, name, A.shape)
        dense = O.binary_ms(A, s, OPS[name], left)
        if name not in INEXACT:
            assert np.array_equal(_dense_of(got.to_host(), A.shape), dense + 0.0), name   # + 0.0: a sparse cell has no sign
        got.free()
    for name, s, left in (("+", 1.0, False), ("-", 1.0, True), ("/", 0.0, False), ("==", 0.0, False), ("<", 0.5, False), ("1-*", 2.0, False),
                          ("^", 0.0, False), ("^", -1.0, False), ("||", 1.0, False), ("max", 0.5, False)):
        with pytest.raises(native.NativeError, match="not sparse-safe"):
            L.sparseScalarOp(gctx, name, dA, s, OPS[name], left)
    for name, code in UN.items():
        src, ad = (dA, a)
        if name == "sqrt":
            src, ad = _csr(gctx, np.abs(A))
        ref = O.csr_unary_op(*ad, rows, code)
        got = L.sparseUnaryOp(gctx, name, src, code)
        _same(got, ref, name, A.shape)
        got.free()
    for code in (0, 4, 8, 10):   # exp, log, sigmoid, cos
        with pytest.raises(native.NativeError, match="not sparse-safe"):
            L.sparseUnaryOp(gctx, "u", dA, code)


@pytest.mark.parametrize("rows,cols,sp", SHAPES)
def test_csr_op_dense_operand(gctx, rows, cols, sp):
    from systemml_b200 import LibMatrixB200 as L, native
    A = O.java_test_matrix(rows, cols, -2, 2, sp, 41)
    dA, a = _csr(gctx, A)
    col = O.java_test_matrix(rows, 1, -1, 1, 0.7, 42)
    col[::5] = 1.0                                      # the ROW COPY branch (LibMatrixBincell.java:372-376)
    row = O.java_test_matrix(1, cols, -1, 1, 0.8, 43)
    full = O.java_test_matrix(rows, cols, 0.5, 2, 1.0, 44)
    holes = O.java_test_matrix(rows, cols, -1, 1, 0.5, 45)
    for name, B in (("*", col), ("*", row), ("*", full), ("*", holes), ("&&", col), ("&&", row), ("&&", holes), ("/", full),
                    ("/", np.abs(col) + 0.5), ("/", np.abs(row) + 0.5)):
        ref = O.csr_dense_op(*a, rows, cols, B, OPS[name])
        got = L.sparseDenseOp(gctx, name, dA, gctx.dense(B), OPS[name])
        _same(got, ref, name, A.shape)
        assert np.array_equal(_dense_of(got.to_host(), A.shape), O.binary_mm(A, B, OPS[name]) + 0.0), name
        got.free()
    # dense on the left (commutes for Multiply; the reference's skip path takes either side, LibMatrixBincell.java:766-770)
    got = L.sparseDenseOp(gctx, "*", dA, gctx.dense(holes), OPS["*"], dense_left=True)
    assert np.array_equal(_dense_of(got.to_host(), A.shape), holes * A + 0.0)
    with pytest.raises(native.NativeError, match="sparse"):       # dense / CSR divides by the cells that are not stored
        L.sparseDenseOp(gctx, "/", dA, gctx.dense(full), OPS["/"], dense_left=True)
    # not sparse-safe for this operand: the reference would fill the cells the rows do not store
    bad = [("+", col), ("-", row), ("||", row), ("max", np.abs(col)), ("+", full), ("%%", full)]
    if np.any(col == 0):
        bad.append(("/", col))
    if rows > 1 and cols > 1 and sp < 1.0:
        nanv = col.copy()
        nanv[int(np.argmax(np.count_nonzero(A, axis=1)))] = np.nan   # 0 * NaN in the gaps of a row that holds cells
        bad.append(("*", nanv))
    for name, B in bad:
        with pytest.raises(native.NativeError, match="sparse"):
            L.sparseDenseOp(gctx, name, dA, gctx.dense(B), OPS[name])


@pytest.mark.parametrize("rows,cols,sp", SHAPES)
def test_csr_op_csr(gctx, rows, cols, sp):
    from systemml_b200 import LibMatrixB200 as L, native
    A = O.java_test_matrix(rows, cols, -2, 2, sp, 51)
    B = O.java_test_matrix(rows, cols, -2, 2, sp, 52)
    if rows > 3:
        B[::3] = A[::3]          # A - B cancels, A != B is 0, min/max meet equal cells
        A[1::4] = 0.0            # rows only B holds
    dA, a = _csr(gctx, A)
    dB, b = _csr(gctx, B)
    for name in ("+", "-", "*", "&&", "||", "!=", "<", ">", "min", "max", "-nz"):
        ref = O.csr_csr_op(a, b, rows, OPS[name])
        got = L.sparseSparseOp(gctx, name, dA, dB, OPS[name])
        _same(got, ref, name, A.shape)
        assert np.array_equal(_dense_of(got.to_host(), A.shape), O.binary_mm(A, B, OPS[name]) + 0.0), name
        got.free()
    for name in ("==", "<=", ">=", "/", "1-*", "^"):
        with pytest.raises(native.NativeError, match="not sparse-safe"):
            L.sparseSparseOp(gctx, name, dA, dB, OPS[name])
    # the csrgeam twin agrees with the generic merge
    s1 = L.sparseSparseAdd(gctx, "+", dA, dB).to_host()
    s2 = L.sparseSparseOp(gctx, "+", dA, dB, OPS["+"]).to_host()
    assert all(np.array_equal(x, y) for x, y in zip(s1, s2))


def test_csr_cellwise_edge_cases(gctx):
    from systemml_b200 import LibMatrixB200 as L, native
    Z = np.zeros((40, 30))
    A = O.java_test_matrix(40, 30, -1, 1, 0.2, 3)
    dZ, z = _csr(gctx, Z)
    dA, a = _csr(gctx, A)
    assert L.sparseScalarOp(gctx, "*", dZ, 2.0, OPS["*"]).csr.nnz == 0
    assert L.sparseSparseOp(gctx, "*", dA, dZ, OPS["*"]).csr.nnz == 0
    assert L.sparseSparseOp(gctx, "-", dA, dA, OPS["-"]).csr.nnz == 0
    assert L.sparseDenseOp(gctx, "*", dA, gctx.dense(np.zeros((40, 1))), OPS["*"]).csr.nnz == 0
    _same(L.sparseSparseOp(gctx, "+", dA, dZ, OPS["+"]), a, "+", A.shape)
    # stored NaN / Inf: a row that meets an EMPTY row goes through appendLeft (0 * NaN is kept), cells inside a merge are joined
    N = A.copy()
    N[0, :] = 0.0
    N[0, 3], N[1, 4], N[2, 5] = np.nan, np.inf, np.nan
    M = A.copy()
    M[0, :] = 0.0                   # empty row against the NaN row
    M[2, 5] = 0.0
    dN, n = _csr(gctx, N)
    dM, m = _csr(gctx, M)
    for name in ("*", "+", "min"):
        _same(L.sparseSparseOp(gctx, name, dN, dM, OPS[name]), O.csr_csr_op(n, m, 40, OPS[name]), name, A.shape)
        _same(L.sparseSparseOp(gctx, name, dM, dN, OPS[name]), O.csr_csr_op(m, n, 40, OPS[name]), name, A.shape)
    _same(L.sparseScalarOp(gctx, "*", dN, 0.0, OPS["*"]), O.csr_scalar_op(*n, 40, 0.0, OPS["*"]), "*", A.shape)   # NaN * 0 stays
    # dimension errors and the protocol
    with pytest.raises(native.NativeError):
        L.sparseSparseOp(gctx, "+", dA, _csr(gctx, np.zeros((30, 40)))[0], OPS["+"])
    with pytest.raises(native.NativeError):
        L.sparseDenseOp(gctx, "*", dA, gctx.dense(np.zeros((3, 3))), OPS["*"])
    import ctypes as C
    rp = gctx.allocate(41 * 4)
    nnz = C.c_int64()
    rc = gctx._lib.sysml_b200_csr_scalar_op(gctx.handle, C.c_void_p(dA.csr.rowPtr), C.c_void_p(dA.csr.colInd), C.c_void_p(dA.csr.val), 40, 30,
                                            2.0, OPS["*"], 0, C.c_void_p(rp), C.c_void_p(rp), None, C.byref(nnz))
    assert rc == native.E_INVALID       # c_colind without c_val


def test_csr_cellwise_at_scale(gctx):
    """500k x 10k, ~10 cells per row (a slice of BASELINE config C3): X * y (column vector), X * 0.5, X * X -- a dense
    image of this block would be 40 GB; checked against the oracle on the structure and bit for bit on the values."""
    from systemml_b200 import LibMatrixB200 as L
    rows, cols = 500_000, 10_000
    X = gctx.rand_csr(rows, cols, 10, 2)
    rp, ci, va = X.to_host()
    rp = np.asarray(rp, dtype=np.int64)
    y = O.rand_uniform(rows, seed=8).reshape(rows, 1)
    y[::9] = 0.0
    got = L.sparseDenseOp(gctx, "*", X, gctx.dense(y), OPS["*"])
    _same(got, O.csr_dense_op(rp, ci, va, rows, cols, y, OPS["*"]), "*", (rows, cols))
    assert got.csr.nnz < X.csr.nnz
    _same(L.sparseScalarOp(gctx, "*", X, 0.5, OPS["*"]), O.csr_scalar_op(rp, ci, va, rows, 0.5, OPS["*"]), "*", (rows, cols))
    _same(L.sparseSparseOp(gctx, "*", X, X, OPS["*"]), O.csr_csr_op((rp, ci, va), (rp, ci, va), rows, OPS["*"]), "*", (rows, cols))
