"""mmchain / MV / VM parity: CUDA path (through the C ABI) vs the CPU oracle on the same inputs.

Shapes follow the reference's own test: MapMultChainTest (src/test/java/org/apache/sysml/test/integration/
functions/binary/matrix/MapMultChainTest.java:50-54,343-349): X 3471x567, sparsity 0.7 / 0.1, seeds 7/3/10,
values U(0,1), all three chain types; plus ragged/odd shapes, the C1 shape (100k x 1000) and, at the full
C2 size (10M x 1000), size-independent properties.
"""
import ctypes as C

import numpy as np
import pytest

from oracle import oracle as O
from tests.util import abs_terms_mmchain, assert_equal_matrices, assert_parity

pytestmark = pytest.mark.gpu

CT = {"XtXv": 0, "XtwXv": 1, "XtXvy": 2}


def _run(gctx, X, v, w, ct):
    from systemml_b200 import LibMatrixB200 as L
    dX, dv = gctx.dense(X), gctx.dense(v)
    dw = gctx.dense(w) if w is not None else None
    out = L.mmchain(gctx, "mmchain", dX, dv, dw, ct)
    res = out.to_host().reshape(-1)
    for o in (dX, dv, dw, out):
        if o is not None:
            o.free()
    return res


@pytest.mark.parametrize("sparsity", [0.7, 0.1])
@pytest.mark.parametrize("ct", ["XtXv", "XtwXv", "XtXvy"])
def test_reference_mapmultchain_shapes(gctx, ct, sparsity):
    X = O.java_test_matrix(3471, 567, 0, 1, sparsity, 7)
    v = O.java_test_matrix(567, 1, 0, 1, 0.7, 3)
    w = O.java_test_matrix(3471, 1, 0, 1, sparsity, 10)
    exp = O.mmchain_dense(X, v, w if ct != "XtXv" else None, CT[ct], k=O.num_threads())
    got = _run(gctx, X, v, w if ct != "XtXv" else None, ct)
    assert_equal_matrices(exp, got, what="mmchain %s sp=%s" % (ct, sparsity))
    # the reference test's own bar: absolute 1e-8 (MapMultChainTest.java:48)
    assert np.max(np.abs(exp - got)) < 1e-8


@pytest.mark.parametrize("n,m", [(1, 1), (2, 3), (7, 64), (24, 65), (25, 129), (1000, 100), (513, 257), (4099, 1000),
                                 (300, 1025), (130, 2049), (70, 4096), (33, 5000), (64, 9001), (3001, 4097), (2000, 5600), (1999, 5601),
                                 (1500, 8192), (700, 8193)])
@pytest.mark.parametrize("ct", [0, 1, 2])
def test_ragged_shapes(gctx, n, m, ct):
    rng = np.random.default_rng(1000 * n + m + ct)
    X = rng.uniform(-1, 1, (n, m))
    v = rng.uniform(-1, 1, (m, 1))
    w = rng.uniform(0.5, 1.5, (n, 1))
    exp = O.mmchain_dense(X, v, w if ct else None, ct, k=1)
    got = _run(gctx, X, v, w if ct else None, ct)
    # 1e-9 element-wise; zero-mean inputs produce a few cancelling cells, which take the rounding-bound exemption of
    # tests/util.py (the count is printed)
    assert_parity(exp, got, abs_terms_mmchain(X, v, w, ct), n + m, what="mmchain %dx%d ct=%d" % (n, m, ct))


def test_c1_shape_100k_x_1000(gctx):
    """BASELINE configs[0]: LinearRegCG 100k x 1000, X ~ U(-1,1) (scripts/datagen/genRandData4LinearRegression.dml)."""
    n, m = 100_000, 1000
    X = O.rand_uniform(n * m, seed=0).reshape(n, m)
    v = O.rand_uniform(m, seed=1).reshape(m, 1)
    w = O.rand_uniform(n, seed=2, lo=0.0, hi=1.0).reshape(n, 1)
    for ct in (0, 1, 2):
        exp = O.mmchain_dense(X, v, w if ct else None, ct, k=O.num_threads())
        got = _run(gctx, X, v, w if ct else None, ct)
        assert_equal_matrices(exp, got, what="C1 mmchain ct=%d" % ct)


def test_device_generator_matches_cpu_generator(gctx):
    d = gctx.rand(1000, 37, seed=5, lo=-1.0, hi=1.0, offset=123)
    h = O.rand_uniform(1000 * 37, seed=5, offset=123).reshape(1000, 37)
    assert np.array_equal(d.to_host(), h)
    d.free()


def test_unaligned_base_pointer(gctx):
    """A row shard may start at an address that is only 8-byte aligned: the bulk-copy path must not be used."""
    from systemml_b200 import LibMatrixB200 as L, GPUObject
    n, m = 501, 333
    rng = np.random.default_rng(3)
    X = rng.uniform(-1, 1, (n + 1, m))
    v = rng.uniform(-1, 1, (m, 1))
    big = gctx.dense(X)
    view = GPUObject.wrap(gctx, big.ptr + 8 * m, n, m)  # skip one (odd-length) row -> 8-byte aligned only
    dv = gctx.dense(v)
    out = L.mmchain(gctx, "mmchain", view, dv, None, "XtXv")
    exp = O.mmchain_dense(np.ascontiguousarray(X[1:]), v, None, 0)
    assert_equal_matrices(exp, out.to_host())
    for o in (big, dv, out):
        o.free()


def test_empty_and_error_paths(gctx):
    from systemml_b200 import LibMatrixB200 as L, native
    dv = gctx.dense(np.ones((5, 1)))
    # empty X (0 rows): the result is the zero vector (LibMatrixMult.java:245-248)
    X0 = gctx.empty(0, 5)
    out = L.mmchain(gctx, "mmchain", X0, dv, None, "XtXv")
    assert np.array_equal(out.to_host().reshape(-1), np.zeros(5))
    # dimension mismatch -> error, not a crash
    X = gctx.dense(np.ones((4, 6)))
    with pytest.raises(native.NativeError) as ei:
        L.mmchain(gctx, "mmchain", X, dv, None, "XtXv")
    assert ei.value.code == native.E_DIM
    # weights missing for XtwXv
    with pytest.raises(native.NativeError):
        L.mmchain(gctx, "mmchain", X, gctx.dense(np.ones((6, 1))), None, "XtwXv")
    # bad chain type through the raw ABI
    rc = gctx._lib.sysml_b200_mmchain_dense(gctx.handle, C.c_void_p(X.ptr), C.c_void_p(dv.ptr), None, C.c_void_p(dv.ptr), 4, 6, 9)
    assert rc == native.E_INVALID and b"chain type" in gctx._lib.sysml_b200_last_error(gctx.handle)


def test_hostvec_entry_point(gctx):
    from systemml_b200 import LibMatrixB200 as L
    n, m = 5000, 1000
    X = O.rand_uniform(n * m, seed=11).reshape(n, m)
    v = O.rand_uniform(m, seed=12)
    w = O.rand_uniform(n, seed=13, lo=0.0, hi=2.0)
    dX = gctx.dense(X)
    for ct in (0, 1, 2):
        got = L.mmchain_hostvec(gctx, dX, v, w if ct else None, ct)
        exp = O.mmchain_dense(X, v, w if ct else None, ct, k=4)
        assert_equal_matrices(exp, got, what="hostvec ct=%d" % ct)
    dw = gctx.dense(w)
    got = L.mmchain_hostvec(gctx, dX, v, None, 1, d_w=dw)
    assert_equal_matrices(O.mmchain_dense(X, v, w, 1, k=4), got)
    dX.free()
    dw.free()


def test_run_to_run_determinism(gctx):
    from systemml_b200 import LibMatrixB200 as L
    dX = gctx.rand(200_000, 1000, seed=21)
    dv = gctx.rand(1000, 1, seed=22)
    a = L.mmchain(gctx, "mmchain", dX, dv, None, 0).to_host()
    b = L.mmchain(gctx, "mmchain", dX, dv, None, 0).to_host()
    assert np.array_equal(a, b)
    dX.free()


@pytest.mark.parametrize("n,m", [(3471, 567), (1000, 1), (1, 1000), (257, 4097), (100_000, 100), (1201, 5600), (999, 8192), (130, 8193)])
def test_mv_and_vm(gctx, n, m):
    """X %*% v and t(X) %*% y (= t(t(y) %*% X)): matrixMultDenseDenseMVShortRHS / VM (LibMatrixMult.java:950-993)."""
    from systemml_b200 import LibMatrixB200 as L
    X = O.rand_uniform(n * m, seed=31).reshape(n, m)
    v = O.rand_uniform(m, seed=32).reshape(m, 1)
    y = O.rand_uniform(n, seed=33).reshape(n, 1)
    dX, dv, dy = gctx.dense(X), gctx.dense(v), gctx.dense(y)
    mv = L.matmult(gctx, "ba+*", dX, dv).to_host()
    exp, _ = O.matmult_dense(X, v, k=4)
    assert_parity(exp, mv, np.abs(X) @ np.abs(v), m, rel=1e-9 if m > 1 else 1e-12, what="MV %dx%d" % (n, m))
    vm = L.matmult(gctx, "ba+*", dX, dy, isLeftTransposed=True).to_host()
    exp2, _ = O.matmult_dense(np.ascontiguousarray(y.T), X, k=4)  # t(y) %*% X, the reference's rewrite
    assert_parity(exp2.reshape(-1), vm.reshape(-1), (np.abs(X).T @ np.abs(y)).reshape(-1), n, what="VM %dx%d" % (n, m))
    for o in (dX, dv, dy):
        o.free()


def test_full_size_properties_10M_x_1000(gctx):
    """BASELINE configs[1] (10M x 1000, 80 GB): size-independent properties instead of a full oracle run.
    (1) additivity over row shards, (2) linearity in v, (3) oracle parity on the first 100k rows."""
    from systemml_b200 import LibMatrixB200 as L, GPUObject
    free, _ = gctx.getAvailableMemory()
    n, m = 10_000_000, 1000
    if free < (n * m * 8) * 1.05:
        n = int(free * 0.8 / (m * 8)) // 1000 * 1000
    X = gctx.rand(n, m, seed=1)
    v1, v2 = gctx.rand(m, 1, seed=2), gctx.rand(m, 1, seed=3)
    w = gctx.rand(n, 1, seed=4, lo=0.0, hi=1.0)
    full = L.mmchain(gctx, "mmchain", X, v1, w, "XtwXv").to_host().reshape(-1)
    # (1) two row shards add up to the whole (this is the multi-GPU plan on one device)
    h = (n // 2) // 1000 * 1000
    top = GPUObject.wrap(gctx, X.ptr, h, m)
    bot = GPUObject.wrap(gctx, X.ptr + 8 * h * m, n - h, m)
    wt = GPUObject.wrap(gctx, w.ptr, h, 1)
    wb = GPUObject.wrap(gctx, w.ptr + 8 * h, n - h, 1)
    parts = L.mmchain(gctx, "mmchain", top, v1, wt, "XtwXv").to_host() + L.mmchain(gctx, "mmchain", bot, v1, wb, "XtwXv").to_host()
    assert_equal_matrices(full, parts.reshape(-1), what="shard additivity")
    # (2) linearity: f(2*v1 - 3*v2) = 2 f(v1) - 3 f(v2)
    f2 = L.mmchain(gctx, "mmchain", X, v2, w, "XtwXv").to_host().reshape(-1)
    comb = gctx.dense(2 * v1.to_host() - 3 * v2.to_host())
    fc = L.mmchain(gctx, "mmchain", X, comb, w, "XtwXv").to_host().reshape(-1)
    lin = 2 * full - 3 * f2
    assert np.max(np.abs(fc - lin)) <= 1e-9 * np.max(np.abs(lin))
    # (3) oracle on the first 100k rows (same generator on the CPU)
    k = 100_000
    Xh = O.rand_uniform(k * m, seed=1).reshape(k, m)
    wh = O.rand_uniform(k, seed=4, lo=0.0, hi=1.0)
    sub = L.mmchain(gctx, "mmchain", GPUObject.wrap(gctx, X.ptr, k, m), v1, GPUObject.wrap(gctx, w.ptr, k, 1), "XtwXv")
    exp = O.mmchain_dense(Xh, v1.to_host(), wh, 1, k=O.num_threads())
    assert_equal_matrices(exp, sub.to_host(), what="first 100k rows")
    X.free()
    w.free()


@pytest.mark.parametrize("n,m", [(3471, 567), (1000, 1), (257, 4097), (20_000, 1000), (999, 8192), (5, 100), (4099, 64), (501, 2000), (303, 3000),
                                 (2001, 200), (130, 5601)])
def test_mmchain_with_a_row_function(gctx, n, m):
    """sysml_b200_mmchain_rowfn_dense: out = t(X) %*% f(X %*% v, w, y) in one pass, f a postfix program.
    (1) With f = the three fixed chain types the result is bit for bit mmchain_dense's (same kernel, same order).
    (2) GLM's outer step for Poisson / log (GLM.dml:702-706): f = y - exp(eta) with mu = exp(eta) as the per-row second
    result, against the unfused evaluation through the oracle: X %*% v (matrixMultDenseDense), exp, minus (LibMatrixBincell),
    t(X) %*% . (VM form) -- 1e-9 per cell; mu itself to 1e-13 (CUDA's exp vs the host's).
    (3) Inputs the program does not read may be NULL; a malformed program or a missing input is refused."""
    from systemml_b200 import LibMatrixB200 as L, native
    E = L.Expr
    X = O.java_test_matrix(n, m, -1, 1, 0.9, 7)
    v = O.java_test_matrix(m, 1, -0.5, 0.5, 1.0, 3) / max(1.0, np.sqrt(m) / 4)
    w = O.java_test_matrix(n, 1, 0, 1, 1.0, 10)
    y = np.floor(O.java_test_matrix(n, 1, 0, 6, 1.0, 11))
    dX, dv, dw, dy = gctx.dense(X), gctx.dense(v), gctx.dense(w), gctx.dense(y)
    for ct, prog in (("XtXv", E.input(0)), ("XtwXv", E.input(0) * E.input(1)), ("XtXvy", E.input(0) - E.input(1))):
        ref = L.mmchain(gctx, "mmchain", dX, dv, dw if ct != "XtXv" else None, ct).to_host()
        got, _ = L.mmchainRowFn(gctx, "mmchain", dX, dv, prog, w=dw if ct != "XtXv" else None)
        assert np.array_equal(ref, got.to_host()), ct
    mu_e = E.input(0).unary("exp")
    grad, mu = L.mmchainRowFn(gctx, "mmchain", dX, dv, [E.input(2) - mu_e, mu_e], y=dy)
    eta = O.matmult_dense(X, v, 4)[0]
    mu_ref = np.exp(eta)
    np.testing.assert_allclose(mu.to_host(), mu_ref, rtol=1e-13, atol=0)
    resid = y - mu_ref
    exp = (X.astype(np.longdouble).T @ resid.astype(np.longdouble)).astype(np.float64)
    assert_parity(exp.reshape(-1), grad.to_host().reshape(-1), (np.abs(X).T @ np.abs(resid)).reshape(-1), n, what="mmchain rowfn %dx%d" % (n, m))
    # a constant in the program, both row vectors, one result
    got, side = L.mmchainRowFn(gctx, "mmchain", dX, dv, (E.input(0) * E.const(0) + E.input(1)) * E.input(2), w=dw, y=dy, consts=(0.25,))
    assert side is None
    t = (eta * 0.25 + w) * y
    exp = (X.astype(np.longdouble).T @ t.astype(np.longdouble)).astype(np.float64)
    assert_parity(exp.reshape(-1), got.to_host().reshape(-1), (np.abs(X).T @ np.abs(t)).reshape(-1), n, what="mmchain rowfn const %dx%d" % (n, m))
    with pytest.raises(native.NativeError, match="null"):
        L.mmchainRowFn(gctx, "mmchain", dX, dv, E.input(0) * E.input(1))          # reads w, which is not given
    with pytest.raises(native.NativeError):
        L.mmchainRowFn(gctx, "mmchain", dX, dv, [E.input(0)] * 3)                  # three results
    with pytest.raises(native.NativeError):
        L.mmchainRowFn(gctx, "mmchain", dX, gctx.dense(np.zeros((m + 1, 1))), E.input(0))


def test_mmchain_row_function_limits(gctx):
    from systemml_b200 import LibMatrixB200 as L, native
    E = L.Expr
    dX, dv = gctx.dense(np.ones((4, 8200))), gctx.dense(np.ones((8200, 1)))
    with pytest.raises(native.NativeError, match="8192"):
        L.mmchainRowFn(gctx, "mmchain", dX, dv, E.input(0))
    # log of a zero-filled padding row must not leak: n = 5 rows in tiles of 4 or more
    X = O.java_test_matrix(5, 40, 0.5, 1, 1.0, 2)
    v = np.ones((40, 1))
    got, _ = L.mmchainRowFn(gctx, "mmchain", gctx.dense(X), gctx.dense(v), E.input(0).unary("log"))
    exp = X.T @ np.log(X @ v)
    np.testing.assert_allclose(got.to_host(), exp, rtol=1e-12)
