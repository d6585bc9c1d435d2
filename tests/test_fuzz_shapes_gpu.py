"""Seeded shape fuzzing around the tile / fragment boundaries of the DMMA kernels (128-wide tiles, 8-wide fragments,
16-deep k blocks, strip mappings for diagonal / ragged / small-output tiles) and of the streamed row kernels,
against extended-precision numpy."""
import numpy as np
import pytest

from tests.util import assert_equal_matrices

pytestmark = pytest.mark.gpu


def _ld(a):
    return a.astype(np.longdouble)


@pytest.mark.parametrize("seed", range(12))
def test_tsmm_boundary_shapes(gctx, seed):
    from systemml_b200 import LibMatrixB200 as L
    rng = np.random.default_rng(100 + seed)
    m = int(rng.choice([2, 8, 64, 120, 128, 130, 136, 248, 256, 258, 264, 384, 392, 500, 1000, 1016, 1030]))
    n = int(rng.choice([1, 15, 16, 17, 100, 513, 2048, 4099, 20011]))
    X = rng.uniform(-1, 1, (n, m))
    X[:, rng.integers(0, m)] = 0.0  # a zero column: exact nnz must notice it
    exp = (_ld(X).T @ _ld(X)).astype(np.float64)
    out, nnz = L.matmultTSMM(gctx, "tsmm", gctx.dense(X), True, with_nnz=True)
    got = out.to_host()
    assert np.array_equal(got, got.T)
    assert_equal_matrices(exp, got, rel=1e-9, what="tsmm %dx%d" % (n, m))
    assert nnz == np.count_nonzero(got)


@pytest.mark.parametrize("seed", range(24))
def test_matmult_boundary_shapes(gctx, seed):
    from systemml_b200 import LibMatrixB200 as L
    rng = np.random.default_rng(200 + seed)
    M = int(rng.choice([1, 2, 7, 8, 50, 56, 63, 64, 66, 100, 128, 130, 256, 300, 1000]))
    N = int(rng.choice([1, 2, 8, 20, 24, 26, 50, 64, 66, 100, 104, 128, 130, 200, 260]))
    K = int(rng.choice([1, 3, 4, 15, 16, 17, 100, 1000, 1024, 4097, 30000]))
    ta, tb = bool(rng.integers(0, 2)), bool(rng.integers(0, 2))
    A = rng.uniform(-1, 1, (K, M) if ta else (M, K))
    B = rng.uniform(-1, 1, (N, K) if tb else (K, N))
    exp = ((_ld(A).T if ta else _ld(A)) @ (_ld(B).T if tb else _ld(B))).astype(np.float64)
    got = L.matmult(gctx, "ba+*", gctx.dense(A), gctx.dense(B), isLeftTransposed=ta, isRightTransposed=tb).to_host()
    scale = max(np.max(np.abs(exp)), 1e-300)
    assert np.max(np.abs(exp - got)) <= 1e-12 * scale * max(1, K ** 0.5), (M, N, K, ta, tb)


@pytest.mark.parametrize("seed", range(10))
def test_mmchain_and_row_aggregates_boundary_shapes(gctx, seed):
    from systemml_b200 import LibMatrixB200 as L
    rng = np.random.default_rng(300 + seed)
    m = int(rng.choice([1, 31, 64, 100, 128, 129, 256, 257, 1000, 1024, 1025, 2049, 4096, 4097, 5600, 5601, 8192]))
    n = int(rng.choice([1, 3, 33, 257, 1000, 4111]))
    X = rng.uniform(-1, 1, (n, m))
    v, w = rng.uniform(-1, 1, m), rng.uniform(0, 1, n)
    dX = gctx.dense(X)
    for ct in (0, 1, 2):
        t = _ld(X) @ _ld(v)
        t = t * w if ct == 1 else (t - w if ct == 2 else t)
        exp = (_ld(X).T @ t).astype(np.float64)
        got = L.mmchain(gctx, "mmchain", dX, gctx.dense(v.reshape(m, 1)), gctx.dense(w.reshape(n, 1)) if ct else None, ct).to_host().reshape(-1)
        assert np.max(np.abs(exp - got)) <= 1e-11 * max(np.max(np.abs(exp)), 1e-300), (n, m, ct)
    for opc, f in (("uark+", lambda a: a.sum(axis=1)), ("uarmin", lambda a: a.min(axis=1)), ("uack+", lambda a: a.sum(axis=0)),
                   ("uacmax", lambda a: a.max(axis=0))):
        got = L.unaryAggregate(gctx, opc, dX, opc).to_host().reshape(-1)
        exp = f(_ld(X)).astype(np.float64)
        assert np.max(np.abs(exp - got)) <= 1e-12 * max(np.max(np.abs(exp)), 1.0), (n, m, opc)
