"""Shared helpers for the parity tests."""
import numpy as np

REL_TOL = 1e-9  # BASELINE.json north_star / GPUTests.DOUBLE_PRECISION_THRESHOLD (gpu/GPUTests.java:57)


def assert_equal_matrices(expected, actual, rel=REL_TOL, what=""):
    """GPUTests.assertEqualMatrices (src/test/java/org/apache/sysml/test/gpu/GPUTests.java:254-302):
    relative error < rel on cells whose expected value is non-zero and finite, |diff| <= rel otherwise
    (NaN must match NaN, +-Inf must match exactly)."""
    e = np.asarray(expected, dtype=np.float64).reshape(-1)
    a = np.asarray(actual, dtype=np.float64).reshape(-1)
    assert e.shape == a.shape, "%s: shape %s vs %s" % (what, e.shape, a.shape)
    nan_e, nan_a = np.isnan(e), np.isnan(a)
    assert np.array_equal(nan_e, nan_a), "%s: NaN pattern differs" % what
    inf_e = np.isinf(e)
    assert np.array_equal(e[inf_e], a[inf_e]), "%s: infinities differ" % what
    reg = ~nan_e & ~inf_e
    nz = reg & (e != 0)
    if nz.any():
        relerr = np.abs((e[nz] - a[nz]) / e[nz])
        worst = int(np.argmax(relerr))
        assert relerr[worst] < rel, "%s: max relative error %.3e (expected %r, actual %r)" % (
            what, relerr[worst], e[nz][worst], a[nz][worst])
    z = reg & (e == 0)
    if z.any():
        assert np.max(np.abs(a[z])) <= rel, "%s: expected 0, got %r" % (what, np.max(np.abs(a[z])))


def max_rel_err(expected, actual):
    e = np.asarray(expected, dtype=np.float64).reshape(-1)
    a = np.asarray(actual, dtype=np.float64).reshape(-1)
    nz = np.isfinite(e) & (e != 0)
    return float(np.max(np.abs((e[nz] - a[nz]) / e[nz]))) if nz.any() else 0.0
