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
        _parity_log({"what": what, "cells": int(e.size), "worst_rel": float(relerr[worst]), "exempt_cells": 0, "tol": rel})
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


# ---------------------------------------------------------------------------------------------------------
# Element-wise parity with the ONE documented exemption ("Numerics").
#
# The bar is GPUTests.assertEqualMatrices at 1e-9 on every non-zero finite cell.  A cell of a sum of products whose
# terms cancel (zero-mean inputs: |sum| << sum of |terms|) cannot meet a RELATIVE bar in any fp64 implementation --
# the reference's own result moves by more than 1e-9 of such a cell when its summation order changes (other thread
# count k, other blocking).  For those cells, and only for them, the test falls back to the a-priori bound on the
# disagreement of two fp64 summations of the same terms in different orders (Higham, Accuracy and Stability of
# Numerical Algorithms, eq. 4.4): |a - b| <= 2 * gamma_n * S with S = sum of |terms| and gamma_n = n*u/(1 - n*u),
# u = 2^-53.  Every use prints the worst relative error it saw and how many cells took the exemption.
# ---------------------------------------------------------------------------------------------------------
def _parity_log(rec):
    import json
    import os
    path = os.environ.get("SYSML_PARITY_LOG")
    print("parity: " + json.dumps(rec))
    if path:
        with open(path, "a") as f:
            f.write(json.dumps(rec) + "\n")


def assert_parity(expected, actual, abs_terms=None, n_terms=None, rel=REL_TOL, what=""):
    """expected/actual: arrays.  abs_terms: per-cell sum of |terms| (same shape) and n_terms: terms per cell -- when
    given, cells above `rel` may pass through the rounding-bound exemption described above."""
    e = np.asarray(expected, dtype=np.float64).reshape(-1)
    a = np.asarray(actual, dtype=np.float64).reshape(-1)
    assert e.shape == a.shape, "%s: shape %s vs %s" % (what, e.shape, a.shape)
    nan_e = np.isnan(e)
    assert np.array_equal(nan_e, np.isnan(a)), "%s: NaN pattern differs" % what
    inf_e = np.isinf(e)
    assert np.array_equal(e[inf_e], a[inf_e]), "%s: infinities differ" % what
    reg = ~nan_e & ~inf_e
    diff = np.abs(e - a)
    with np.errstate(divide="ignore", invalid="ignore"):
        relerr = np.where(reg & (e != 0), diff / np.abs(e), 0.0)
    bad = reg & (((e != 0) & (relerr >= rel)) | ((e == 0) & (diff > rel)))
    n_exempt = 0
    if bad.any() and abs_terms is not None:
        S = np.asarray(abs_terms, dtype=np.float64).reshape(-1)
        u = 2.0 ** -53
        gamma = n_terms * u / (1.0 - n_terms * u)
        within = diff <= 2.0 * gamma * S
        n_exempt = int(np.count_nonzero(bad & within))
        bad = bad & ~within
    worst = float(relerr.max()) if relerr.size else 0.0
    _parity_log({"what": what, "cells": int(e.size), "worst_rel": worst, "exempt_cells": n_exempt, "tol": rel})
    if bad.any():
        i = int(np.argmax(np.where(bad, relerr + 1.0, 0.0)))
        raise AssertionError("%s: cell %d expected %r actual %r (relative error %.3e, %d cells beyond %.0e with no exemption)"
                             % (what, i, e[i], a[i], relerr[i], int(np.count_nonzero(bad)), rel))
    return worst


def abs_terms_mmchain(X, v, w, ct):
    """Sum of |terms| per output cell of t(X) %*% (w * (X %*% v)) (ct 1), ... - w (ct 2), plain (ct 0)."""
    aX = np.abs(X)
    t = aX @ np.abs(np.asarray(v).reshape(-1))
    if ct == 1:
        t = t * np.abs(np.asarray(w).reshape(-1))
    elif ct == 2:
        t = t + np.abs(np.asarray(w).reshape(-1))
    return aX.T @ t
