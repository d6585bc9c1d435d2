"""End-to-end workloads (LinearRegCG, GLM Poisson/log) on the GPU against the same algorithms run on the CPU
with the oracle's kernels: BASELINE configs[0] (100k x 1000) and a smaller shape; beta within 1e-9 relative
(north_star: 'run end-to-end ... within 1e-9 of the CPU reference')."""
import numpy as np
import pytest

from oracle import oracle as O
from tests.util import assert_equal_matrices

pytestmark = pytest.mark.gpu


def cpu_linreg_cg(X, y, reg, tol, maxi, k):
    """scripts/algorithms/LinearRegCG.dml:160-200 with the oracle's mmchain / VM kernels."""
    n, m = X.shape
    if maxi == 0:
        maxi = m
    r = -O.matmult_dense(np.ascontiguousarray(y.reshape(1, n)), X, k)[0].reshape(m)
    p = -r
    beta = np.zeros(m)
    nr2 = float(O.agg_unary_dense(r.reshape(m, 1), O.SUMSQ, O.ALL)[0])
    target = nr2 * tol ** 2
    i = 0
    while i < maxi and nr2 > target:
        q = O.mmchain_dense(X, p, None, O.XtXv, k) + reg * p
        a = nr2 / float(np.sum(p * q))
        beta = beta + a * p
        r = r + a * q
        old = nr2
        nr2 = float(np.sum(r * r))
        p = -r + (nr2 / old) * p
        i += 1
    return beta, i


@pytest.mark.parametrize("n,m", [(20_000, 200), (100_000, 1000)])
def test_linreg_cg_end_to_end(gctx, n, m):
    from systemml_b200.algorithms import linreg_cg
    X = O.rand_uniform(n * m, seed=0).reshape(n, m)           # genRandData4LinearRegression: X ~ U(-1,1)
    wtrue = O.rand_uniform(m, seed=1)
    y = X @ wtrue + 0.1 * O.rand_uniform(n, seed=2)
    reg, tol = 0.01, 1e-4                                       # scripts/perftest/runLinearRegCG.sh
    exp, it_cpu = cpu_linreg_cg(X, y, reg, tol, 0, O.num_threads())
    dX, dy = gctx.dense(X), gctx.dense(y.reshape(n, 1))
    got, it_gpu = linreg_cg(gctx, dX, dy, reg=reg, tol=tol)
    assert it_gpu == it_cpu
    assert_equal_matrices(exp, got, rel=1e-9, what="LinearRegCG beta %dx%d" % (n, m))
    assert np.max(np.abs(got - wtrue)) < 0.05                   # and it actually solves the regression
    got2, _ = linreg_cg(gctx, dX, dy, reg=reg, tol=tol, fused_tail=False)
    assert_equal_matrices(exp, got2, rel=1e-9, what="unfused tail")
    dX.free()
    dy.free()


def cpu_glm_poisson_log(X, y, reg, outer_iters, inner_iters, inner_tol, k):
    n, m = X.shape
    beta = np.zeros(m)
    for _ in range(outer_iters):
        eta = O.matmult_dense(X, beta.reshape(m, 1), k)[0].reshape(n)
        mu = np.exp(eta)
        grad = O.matmult_dense(np.ascontiguousarray((y - mu).reshape(1, n)), X, k)[0].reshape(m)
        if reg:
            grad = grad + (-reg) * beta
        d = np.zeros(m)
        r = -grad
        p = grad.copy()
        nr2 = float(np.sum(r * r))
        target = nr2 * inner_tol ** 2
        it = 0
        while it < inner_iters and nr2 > target and nr2 > 0:
            q = O.mmchain_dense(X, p, mu, O.XtwXv, k) + reg * p
            a = nr2 / float(np.sum(p * q))
            d = d + a * p
            r = r + a * q
            old = nr2
            nr2 = float(np.sum(r * r))
            p = -r + (nr2 / old) * p
            it += 1
        beta = beta + d
    return beta


def test_glm_poisson_log_end_to_end(gctx):
    from systemml_b200.algorithms import glm_poisson_log
    n, m = 50_000, 100
    X = O.rand_uniform(n * m, seed=1).reshape(n, m)
    btrue = 0.05 * O.rand_uniform(m, seed=5)
    rng = np.random.default_rng(7)
    y = rng.poisson(np.exp(X @ btrue)).astype(np.float64)
    exp = cpu_glm_poisson_log(X, y, 0.0, 6, 15, 1e-10, O.num_threads())
    dX, dy = gctx.dense(X), gctx.dense(y.reshape(n, 1))
    got, inner = glm_poisson_log(gctx, dX, dy, reg=0.0, outer_iters=6, inner_iters=15, inner_tol=1e-10)
    assert_equal_matrices(exp, got, rel=1e-8, what="GLM beta")
    assert np.max(np.abs(got - btrue)) < 0.05
    dX.free()
    dy.free()


def cpu_l2svm(Xd_, Y, reg, tol, maxiter):
    """scripts/algorithms/l2-svm.dml:103-154 in numpy (intercept 0), dense X."""
    n, m = Xd_.shape
    w, Xw = np.zeros(m), np.zeros(n)
    g_old = Xd_.T @ Y
    s = g_old.copy()
    it, cont, obj = 0, True, 0.0
    while cont and it < maxiter:
        step = 0.0
        Xd = Xd_ @ s
        wd, dd = reg * np.sum(w * s), reg * np.sum(s * s)
        while True:
            out = 1 - Y * (Xw + step * Xd)
            sv = (out > 0).astype(np.float64)
            out = out * sv
            g = wd + step * dd - np.sum(out * Y * Xd)
            h = dd + np.sum(Xd * sv * Xd)
            step = step - g / h
            if not (g * g / h >= 1e-10):
                break
        w = w + step * s
        Xw = Xw + step * Xd
        out = 1 - Y * Xw
        out = (out > 0) * out
        obj = 0.5 * np.sum(out * out) + reg / 2 * np.sum(w * w)
        g_new = Xd_.T @ (out * Y) - reg * w
        tmp = np.sum(s * g_old)
        be = np.sum(g_new * g_new) / np.sum(g_old * g_old)
        s = be * s + g_new
        g_old = g_new
        cont = (step * tmp >= tol * obj) and np.sum(s * s) != 0
        it += 1
    return w, it, obj


@pytest.mark.parametrize("sparse", [True, False])
def test_l2svm_end_to_end(gctx, sparse):
    """BASELINE config C3's workload (L2SVM on a CSR matrix) at test size: SpMV / t(X)%*%y on the CSR kernels,
    every n-vector chain of the line search as ONE fused launch; same iterations and weights as the script run in
    numpy, and as the operator-by-operator plan."""
    from systemml_b200.algorithms import l2svm
    n, m = 30_000, 300
    X = O.java_test_matrix(n, m, -1, 1, 0.05 if sparse else 1.0, 7)
    wtrue = O.rand_uniform(m, seed=3)
    Y = np.where(X @ wtrue + 0.05 * O.rand_uniform(n, seed=4) > 0, 1.0, -1.0)
    exp_w, exp_it, exp_obj = cpu_l2svm(X, Y, 1.0, 0.001, 20)
    if sparse:
        rp, ci, va = O.dense_to_csr(X)
        dX = gctx.csr(rp, ci, va, X.shape)
    else:
        dX = gctx.dense(X)
    dY = gctx.dense(Y.reshape(n, 1))
    for fused in (True, False):
        w, it, obj = l2svm(gctx, dX, dY, reg=1.0, tol=0.001, maxiter=20, fused=fused)
        assert it == exp_it, (fused, it, exp_it)
        assert abs(obj - exp_obj) <= 1e-9 * abs(exp_obj)
        assert_equal_matrices(exp_w, w, rel=1e-7, what="l2svm weights (fused=%s)" % fused)
    assert np.mean(np.sign(X @ w) == Y) > 0.9  # and it separates the classes
