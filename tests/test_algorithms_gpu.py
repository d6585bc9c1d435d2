"""End-to-end workloads (LinearRegCG, GLM Poisson/log) on the GPU against the same algorithms run on the CPU
with the oracle's kernels: BASELINE configs[0] (100k x 1000) and a smaller shape; beta within 1e-9 relative
(north_star: 'run end-to-end ... within 1e-9 of the CPU reference')."""
import numpy as np
import pytest

from oracle import oracle as O
from oracle.algorithms_cpu import cpu_glm_poisson_log, cpu_kmeans_lloyd, cpu_l2svm, cpu_linreg_cg, cpu_multilogreg  # noqa: F401
from tests.util import assert_equal_matrices, assert_parity

pytestmark = pytest.mark.gpu


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
    assert_equal_matrices(exp, got, rel=1e-9, what="GLM beta")
    assert np.max(np.abs(got - btrue)) < 0.05
    dX.free()
    dy.free()


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
        assert_equal_matrices(exp_w, w, rel=1e-9, what="l2svm weights (fused=%s)" % fused)
    if sparse:   # t(X) %*% y as a gather over the column-ordered copy: same iterations, same weights, and the same bits twice
        runs = [l2svm(gctx, dX, dY, reg=1.0, tol=0.001, maxiter=20, column_copy=True) for _ in range(2)]
        assert dX.colcopy is not None and runs[0][1] == exp_it
        assert_equal_matrices(exp_w, runs[0][0], rel=1e-9, what="l2svm weights (column-ordered copy)")
        assert np.array_equal(runs[0][0], runs[1][0]) and runs[0][2] == runs[1][2]
    assert np.mean(np.sign(X @ w) == Y) > 0.9  # and it separates the classes


@pytest.mark.parametrize("fused", ["step", "chains", False])
def test_kmeans_end_to_end(gctx, fused):
    """BASELINE config C5's workload at test size: Gaussian blobs, k = 12."""
    from systemml_b200.algorithms import kmeans_lloyd
    rng = np.random.default_rng(4)
    n, m, k = 60_000, 20, 12
    centers = rng.uniform(-10, 10, (k, m))
    X = centers[rng.integers(0, k, n)] + rng.normal(size=(n, m))
    C0 = X[rng.choice(n, k, replace=False)].copy()
    expC, exp_it, exp_wcss, exp_term = cpu_kmeans_lloyd(X, C0, 25, 1e-6)
    gotC, it, wcss, term = kmeans_lloyd(gctx, gctx.dense(X), gctx.dense(C0), max_iter=25, eps=1e-6, fused=fused)
    assert (it, term) == (exp_it, exp_term)
    assert abs(wcss - exp_wcss) <= 1e-9 * abs(exp_wcss)
    assert_equal_matrices(expC, gotC, rel=1e-9, what="Kmeans centroids")


def test_multilogreg_end_to_end(gctx):
    """BASELINE config C4's workload at test size (K + 1 = 5 classes): same trust-region path, same B."""
    from systemml_b200.algorithms import multilogreg
    rng = np.random.default_rng(3)
    N, D, K = 30_000, 40, 4
    X = rng.uniform(-1, 1, (N, D))
    Wt = rng.normal(size=(D, K + 1))
    labels = np.argmax(X @ Wt + 0.5 * rng.gumbel(size=(N, K + 1)), axis=1)
    Yfull = np.eye(K + 1)[labels]
    Y, y_last = np.ascontiguousarray(Yfull[:, :K]), np.ascontiguousarray(Yfull[:, K:])
    expB, exp_it, exp_obj = cpu_multilogreg(X, Y, y_last, 1.0, 1e-6, 12, 0)
    gotB, it, obj = multilogreg(gctx, gctx.dense(X), gctx.dense(Y), gctx.dense(y_last), reg=1.0, tol=1e-6, maxiter=12)
    assert it == exp_it
    assert abs(obj - exp_obj) <= 1e-9 * abs(exp_obj)
    assert_equal_matrices(expB, gotB, rel=1e-9, what="MultiLogReg B")
    pred = np.argmax(np.hstack([X @ gotB, np.zeros((N, 1))]), axis=1)
    assert np.mean(pred == labels) > 0.6


@pytest.mark.parametrize("n,m,k", [(5000, 100, 50), (63, 20, 12), (4097, 22, 64), (1000, 6, 3), (20_000, 100, 50)])
def test_kmeans_one_pass_step_kernel(gctx, n, m, k):
    """sysml_b200_kmeans_step against the script's step in numpy (Kmeans.dml:146-166): sums = t(P) %*% X, counts =
    colSums(P), sum(minD); duplicate centroids give exact ties, which must share their records equally; shapes cover a
    ragged last tile, fewer rows than a tile, m not a multiple of 4 and the largest k."""
    import ctypes as C
    rng = np.random.default_rng(n + m + k)
    centers = rng.uniform(-5, 5, (k, m))
    X = centers[rng.integers(0, k, n)] + 0.5 * rng.normal(size=(n, m))
    Cm = centers + 0.1 * rng.normal(size=(k, m))
    if k >= 3:
        Cm[k - 1] = Cm[0]                      # an exact tie for every record closest to centroid 0
    c2 = np.sum(Cm ** 2, axis=1)
    D = -2 * (X @ Cm.T) + c2[None, :]
    minD = D.min(axis=1, keepdims=True)
    P = (D <= minD).astype(np.float64)
    P = P / P.sum(axis=1, keepdims=True)
    dX, dC = gctx.dense(X), gctx.dense(Cm)
    dc2 = gctx.empty(k, 1)
    gctx.call("sysml_b200_agg_unary_dense", C.c_void_p(dC.ptr), k, m, 1, 1, C.c_void_p(dc2.ptr))     # rowSums(C^2)
    sums, cnt, smin = gctx.empty(k, m), gctx.empty(k, 1), gctx.empty(1, 1)
    runs = []
    for _ in range(2):
        gctx.call("sysml_b200_kmeans_step", C.c_void_p(dX.ptr), C.c_void_p(dC.ptr), C.c_void_p(dc2.ptr), n, m, k, C.c_void_p(sums.ptr),
                  C.c_void_p(cnt.ptr), C.c_void_p(smin.ptr))
        runs.append((sums.to_host(), cnt.to_host().reshape(-1), float(smin.to_host()[0, 0])))
    assert np.array_equal(runs[0][0], runs[1][0]) and np.array_equal(runs[0][1], runs[1][1]) and runs[0][2] == runs[1][2], "not reproducible"
    got_sums, got_cnt, got_smin = runs[0]
    assert_equal_matrices(P.sum(axis=0), got_cnt, what="kmeans counts %dx%d k=%d" % (n, m, k))
    if k >= 3:
        assert got_cnt[0] == got_cnt[k - 1] and got_cnt[0] > 0, "tied centroids share their records"
    assert_parity(P.T @ X, got_sums, np.abs(P).T @ np.abs(X), n, what="kmeans sums %dx%d k=%d" % (n, m, k))
    assert abs(got_smin - float(np.sum(minD))) <= 1e-9 * abs(float(np.sum(minD)))
    for o in (dX, dC, dc2, sums, cnt, smin):
        o.free()
