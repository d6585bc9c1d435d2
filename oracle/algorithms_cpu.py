"""CPU twins of the five workloads (scripts/algorithms/{LinearRegCG,GLM,l2-svm,Kmeans,MultiLogReg}.dml) in numpy over
the oracle's kernels where the reference has a dedicated kernel (mmchain) and plain numpy elsewhere.

TEST INFRASTRUCTURE, like the rest of oracle/: used by tests/ as the expected result and by bench.py's CPU legs as the
timed host-side twin of each config's end-to-end run.  X may be a scipy.sparse matrix where noted.
"""
import numpy as np

from . import oracle as O


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


def cpu_kmeans_lloyd(X, C, max_iter, eps):
    """scripts/algorithms/Kmeans.dml:130-176 in numpy."""
    sumXsq = np.sum(X ** 2)
    wcss, it, term = np.inf, 0, 0
    while term == 0:
        D = -2 * (X @ C.T) + np.sum(C ** 2, axis=1)[None, :]
        minD = D.min(axis=1, keepdims=True)
        wcss_old, wcss = wcss, sumXsq + np.sum(minD)
        P = (D <= minD).astype(np.float64)
        P = P / P.sum(axis=1, keepdims=True)
        den = P.sum(axis=0)
        C_new = (P.T @ X) / den[:, None]
        it += 1
        if wcss_old - wcss < eps * wcss:
            term = 1
        elif it >= max_iter:
            term = 2
        elif np.sum(den <= 0) > 0:
            term = 3
        else:
            C = C_new
    return C, it, wcss, term


def cpu_multilogreg(X, Y, y_last, reg, tol, maxiter, maxinneriter):
    """scripts/algorithms/MultiLogReg.dml:150-340 in numpy, icpt = 0 (the script's N x (K+1) matrices carried explicitly)."""
    N, D = X.shape
    K = Y.shape[1]
    Yf = np.hstack([Y, y_last])
    eta0, eta1, eta2, sigma1, sigma2, sigma3, psi = 1e-4, 0.25, 0.75, 0.25, 0.5, 4.0, 0.1
    delta = 0.5 * np.sqrt(D) / np.max(np.sqrt(np.sum(X ** 2, axis=1)))
    B = np.zeros((D, K))
    P = np.full((N, K + 1), 1.0 / (K + 1))
    obj = N * np.log(K + 1)
    Grad = X.T @ (P[:, :K] - Y) + reg * B
    norm_Grad = np.sqrt(np.sum(Grad ** 2))
    norm_Grad_initial = norm_Grad
    if maxinneriter == 0:
        maxinneriter = D * K
    it = 1
    converge = (norm_Grad < tol) or (it > maxiter)
    while not converge:
        S = np.zeros((D, K))
        R = -Grad
        V = R.copy()
        delta2 = delta ** 2
        inneriter = 1
        norm_R2 = np.sum(R ** 2)
        innerconverge = np.sqrt(norm_R2) <= psi * norm_Grad
        reached = 0
        while not innerconverge:
            Q = P[:, :K] * (X @ V)
            HV = X.T @ (Q - P[:, :K] * Q.sum(axis=1, keepdims=True)) + reg * V
            alpha = norm_R2 / np.sum(V * HV)
            Snew = S + alpha * V
            if np.sum(Snew ** 2) <= delta2:
                S = Snew
                R = R - alpha * HV
                old = norm_R2
                norm_R2 = np.sum(R ** 2)
                V = R + (norm_R2 / old) * V
                innerconverge = np.sqrt(norm_R2) <= psi * norm_Grad
            else:
                reached = 1
                sv, v2, s2 = np.sum(S * V), np.sum(V ** 2), np.sum(S ** 2)
                rad = np.sqrt(sv ** 2 + v2 * (delta2 - s2))
                alpha = (delta2 - s2) / (sv + rad) if sv >= 0 else (rad - sv) / v2
                S = S + alpha * V
                R = R - alpha * HV
                innerconverge = True
            inneriter += 1
            innerconverge = innerconverge or inneriter > maxinneriter
        gs = np.sum(S * Grad)
        qk = -0.5 * (gs - np.sum(S * R))
        B_new = B + S
        LT = np.hstack([X @ B_new, np.zeros((N, 1))])
        LT = LT - LT.max(axis=1, keepdims=True)
        exp_LT = np.exp(LT)
        P_new = exp_LT / exp_LT.sum(axis=1, keepdims=True)
        obj_new = -np.sum(Yf * LT) + np.sum(np.log(exp_LT.sum(axis=1))) + 0.5 * np.sum(reg * B_new ** 2)
        actred = obj - obj_new
        rho = actred / qk
        snorm = np.sqrt(np.sum(S ** 2))
        if it == 1:
            delta = min(delta, snorm)
        alpha2 = obj_new - obj - gs
        alpha = sigma3 if alpha2 <= 0 else max(sigma1, -0.5 * gs / alpha2)
        if rho < eta0:
            delta = min(max(alpha, sigma1) * snorm, sigma2 * delta)
        elif rho < eta1:
            delta = max(sigma1 * delta, min(alpha * snorm, sigma2 * delta))
        elif rho < eta2:
            delta = max(sigma1 * delta, min(alpha * snorm, sigma3 * delta))
        else:
            delta = max(delta, min(alpha * snorm, sigma3 * delta))
        if rho > eta0:
            B, P = B_new, P_new
            Grad = X.T @ (P[:, :K] - Y) + reg * B
            norm_Grad = np.sqrt(np.sum(Grad ** 2))
            obj = obj_new
        it += 1
        converge = (norm_Grad < tol * norm_Grad_initial) or (it > maxiter) or \
                   (reached == 0 and abs(actred) < (abs(obj) + abs(obj_new)) * 1e-14)
    return B, it - 1, obj
