"""The two workloads BASELINE.json's north_star names, written against the library the way SystemML's
runtime executes the DML scripts: host control flow and scalars on the host, every matrix op on the GPU.

  linreg_cg          <- scripts/algorithms/LinearRegCG.dml:160-200 (icpt = 0)
  glm_poisson_log    <- the mmchain-dominated core of scripts/algorithms/GLM.dml for dfam=1 (Poisson), link=log:
                        outer Newton / IRLS iterations, each solved with conjugate gradients whose only O(n m)
                        op is  t(X) %*% (w * (X %*% p))  (GLM.dml:1011), plus  X %*% beta  (:381),
                        t(X) %*% g  (:365) and the cellwise  exp / * / -  chain over n-vectors (:702-706).
                        GLM.dml wraps this in a trust-region step control that is host-side scalar logic; the
                        plain Newton-CG below keeps the device work identical and the test comparable.

With a Communicator, X / y are this rank's row shard and every m-sized partial result is summed over the
ranks; all ranks return the same beta.
"""
import ctypes as C

import numpy as np

from . import lib_matrix as L
from .gpu_context import GPUObject


def _p(o):
    return C.c_void_p(o.ptr)


def _allreduce(comm, obj, length):
    if comm is not None:
        comm.allreduce_sum(obj.ptr, length)


def _scalar(g, obj):
    return float(g.d2h(obj.ptr, (1,))[0])


def linreg_cg(g, X, y, reg=1e-6, tol=1e-6, maxi=0, comm=None, fused_tail=True, log=None):
    """LinearRegCG.dml with icpt=0.  X: GPUObject n x m (dense), y: GPUObject n x 1.  Returns (beta ndarray, iterations)."""
    n, m = X.rows, X.cols
    if maxi == 0:
        maxi = m
    r = GPUObject.allocate_dense(g, m, 1)
    # r = - t(X) %*% y     (compiled as t(t(y) %*% X): the VM kernel)
    g.call("sysml_b200_matmult_dense", _p(X), _p(y), _p(r), n, m, n, 1, 1, 0)
    _allreduce(comm, r, m)
    g.call("sysml_b200_binary_ms", _p(r), -1.0, _p(r), m, L.BINARY_OPS["*"], 0)
    p = GPUObject.allocate_dense(g, m, 1)
    g.call("sysml_b200_binary_ms", _p(r), -1.0, _p(p), m, L.BINARY_OPS["*"], 0)   # p = -r
    beta = GPUObject.allocate_dense(g, m, 1)
    g.call("sysml_b200_fill", _p(beta), m, 0.0)
    q = GPUObject.allocate_dense(g, m, 1)
    nr2_d = GPUObject.allocate_dense(g, 1, 1)
    g.call("sysml_b200_agg_unary_dense", _p(r), m, 1, L.AGG_SUMSQ, L.DIR_ALL, _p(nr2_d))  # norm_r2 = sum(r^2)
    nr2 = _scalar(g, nr2_d)
    nr2_init = nr2
    target = nr2_init * tol ** 2
    i = 0
    while i < maxi and nr2 > target:
        # q = t(X) %*% (X %*% p)
        if comm is not None:
            g.call("sysml_b200_mmchain_dense_allreduce", comm.handle, _p(X), _p(p), None, _p(q), n, m, L.XtXv)
        else:
            g.call("sysml_b200_mmchain_dense", _p(X), _p(p), None, _p(q), n, m, L.XtXv)
        if fused_tail:
            # q += lambda*p; a = nr2/sum(p*q); beta += a*p; r += a*q; nr2 = sum(r^2); p = -r + (nr2/old)*p
            g.call("sysml_b200_cg_vector_update", _p(p), _p(q), _p(r), _p(beta), _p(nr2_d), m, float(reg))
            nr2 = _scalar(g, nr2_d)
        else:
            ph, qh, rh, bh = (g.d2h(o.ptr, (m,)) for o in (p, q, r, beta))
            qh = qh + reg * ph
            a = nr2 / float(np.sum(ph * qh))
            bh = bh + a * ph
            rh = rh + a * qh
            old = nr2
            nr2 = float(np.sum(rh * rh))
            ph = -rh + (nr2 / old) * ph
            for o, h in ((p, ph), (r, rh), (beta, bh)):
                g.h2d(o.ptr, h)
        i += 1
        if log is not None:
            log.append((i, (nr2 / nr2_init) ** 0.5))
    out = beta.to_host().reshape(-1)
    for o in (r, p, beta, q, nr2_d):
        o.free()
    return out, i


def glm_poisson_log(g, X, y, reg=0.0, outer_iters=8, inner_iters=20, inner_tol=1e-10, comm=None):
    """Poisson regression with log link by Newton / IRLS; every inner CG step is mmchain XtwXv.
    Returns (beta ndarray, list of inner iteration counts)."""
    n, m = X.rows, X.cols
    alloc = lambda r_, c_: GPUObject.allocate_dense(g, r_, c_)
    beta, eta, mu, resid = alloc(m, 1), alloc(n, 1), alloc(n, 1), alloc(n, 1)
    grad, d, r, p, q, nr2_d = alloc(m, 1), alloc(m, 1), alloc(m, 1), alloc(m, 1), alloc(m, 1), alloc(1, 1)
    g.call("sysml_b200_fill", _p(beta), m, 0.0)
    inner_counts = []
    for _ in range(outer_iters):
        g.call("sysml_b200_matmult_dense", _p(X), _p(beta), _p(eta), n, m, m, 1, 0, 0)        # eta = X %*% beta
        g.call("sysml_b200_unary", _p(eta), _p(mu), n, L.UNARY_OPS["exp"])                    # mu = exp(eta) (= weights w)
        g.call("sysml_b200_binary_mm", _p(y), _p(mu), _p(resid), n, 1, 0, L.BINARY_OPS["-"])  # y - mu
        g.call("sysml_b200_matmult_dense", _p(X), _p(resid), _p(grad), n, m, n, 1, 1, 0)      # t(X) %*% (y - mu)
        _allreduce(comm, grad, m)
        if reg:
            g.call("sysml_b200_axpy", _p(grad), _p(beta), -float(reg), _p(grad), m, 1, 0)
        # solve (t(X) W X + reg I) d = grad by CG, d0 = 0
        g.call("sysml_b200_fill", _p(d), m, 0.0)
        g.call("sysml_b200_binary_ms", _p(grad), -1.0, _p(r), m, L.BINARY_OPS["*"], 0)        # r = -grad
        g.call("sysml_b200_d2d", _p(p), _p(grad), 8 * m)                                      # p = -r = grad
        g.call("sysml_b200_agg_unary_dense", _p(r), m, 1, L.AGG_SUMSQ, L.DIR_ALL, _p(nr2_d))
        nr2 = _scalar(g, nr2_d)
        target = nr2 * inner_tol ** 2
        it = 0
        while it < inner_iters and nr2 > target and nr2 > 0.0:
            if comm is not None:
                g.call("sysml_b200_mmchain_dense_allreduce", comm.handle, _p(X), _p(p), _p(mu), _p(q), n, m, L.XtwXv)
            else:
                g.call("sysml_b200_mmchain_dense", _p(X), _p(p), _p(mu), _p(q), n, m, L.XtwXv)
            g.call("sysml_b200_cg_vector_update", _p(p), _p(q), _p(r), _p(d), _p(nr2_d), m, float(reg))
            nr2 = _scalar(g, nr2_d)
            it += 1
        inner_counts.append(it)
        g.call("sysml_b200_binary_mm", _p(beta), _p(d), _p(beta), m, 1, 0, L.BINARY_OPS["+"])  # beta += d
    out = beta.to_host().reshape(-1)
    for o in (beta, eta, mu, resid, grad, d, r, p, q, nr2_d):
        o.free()
    return out, inner_counts


def l2svm(g, X, Y, reg=1.0, tol=0.001, maxiter=100, fused=True, log=None):
    """scripts/algorithms/l2-svm.dml:103-154 (intercept = 0, labels already -1/+1): non-linear CG on the squared-hinge
    primal with an exact Newton line search.  X: GPUObject n x m, dense or CSR (BASELINE config C3 is CSR 50M x 10k);
    Y: GPUObject n x 1.  Per outer iteration the O(nnz) work is X %*% s and t(X) %*% (out * Y); everything else is
    chains over n-vectors, each evaluated in one pass by fusedCellwise (fused=True) or operator by operator
    (fused=False, the reference's plan).  Returns (w ndarray, iterations, final objective)."""
    n, m = X.rows, X.cols
    E = L.Expr
    alloc = lambda r_, c_: GPUObject.allocate_dense(g, r_, c_)
    sparse = X.csr is not None

    def x_times(v, out):        # out = X %*% v
        if sparse:
            g.call("sysml_b200_matmult_csr_dense", C.c_void_p(X.csr.rowPtr), C.c_void_p(X.csr.colInd), C.c_void_p(X.csr.val),
                   _p(v), _p(out), n, m, m, 1, 0)
        else:
            g.call("sysml_b200_matmult_dense", _p(X), _p(v), _p(out), n, m, m, 1, 0, 0)

    def xt_times(yv, out):      # out = t(X) %*% yv
        if sparse:
            g.call("sysml_b200_matmult_csr_dense", C.c_void_p(X.csr.rowPtr), C.c_void_p(X.csr.colInd), C.c_void_p(X.csr.val),
                   _p(yv), _p(out), n, m, n, 1, 1)
        else:
            g.call("sysml_b200_matmult_dense", _p(X), _p(yv), _p(out), n, m, n, 1, 1, 0)

    def fsum(inputs, expr, consts=()):
        return L.fusedCellwise(g, "spoof", inputs, expr, consts, "sum")

    def dot(a, b):
        return fsum([a, b], E.input(0) * E.input(1))

    w, Xw, Xd, s, g_old, g_new = alloc(m, 1), alloc(n, 1), alloc(n, 1), alloc(m, 1), alloc(m, 1), alloc(m, 1)
    t1, t2, t3 = alloc(n, 1), alloc(n, 1), alloc(n, 1)
    g.call("sysml_b200_fill", _p(w), m, 0.0)
    g.call("sysml_b200_fill", _p(Xw), n, 0.0)
    xt_times(Y, g_old)
    g.call("sysml_b200_d2d", _p(s), _p(g_old), 8 * m)
    # inputs of the n-vector chains: 0 = Y, 1 = Xw, 2 = Xd; constants: 0 = 1.0, 1 = 0.0, 2 = step_sz
    out_e = E.const(0) - E.input(0) * (E.input(1) + E.const(2) * E.input(2))          # 1 - Y * (Xw + step_sz*Xd)
    g_e = E.let(out_e, lambda o: o * o.binary(E.const(1), ">") * E.input(0) * E.input(2))   # (out*sv) * Y * Xd
    h_e = E.input(2) * out_e.binary(E.const(1), ">") * E.input(2)                     # Xd * sv * Xd
    it, cont, obj = 0, True, 0.0
    while cont and it < maxiter:
        step = 0.0
        x_times(s, Xd)
        wd = reg * dot(w, s)
        dd = reg * dot(s, s)
        while True:
            if fused:
                gsum = fsum([Y, Xw, Xd], g_e, (1.0, 0.0, step))
                hsum = fsum([Y, Xw, Xd], h_e, (1.0, 0.0, step))
            else:
                g.call("sysml_b200_axpy", _p(Xw), _p(Xd), step, _p(t1), n, 1, 0)                       # tmp_Xw
                g.call("sysml_b200_binary_mm", _p(Y), _p(t1), _p(t1), n, 1, 0, L.BINARY_OPS["1-*"])    # out = 1 - Y*tmp_Xw
                g.call("sysml_b200_binary_ms", _p(t1), 0.0, _p(t2), n, L.BINARY_OPS[">"], 0)           # sv
                g.call("sysml_b200_binary_mm", _p(t1), _p(t2), _p(t1), n, 1, 0, L.BINARY_OPS["*"])     # out * sv
                g.call("sysml_b200_binary_mm", _p(t1), _p(Y), _p(t1), n, 1, 0, L.BINARY_OPS["*"])
                g.call("sysml_b200_binary_mm", _p(t1), _p(Xd), _p(t1), n, 1, 0, L.BINARY_OPS["*"])
                gsum = L.unaryAggregate(g, "uak+", t1, "uak+")
                g.call("sysml_b200_binary_mm", _p(Xd), _p(t2), _p(t3), n, 1, 0, L.BINARY_OPS["*"])
                g.call("sysml_b200_binary_mm", _p(t3), _p(Xd), _p(t3), n, 1, 0, L.BINARY_OPS["*"])
                hsum = L.unaryAggregate(g, "uak+", t3, "uak+")
            gg = wd + step * dd - gsum
            hh = dd + hsum
            step = step - gg / hh
            if not (gg * gg / hh >= 1e-10):
                break
        g.call("sysml_b200_axpy", _p(w), _p(s), step, _p(w), m, 1, 0)          # w += step_sz * s
        g.call("sysml_b200_axpy", _p(Xw), _p(Xd), step, _p(Xw), n, 1, 0)       # Xw += step_sz * Xd
        # out = sv * (1 - Y*Xw); obj = 0.5*sum(out*out) + reg/2*sum(w*w); t1 = out * Y
        hinge = E.let(E.const(0) - E.input(0) * E.input(1), lambda o: o.binary(E.const(1), ">") * o)
        if fused:
            osum = fsum([Y, Xw], E.let(hinge, lambda o: o * o), (1.0, 0.0))
            oy = L.fusedCellwise(g, "spoof", [Y, Xw], hinge * E.input(0), (1.0, 0.0))
        else:
            g.call("sysml_b200_binary_mm", _p(Y), _p(Xw), _p(t1), n, 1, 0, L.BINARY_OPS["1-*"])
            g.call("sysml_b200_binary_ms", _p(t1), 0.0, _p(t2), n, L.BINARY_OPS[">"], 0)
            g.call("sysml_b200_binary_mm", _p(t2), _p(t1), _p(t1), n, 1, 0, L.BINARY_OPS["*"])
            osum = L.unaryAggregate(g, "uasqk+", t1, "uasqk+")
            oy = alloc(n, 1)
            g.call("sysml_b200_binary_mm", _p(t1), _p(Y), _p(oy), n, 1, 0, L.BINARY_OPS["*"])
        obj = 0.5 * osum + reg / 2.0 * dot(w, w)
        xt_times(oy, g_new)
        oy.free()
        g.call("sysml_b200_axpy", _p(g_new), _p(w), -float(reg), _p(g_new), m, 1, 0)     # g_new = t(X)(out*Y) - reg*w
        tmp = dot(s, g_old)
        be = dot(g_new, g_new) / dot(g_old, g_old)
        g.call("sysml_b200_axpy", _p(g_new), _p(s), be, _p(s), m, 1, 0)                  # s = be*s + g_new
        g.call("sysml_b200_d2d", _p(g_old), _p(g_new), 8 * m)
        cont = (step * tmp >= tol * obj) and dot(s, s) != 0.0
        it += 1
        if log is not None:
            log.append((it, obj))
    res = w.to_host().reshape(-1)
    for o in (w, Xw, Xd, s, g_old, g_new, t1, t2, t3):
        o.free()
    return res, it, obj
