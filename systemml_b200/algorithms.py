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
        if fused_tail and m <= 8192:
            # the whole iteration in two launches: q = t(X) %*% (X %*% p) [summed over ranks], then the m-vector tail
            g.call("sysml_b200_cg_step_dense", comm.handle if comm is not None else None, _p(X), None, _p(p), _p(q), _p(r), _p(beta),
                   _p(nr2_d), n, m, L.XtXv, float(reg))
            nr2 = _scalar(g, nr2_d)
            i += 1
            if log is not None:
                log.append((i, (nr2 / nr2_init) ** 0.5))
            continue
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


def glm_poisson_log(g, X, y, reg=0.0, outer_iters=8, inner_iters=20, inner_tol=1e-10, comm=None, fused_outer=True):
    """Poisson regression with log link by Newton / IRLS; every inner CG step is mmchain XtwXv.
    fused_outer: the outer step's  eta = X %*% beta; mu = exp(eta); grad = t(X) %*% (y - mu)  as ONE pass over X
    (mmchain with a row function) instead of two products and two cellwise passes.
    Returns (beta ndarray, list of inner iteration counts)."""
    n, m = X.rows, X.cols
    alloc = lambda r_, c_: GPUObject.allocate_dense(g, r_, c_)
    beta, eta, mu, resid = alloc(m, 1), alloc(n, 1), alloc(n, 1), alloc(n, 1)
    grad, d, r, p, q, nr2_d = alloc(m, 1), alloc(m, 1), alloc(m, 1), alloc(m, 1), alloc(m, 1), alloc(1, 1)
    g.call("sysml_b200_fill", _p(beta), m, 0.0)
    inner_counts = []
    for _ in range(outer_iters):
        if fused_outer and m <= 8192:
            E = L.Expr
            mu_e = E.input(0).unary("exp")
            L.mmchainRowFn(g, "mmchain", X, beta, [E.input(1) - mu_e, mu_e], w=y, out=grad, side=mu)  # grad = t(X) %*% (y - exp(X beta)); mu
        else:
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
            if m <= 8192:
                g.call("sysml_b200_cg_step_dense", comm.handle if comm is not None else None, _p(X), _p(mu), _p(p), _p(q), _p(r), _p(d),
                       _p(nr2_d), n, m, L.XtwXv, float(reg))
            else:
                if comm is not None:
                    g.call("sysml_b200_mmchain_dense", _p(X), _p(p), _p(mu), _p(q), n, m, L.XtwXv)
                    comm.allreduce_sum(q.ptr, m)
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


def l2svm(g, X, Y, reg=1.0, tol=0.001, maxiter=100, fused=True, log=None, column_copy=False, comm=None):
    """scripts/algorithms/l2-svm.dml:103-154 (intercept = 0, labels already -1/+1): non-linear CG on the squared-hinge
    primal with an exact Newton line search.  X: GPUObject n x m, dense or CSR (BASELINE config C3 is CSR 50M x 10k);
    Y: GPUObject n x 1.  Per outer iteration the O(nnz) work is X %*% s and t(X) %*% (out * Y); everything else is
    chains over n-vectors, each evaluated in one pass by fusedCellwise (fused=True) or operator by operator
    (fused=False, the reference's plan).  column_copy=True runs t(X) %*% y as a gather over the column-ordered copy of a
    CSR X (csrc/psell.cu; deterministic), False as the scatter kernel.  comm: X and Y are this rank's row shard:
    X %*% s needs no exchange, t(X) %*% (out*Y) and every sum over an n-vector are a local result + one exchange, the
    m-vectors are replicated; every rank returns the same w (fused plan only).  Returns (w ndarray, iterations, final objective)."""
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

    if sparse and column_copy:
        X.transposed_uses = max(X.transposed_uses, 1)   # X is constant across the iterations: build the copy at the first use

    def xt_times(yv, out):      # out = t(X) %*% yv
        pc = X.columnOrderedCopy() if (sparse and column_copy) else None
        if pc is not None:          # gather over the column-ordered copy (bit-identical run to run)
            pc.vm(yv.ptr, out.ptr)
        elif sparse:
            g.call("sysml_b200_matmult_csr_dense", C.c_void_p(X.csr.rowPtr), C.c_void_p(X.csr.colInd), C.c_void_p(X.csr.val),
                   _p(yv), _p(out), n, m, n, 1, 1)
        else:
            g.call("sysml_b200_matmult_dense", _p(X), _p(yv), _p(out), n, m, n, 1, 1, 0)

    fs, fs4 = alloc(1, 1), alloc(4, 1)

    if comm is not None and not fused:
        raise ValueError("l2svm: the row-sharded plan runs the fused chains")

    def fsum(inputs, expr, consts=()):
        if comm is not None and inputs[0].rows == n:      # a sum over row-sharded n-vectors: local sum, then one exchange
            return fsums(inputs, [expr], consts)[0]
        return L.fusedCellwise(g, "spoof", inputs, expr, consts, "sum", out=fs)

    def fsums(inputs, exprs, consts=()):
        # several sums over the same operands in one pass (the code generator's multi-aggregate template); with fused=False
        # one pass per sum, as separate instructions would run
        if fused:
            if comm is not None and inputs[0].rows == n:
                L.fusedMultiAggregate(g, "spoof", inputs, exprs, consts, out=fs4, read_back=False)
                _allreduce(comm, fs4, len(exprs))
                return [float(v) for v in g.d2h(fs4.ptr, (len(exprs),))]
            return L.fusedMultiAggregate(g, "spoof", inputs, exprs, consts, out=fs4)
        return [fsum(inputs, e, consts) for e in exprs]

    def dot(a, b):
        return fsum([a, b], E.input(0) * E.input(1))

    w, Xw, Xd, s, g_old, g_new = alloc(m, 1), alloc(n, 1), alloc(n, 1), alloc(m, 1), alloc(m, 1), alloc(m, 1)
    t1, t2, t3 = (alloc(n, 1), alloc(n, 1), alloc(n, 1)) if not fused else (None, None, None)   # temporaries of the unfused plan
    oy = alloc(n, 1)      # out * Y: one buffer for all iterations (a 400 MB cudaMalloc + cudaFree per iteration costs more than the kernel)
    g.call("sysml_b200_fill", _p(w), m, 0.0)
    g.call("sysml_b200_fill", _p(Xw), n, 0.0)
    xt_times(Y, g_old)
    _allreduce(comm, g_old, m)
    g.call("sysml_b200_d2d", _p(s), _p(g_old), 8 * m)
    # inputs of the n-vector chains: 0 = Y, 1 = Xw, 2 = Xd; constants: 0 = 1.0, 1 = 0.0, 2 = step_sz
    out_e = E.const(0) - E.input(0) * (E.input(1) + E.const(2) * E.input(2))          # 1 - Y * (Xw + step_sz*Xd)
    g_e = E.let(out_e, lambda o: o * o.binary(E.const(1), ">") * E.input(0) * E.input(2))   # (out*sv) * Y * Xd
    h_e = E.input(2) * out_e.binary(E.const(1), ">") * E.input(2)                     # Xd * sv * Xd
    it, cont, obj = 0, True, 0.0
    while cont and it < maxiter:
        step = 0.0
        x_times(s, Xd)
        wd, dd = fsums([w, s], [E.input(0) * E.input(1), E.input(1) * E.input(1)])
        wd, dd = reg * wd, reg * dd
        while True:
            if fused:
                gsum, hsum = fsums([Y, Xw, Xd], [g_e, h_e], (1.0, 0.0, step))
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
            L.fusedCellwise(g, "spoof", [Y, Xw], hinge * E.input(0), (1.0, 0.0), out=oy)
        else:
            g.call("sysml_b200_binary_mm", _p(Y), _p(Xw), _p(t1), n, 1, 0, L.BINARY_OPS["1-*"])
            g.call("sysml_b200_binary_ms", _p(t1), 0.0, _p(t2), n, L.BINARY_OPS[">"], 0)
            g.call("sysml_b200_binary_mm", _p(t2), _p(t1), _p(t1), n, 1, 0, L.BINARY_OPS["*"])
            osum = L.unaryAggregate(g, "uasqk+", t1, "uasqk+")
            g.call("sysml_b200_binary_mm", _p(t1), _p(Y), _p(oy), n, 1, 0, L.BINARY_OPS["*"])
        obj = 0.5 * osum + reg / 2.0 * dot(w, w)
        xt_times(oy, g_new)
        _allreduce(comm, g_new, m)
        g.call("sysml_b200_axpy", _p(g_new), _p(w), -float(reg), _p(g_new), m, 1, 0)     # g_new = t(X)(out*Y) - reg*w
        tmp, gn2, go2 = fsums([s, g_old, g_new], [E.input(0) * E.input(1), E.input(2) * E.input(2), E.input(1) * E.input(1)])
        be = gn2 / go2
        g.call("sysml_b200_axpy", _p(g_new), _p(s), be, _p(s), m, 1, 0)                  # s = be*s + g_new
        g.call("sysml_b200_d2d", _p(g_old), _p(g_new), 8 * m)
        cont = (step * tmp >= tol * obj) and dot(s, s) != 0.0
        it += 1
        if log is not None:
            log.append((it, obj))
    res = w.to_host().reshape(-1)
    for o in (w, Xw, Xd, s, g_old, g_new, t1, t2, t3, fs, fs4, oy):
        if o is not None:
            o.free()
    return res, it, obj


def _kmeans_lloyd_one_pass(g, X, C0, max_iter, eps, log, comm=None):
    """The same iterations with the whole step in one pass over X (sysml_b200_kmeans_step, csrc/kmeans.cu): distances on the
    FP64 tensor cores, row minimum with the script's tie rule, centroid sums, counts and sum(minD) -- D, minD and P never
    exist in memory.  The k x m algebra around it (rowSums(C^2), the division by the counts) stays as instructions."""
    n, m, k = X.rows, X.cols, C0.rows
    alloc = lambda r_, c_: GPUObject.allocate_dense(g, r_, c_)
    Cc, Cn, c2, den, s1 = alloc(k, m), alloc(k, m), alloc(k, 1), alloc(k, 1), alloc(1, 1)
    g.call("sysml_b200_d2d", _p(Cc), _p(C0), 8 * k * m)
    g.call("sysml_b200_agg_unary_dense", _p(X), n, m, L.AGG_SUMSQ, L.DIR_ALL, _p(s1))
    _allreduce(comm, s1, 1)           # X row-sharded over the ranks of comm: every n-sum is a local sum + one exchange
    sumXsq = _scalar(g, s1)
    wcss, it, term = float("inf"), 0, 0
    while term == 0:
        g.call("sysml_b200_agg_unary_dense", _p(Cc), k, m, L.AGG_SUMSQ, L.DIR_ROW, _p(c2))           # rowSums(C^2)
        try:
            g.call("sysml_b200_kmeans_step", _p(X), _p(Cc), _p(c2), n, m, k, _p(Cn), _p(den), _p(s1))   # t(P)%*%X, colSums(P), sum(minD)
        except Exception:
            for o in (Cc, Cn, c2, den, s1):
                o.free()
            raise
        _allreduce(comm, Cn, k * m)   # t(P) %*% X: k x m doubles (40 KB at C5)
        _allreduce(comm, den, k)      # colSums(P)
        _allreduce(comm, s1, 1)       # sum(minD)
        wcss_old, wcss = wcss, sumXsq + _scalar(g, s1)
        g.call("sysml_b200_binary_mm", _p(Cn), _p(den), _p(Cn), k, m, L.BCAST_COL, L.BINARY_OPS["/"])  # / t(P_denom)
        it += 1
        if log is not None:
            log.append((it, wcss))
        if wcss_old - wcss < eps * wcss:
            term = 1
        elif it >= max_iter:
            term = 2
        elif float(np.sum(den.to_host() <= 0)) > 0:
            term = 3
        else:
            Cc, Cn = Cn, Cc
    out = Cc.to_host()
    for o in (Cc, Cn, c2, den, s1):
        o.free()
    return out, it, wcss, term


def kmeans_lloyd(g, X, C0, max_iter=20, eps=1e-6, fused=True, log=None, comm=None):
    """One run of the Lloyd iterations of scripts/algorithms/Kmeans.dml:130-176 from given initial centroids
    (BASELINE config C5: X 10M x 100, k = 50).  X: GPUObject n x m dense, C0: GPUObject k x m.
    Per iteration: D = -2*(X %*% t(C)) + t(rowSums(C^2)); minD = rowMins(D); P = (D <= minD) / rowSums(D <= minD);
    C_new = (t(P) %*% X) / t(colSums(P)).  Returns (C ndarray, iterations, wcss, termination code 1/2/3).
    fused: True = the one-pass step kernel when the shape fits it, else the fused cellwise chains; "chains" = the chains;
    False = one launch per operator (the reference's plan); "step" = the one-pass kernel or an error.
    comm: X is this rank's row shard; the k x m sums, the k counts and sum(minD) are exchanged once per iteration
    (t(P) %*% X "yes -- one exchange"); every rank returns the same centroids. One-pass plan only."""
    n, m, k = X.rows, X.cols, C0.rows
    E = L.Expr
    alloc = lambda r_, c_: GPUObject.allocate_dense(g, r_, c_)
    if comm is not None:
        fused = "step"
    if fused == "step":
        return _kmeans_lloyd_one_pass(g, X, C0, max_iter, eps, log, comm)
    if fused is True:
        from . import native
        try:
            return _kmeans_lloyd_one_pass(g, X, C0, max_iter, eps, log)
        except native.NativeError as e:   # the shape does not fit the one-pass kernel: the chains do the same step
            if e.code != native.E_UNSUPPORTED:
                raise
    fused = bool(fused)   # "chains" -> True
    Cc, Cn, XC, P = alloc(k, m), alloc(k, m), alloc(n, k), alloc(n, k)
    D = None if fused else alloc(n, k)   # the fused plan never materialises D
    c2, minD, rs, den, s1 = alloc(k, 1), alloc(n, 1), alloc(n, 1), alloc(k, 1), alloc(1, 1)
    g.call("sysml_b200_d2d", _p(Cc), _p(C0), 8 * k * m)
    g.call("sysml_b200_agg_unary_dense", _p(X), n, m, L.AGG_SUMSQ, L.DIR_ALL, _p(s1))
    sumXsq = _scalar(g, s1)
    c2_row = GPUObject.wrap(g, c2.ptr, 1, k)      # t(rowSums(C^2)): the same k doubles seen as a row vector
    den_col = den                                  # t(colSums(P)) as a k x 1 column
    wcss, it, term = float("inf"), 0, 0
    while term == 0:
        g.call("sysml_b200_matmult_dense", _p(X), _p(Cc), _p(XC), n, m, k, m, 0, 1)                 # X %*% t(C)
        g.call("sysml_b200_agg_unary_dense", _p(Cc), k, m, L.AGG_SUMSQ, L.DIR_ROW, _p(c2))           # rowSums(C^2)
        if fused:
            # D is never written: every consumer re-evaluates -2*XC + t(c2) from XC on the fly
            d_e = E.const(0) * E.input(0) + E.input(1)
            L.fusedCellwise(g, "spoof", [XC, c2_row], d_e, (-2.0,), "rowMins", out=minD)              # rowMins(D)
        else:
            g.call("sysml_b200_binary_ms", _p(XC), -2.0, _p(D), n * k, L.BINARY_OPS["*"], 1)
            g.call("sysml_b200_binary_mm", _p(D), _p(c2), _p(D), n, k, L.BCAST_ROW, L.BINARY_OPS["+"])
            g.call("sysml_b200_agg_unary_dense", _p(D), n, k, L.AGG_MIN, L.DIR_ROW, _p(minD))        # rowMins(D)
        g.call("sysml_b200_agg_unary_dense", _p(minD), n, 1, L.AGG_SUM, L.DIR_ALL, _p(s1))
        wcss_old, wcss = wcss, sumXsq + _scalar(g, s1)
        if fused:
            le = d_e.binary(E.input(2), "<=")                                                       # D <= minD
            L.fusedCellwise(g, "spoof", [XC, c2_row, minD], le, (-2.0,), "rowSums", out=rs)
            # P / rowSums(P) with P in {0, 1}: 1/rs once per row, then a product per cell -- the same bits as the
            # script's division (0/rs = 0*(1/rs) = 0, 1/rs = 1*(1/rs)) at a fraction of its cost
            g.call("sysml_b200_binary_ms", _p(rs), 1.0, _p(rs), n, L.BINARY_OPS["/"], 1)
            L.fusedCellwise(g, "spoof", [XC, c2_row, minD, rs], le * E.input(3), (-2.0,), out=P)
        else:
            g.call("sysml_b200_binary_mm", _p(D), _p(minD), _p(P), n, k, L.BCAST_COL, L.BINARY_OPS["<="])
            g.call("sysml_b200_agg_unary_dense", _p(P), n, k, L.AGG_SUM, L.DIR_ROW, _p(rs))
            g.call("sysml_b200_binary_mm", _p(P), _p(rs), _p(P), n, k, L.BCAST_COL, L.BINARY_OPS["/"])
        g.call("sysml_b200_agg_unary_dense", _p(P), n, k, L.AGG_SUM, L.DIR_COL, _p(den))             # colSums(P)
        g.call("sysml_b200_matmult_dense", _p(P), _p(X), _p(Cn), n, k, n, m, 1, 0)                   # t(P) %*% X
        g.call("sysml_b200_binary_mm", _p(Cn), _p(den_col), _p(Cn), k, m, L.BCAST_COL, L.BINARY_OPS["/"])
        it += 1
        if log is not None:
            log.append((it, wcss))
        if wcss_old - wcss < eps * wcss:
            term = 1
        elif it >= max_iter:
            term = 2
        elif float(np.sum(den.to_host() <= 0)) > 0:
            term = 3
        else:
            Cc, Cn = Cn, Cc
    out = Cc.to_host()
    for o in (Cc, Cn, XC, D, P, c2, minD, rs, den, s1):
        if o is not None:
            o.free()
    return out, it, wcss, term


def multilogreg(g, X, Y, y_last, reg=0.0, tol=1e-6, maxiter=100, maxinneriter=0, log=None, comm=None, n_total=None):
    """scripts/algorithms/MultiLogReg.dml:150-340 with icpt = 0 (BASELINE config C4: X 1M x 5000, K = 20):
    trust-region Newton with a conjugate-gradient inner solver.  X: GPUObject N x D; Y: GPUObject N x K, the one-hot
    labels of the K non-baseline classes; y_last: GPUObject N x 1, the indicator of the baseline class (the script's
    Y[, K+1]).  The script carries N x (K+1) matrices whose last column of LT is identically zero; here that column
    is implicit: rowMaxs(LT) = max(rowMaxs(X B), 0), rowSums(exp_LT) = rowSums(exp(X B - mx)) + exp(-mx), which are
    the same operations on the same values.  The O(N D K) work per inner iteration is X %*% V and t(X) %*% (N x K).
    comm: X, Y, y_last are this rank's row shard of an n_total-row problem: X %*% V and the row-wise chains need no
    exchange; t(X) %*% (N x K) and every sum over N rows are a local result + one exchange; the D x K matrices are replicated.
    Returns (B ndarray D x K, outer iterations, objective)."""
    N, Dm, K = X.rows, X.cols, Y.cols
    Ntot = int(n_total) if n_total is not None else N
    E = L.Expr
    alloc = lambda r_, c_: GPUObject.allocate_dense(g, r_, c_)
    eta0, eta1, eta2, sigma1, sigma2, sigma3, psi = 1e-4, 0.25, 0.75, 0.25, 0.5, 4.0, 0.1

    fs = alloc(1, 1)

    def fsum(inputs, expr, consts=()):
        if comm is not None and inputs[0].rows == N and N != Dm:   # a sum over the row-sharded operands: local sum + one exchange
            L.fusedCellwise(g, "spoof", inputs, expr, consts, "sum", out=fs, read_back=False)
            _allreduce(comm, fs, 1)
            return _scalar(g, fs)
        return L.fusedCellwise(g, "spoof", inputs, expr, consts, "sum", out=fs)

    def dotm(a, b):
        return fsum([a, b], E.input(0) * E.input(1))

    def xt_times(M_, out):     # out (D x K) = t(X) %*% M_ (N x K)
        g.call("sysml_b200_matmult_dense", _p(X), _p(M_), _p(out), N, Dm, N, K, 1, 0)
        _allreduce(comm, out, Dm * K)

    def x_times(V_, out):      # out (N x K) = X %*% V_ (D x K)
        g.call("sysml_b200_matmult_dense", _p(X), _p(V_), _p(out), N, Dm, Dm, K, 0, 0)

    def axpy(a, b, alpha, out):  # out = a + alpha*b
        g.call("sysml_b200_axpy", _p(a), _p(b), float(alpha), _p(out), a.rows, a.cols, 0)

    rsx = alloc(N, 1)
    g.call("sysml_b200_agg_unary_dense", _p(X), N, Dm, L.AGG_SUMSQ, L.DIR_ROW, _p(rsx))
    if comm is not None:
        mxd = alloc(1, 1)
        g.call("sysml_b200_agg_unary_dense", _p(rsx), N, 1, L.AGG_MAX, L.DIR_ALL, _p(mxd))
        comm.allreduce(mxd.ptr, 1, "max")
        mxs = _scalar(g, mxd)
        mxd.free()
    else:
        mxs = L.unaryAggregate(g, "uamax", rsx, "uamax")
    rsx.free()
    delta = 0.5 * np.sqrt(Dm) / np.sqrt(mxs)
    lam = float(reg)
    B, Bn, S, Sn, R, V, HV, Grad = (alloc(Dm, K) for _ in range(8))
    P, Pn, Q, LT, T1 = (alloc(N, K) for _ in range(5))
    mx, rs, rs2b = alloc(N, 1), alloc(N, 1), alloc(N, 1)
    g.call("sysml_b200_fill", _p(B), Dm * K, 0.0)
    g.call("sysml_b200_fill", _p(P), N * K, 1.0 / (K + 1))
    obj = Ntot * np.log(K + 1)

    def gradient(Pm, Bm):
        g.call("sysml_b200_binary_mm", _p(Pm), _p(Y), _p(T1), N, K, 0, L.BINARY_OPS["-"])    # P - Y
        xt_times(T1, Grad)
        axpy(Grad, Bm, lam, Grad)                                                              # + lambda * B
        return np.sqrt(dotm(Grad, Grad))

    norm_Grad = gradient(P, B)
    norm_Grad_initial = norm_Grad
    if maxinneriter == 0:
        maxinneriter = Dm * K
    it = 1
    converge = (norm_Grad < tol) or (it > maxiter)
    while not converge:
        g.call("sysml_b200_fill", _p(S), Dm * K, 0.0)
        g.call("sysml_b200_binary_ms", _p(Grad), -1.0, _p(R), Dm * K, L.BINARY_OPS["*"], 0)    # R = -Grad
        g.call("sysml_b200_d2d", _p(V), _p(R), 8 * Dm * K)
        delta2 = delta ** 2
        inneriter = 1
        norm_R2 = dotm(R, R)
        innerconverge = np.sqrt(norm_R2) <= psi * norm_Grad
        reached = 0
        while not innerconverge:
            x_times(V, Q)
            # Q = P * (X V);  T1 = Q - P * rowSums(Q)
            g.call("sysml_b200_binary_mm", _p(P), _p(Q), _p(Q), N, K, 0, L.BINARY_OPS["*"])
            g.call("sysml_b200_agg_unary_dense", _p(Q), N, K, L.AGG_SUM, L.DIR_ROW, _p(rs))
            L.fusedCellwise(g, "spoof", [Q, P, rs], E.input(0) - E.input(1) * E.input(2), out=T1)
            xt_times(T1, HV)
            axpy(HV, V, lam, HV)
            alpha = norm_R2 / dotm(V, HV)
            axpy(S, V, alpha, Sn)
            norm_Snew2 = dotm(Sn, Sn)
            if norm_Snew2 <= delta2:
                S, Sn = Sn, S
                axpy(R, HV, -alpha, R)
                old = norm_R2
                norm_R2 = dotm(R, R)
                axpy(R, V, norm_R2 / old, V)
                innerconverge = np.sqrt(norm_R2) <= psi * norm_Grad
            else:
                reached = 1
                sv, v2, s2 = dotm(S, V), dotm(V, V), dotm(S, S)
                rad = np.sqrt(sv ** 2 + v2 * (delta2 - s2))
                alpha = (delta2 - s2) / (sv + rad) if sv >= 0 else (rad - sv) / v2
                axpy(S, V, alpha, S)
                axpy(R, HV, -alpha, R)
                innerconverge = True
            inneriter += 1
            innerconverge = innerconverge or (inneriter > maxinneriter)
        gs = dotm(S, Grad)
        qk = -0.5 * (gs - dotm(S, R))
        axpy(B, S, 1.0, Bn)
        x_times(Bn, LT)
        g.call("sysml_b200_agg_unary_dense", _p(LT), N, K, L.AGG_MAX, L.DIR_ROW, _p(mx))
        g.call("sysml_b200_binary_ms", _p(mx), 0.0, _p(mx), N, L.BINARY_OPS["max"], 0)           # the implicit zero column
        g.call("sysml_b200_binary_mm", _p(LT), _p(mx), _p(LT), N, K, L.BCAST_COL, L.BINARY_OPS["-"])
        g.call("sysml_b200_unary", _p(LT), _p(Pn), N * K, L.UNARY_OPS["exp"])                    # exp_LT
        g.call("sysml_b200_agg_unary_dense", _p(Pn), N, K, L.AGG_SUM, L.DIR_ROW, _p(rs))
        # rs += exp(0 - mx)
        rs2 = L.fusedCellwise(g, "spoof", [rs, mx], E.input(0) + (E.const(0) - E.input(1)).unary("exp"), (0.0,), out=rs2b)
        g.call("sysml_b200_binary_mm", _p(Pn), _p(rs2), _p(Pn), N, K, L.BCAST_COL, L.BINARY_OPS["/"])
        sum_y_lt = dotm(Y, LT) + fsum([y_last, mx], E.input(0) * (E.const(0) - E.input(1)), (0.0,))
        obj_new = -sum_y_lt + fsum([rs2], E.input(0).unary("log")) + 0.5 * lam * dotm(Bn, Bn)
        actred = obj - obj_new
        rho = actred / qk
        accepted = rho > eta0
        snorm = np.sqrt(dotm(S, S))
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
        if accepted:
            B, Bn = Bn, B
            P, Pn = Pn, P
            norm_Grad = gradient(P, B)
            obj = obj_new
        if log is not None:
            log.append((it, inneriter - 1, obj, norm_Grad))
        it += 1
        converge = (norm_Grad < tol * norm_Grad_initial) or (it > maxiter) or \
                   (reached == 0 and abs(actred) < (abs(obj) + abs(obj_new)) * 1e-14)
    out = B.to_host()
    for o in (B, Bn, S, Sn, R, V, HV, Grad, P, Pn, Q, LT, T1, mx, rs, rs2b, fs):
        o.free()
    return out, it - 1, obj
