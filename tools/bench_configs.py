"""BASELINE configs C3, C4, C5 on one GPU at their ACTUAL sizes: per-op CUDA-event times with the algorithmic bytes / flops
of SURVEY section 8(d) and the roofline fraction, plus one end-to-end run of the config's algorithm and its CPU twin on a
stated row sample.  bench.py calls run_configs() at N = 1 and puts the result under "configs" in its JSON line;
`python tools/bench_configs.py [C3,C4,C5] [scale]` prints it on its own.

  C3  L2SVM, CSR 50M x 10k, ~Poisson(10) entries per row (nnz ~ 5e8): SpMV, t(X)%*%y, mmchain, rowSums; l2-svm.dml
  C4  MultiLogReg, dense 1M x 5000, 20 classes: X%*%V, t(X)%*%P, tsmm, rowSums / rowIndexMax on 1M x 21; MultiLogReg.dml
  C5  Kmeans, dense 10M x 100, k = 50: X%*%t(C), t(P)%*%X, rowMins / rowIndexMax / colSums on 10M x 50; Kmeans.dml
"""
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _p(o):
    return C.c_void_p(o.ptr)


def _timeit(g, fn, reps):
    fn()
    g.sync()
    tot, best = 0.0, 1e30
    for _ in range(reps):
        g.timer_start()
        fn()
        ms = g.timer_stop_ms()
        tot += ms
        best = min(best, ms)
    return tot / reps, best


def _op(name, ms, best, hbm, dmma, nbytes=None, flops=None):
    r = {"op": name, "ms": round(ms, 4), "ms_best": round(best, 4)}
    if nbytes is not None:
        r["algorithmic_bytes"] = nbytes
        r["GBs"] = round(nbytes / ms / 1e6, 1)
        r["frac_hbm"] = round(nbytes / ms / 1e6 / hbm, 3)
    if flops is not None:
        r["flops"] = flops
        r["TFLOPs"] = round(flops / ms / 1e9, 2)
        r["frac_dmma"] = round(flops / ms / 1e9 / dmma, 3)
    return r


def config_c3(g, hbm, dmma, scale=1.0, cpu=True, reps=5):
    from systemml_b200 import LibMatrixB200 as L
    from systemml_b200.algorithms import l2svm
    n, m, mean = int(50_000_000 * scale), 10_000, 10
    X = g.rand_csr(n, m, mean, seed=2)
    nnz = X.csr.nnz
    csr = (C.c_void_p(X.csr.rowPtr), C.c_void_p(X.csr.colInd), C.c_void_p(X.csr.val))
    v, y, w = g.rand(m, 1, seed=2), g.rand(n, 1, seed=3), g.rand(n, 1, seed=4, lo=0.0, hi=1.0)
    o_r, o_c = g.empty(n, 1), g.empty(m, 1)
    spb = 12.0 * nnz + 4.0 * (n + 1) + 8.0 * m + 8.0 * n
    ops = []
    a, b = _timeit(g, lambda: g.call("sysml_b200_matmult_csr_dense", *csr, _p(v), _p(o_r), n, m, m, 1, 0), reps)
    ops.append(_op("SpMV X%*%v", a, b, hbm, dmma, nbytes=spb, flops=2.0 * nnz))
    a, b = _timeit(g, lambda: g.call("sysml_b200_matmult_csr_dense", *csr, _p(y), _p(o_c), n, m, n, 1, 1), reps)
    ops.append(_op("t(X)%*%y", a, b, hbm, dmma, nbytes=spb, flops=2.0 * nnz))
    a, b = _timeit(g, lambda: g.call("sysml_b200_mmchain_csr", *csr, _p(v), _p(w), _p(o_c), n, m, 1), reps)
    ops.append(_op("mmchain XtwXv", a, b, hbm, dmma, nbytes=spb + 8.0 * n + 8.0 * m, flops=4.0 * nnz))
    a, b = _timeit(g, lambda: g.call("sysml_b200_agg_unary_csr", *csr, n, m, 0, 1, _p(o_r)), reps)
    ops.append(_op("rowSums", a, b, hbm, dmma, nbytes=12.0 * nnz + 4.0 * (n + 1) + 8.0 * n))
    # a sparse-safe cellwise operator that keeps X in CSR (csrc/sparse_cellwise.cu): X * y with y a column vector, both passes
    # (symbolic: sizes; numeric: values) into preallocated arrays
    c_rp, c_ci, c_va = g.allocate((n + 1) * 4), g.allocate(max(nnz, 1) * 4), g.allocate(max(nnz, 1) * 8)
    c_nnz = C.c_int64()

    def x_times_colvec():
        for fill in (False, True):
            g.call("sysml_b200_csr_dense_op", *csr, n, m, _p(y), 1, L.BINARY_OPS["*"], 0, C.c_void_p(c_rp), C.c_void_p(c_ci) if fill else None,
                   C.c_void_p(c_va) if fill else None, C.byref(c_nnz))
    a, b = _timeit(g, x_times_colvec, reps)
    ops.append(_op("X * y (column vector), CSR result, symbolic + numeric pass", a, b, hbm, dmma, nbytes=2 * (12.0 * nnz + 4.0 * n + 8.0 * n) + 12.0 * nnz + 8.0 * n))
    for q in (c_rp, c_ci, c_va):
        g.cudaFreeHelper(q)
    # the same two products through the column-ordered copy (csrc/psell.cu): what the iterative algorithms use from the
    # second iteration on
    from systemml_b200.gpu_context import PSELLPointer
    g.sync()
    t0 = time.perf_counter()
    pc = PSELLPointer(g, X.csr, n, m)
    g.sync()
    build_s = time.perf_counter() - t0
    a, b = _timeit(g, lambda: pc.vm(y.ptr, o_c.ptr), reps)
    ops.append(_op("t(X)%*%y [column-ordered copy]", a, b, hbm, dmma, nbytes=12.0 * nnz + 8.0 * n + 8.0 * m, flops=2.0 * nnz))
    a, b = _timeit(g, lambda: pc.mmchain(X.csr, v.ptr, w.ptr, o_r.ptr, o_c.ptr, 1), reps)
    ops.append(_op("mmchain XtwXv [CSR + column-ordered copy]", a, b, hbm, dmma, nbytes=spb + 8.0 * n + 8.0 * m, flops=4.0 * nnz))
    ops[-1]["note"] = "algorithmic bytes as for the one-pass form; this form reads X twice (24 bytes per non-zero) and writes/reads an n-vector"
    pc.deallocate()
    # labels from a planted separator, then l2-svm.dml (icpt=0 reg=1 tol=0.001 maxiter=20, scripts/perftest/runL2SVM.sh style)
    g.call("sysml_b200_matmult_csr_dense", *csr, _p(v), _p(o_r), n, m, m, 1, 0)
    Y = g.empty(n, 1)
    g.call("sysml_b200_binary_ms", _p(o_r), 0.0, _p(Y), n, L.BINARY_OPS[">"], 0)
    g.call("sysml_b200_binary_ms", _p(Y), 2.0, _p(Y), n, L.BINARY_OPS["*"], 0)
    g.call("sysml_b200_binary_ms", _p(Y), 1.0, _p(Y), n, L.BINARY_OPS["-"], 0)
    for o in (y, w, o_r, o_c):
        o.free()
    l2svm(g, X, Y, reg=1.0, tol=0.001, maxiter=1)      # warm-up: compiles the fused chains
    g.sync()
    t0 = time.perf_counter()
    wgt, iters, obj = l2svm(g, X, Y, reg=1.0, tol=0.001, maxiter=20)
    g.sync()
    dt = time.perf_counter() - t0
    l2svm(g, X, Y, reg=1.0, tol=0.001, maxiter=1, column_copy=True)   # builds the column-ordered copy
    g.sync()
    t0 = time.perf_counter()
    _, iters2, obj2 = l2svm(g, X, Y, reg=1.0, tol=0.001, maxiter=20, column_copy=True)
    g.sync()
    dt2 = time.perf_counter() - t0
    res = {"shape": [n, m], "nnz": nnz, "layout": "CSR int32, ~Poisson(10) entries per row (sysml_b200_rand_csr seed 2)", "ops": ops,
           "column_ordered_copy_build_seconds": round(build_s, 3),
           "algorithm": {"name": "L2SVM (l2-svm.dml, icpt=0 reg=1 tol=0.001 maxiter=20)", "outer_iterations": iters, "seconds": round(dt, 4),
                         "objective": obj, "plan": "t(X)%*%y as the scatter kernel",
                         "seconds_with_column_ordered_copy": round(dt2, 4), "objective_with_column_ordered_copy": obj2,
                         "outer_iterations_with_column_ordered_copy": iters2}}
    if cpu:
        import scipy.sparse as sp
        from oracle import oracle as O
        from oracle.algorithms_cpu import cpu_l2svm
        ns = min(n, 1_000_000)
        rp, ci, va = O.rand_csr(ns, m, mean, 2)
        Xs = sp.csr_matrix((va, ci, rp), shape=(ns, m))
        Ys = np.where(Xs @ O.rand_uniform(m, seed=2) > 0, 1.0, -1.0)
        t0 = time.perf_counter()
        _, it_c, _ = cpu_l2svm(Xs, Ys, 1.0, 0.001, 20)
        dc = time.perf_counter() - t0
        res["algorithm"]["cpu_twin"] = {"kind": "numpy + scipy.sparse twin of the same script (oracle/algorithms_cpu.py)", "sample_rows": ns,
                                        "outer_iterations": it_c, "seconds_on_sample": round(dc, 3),
                                        "seconds_scaled_to_full_rows": round(dc * n / ns, 2)}
    X.free()
    Y.free()
    v.free()
    return res


def config_c4(g, hbm, dmma, scale=1.0, cpu=True, reps=3):
    from systemml_b200 import LibMatrixB200 as L
    from systemml_b200.algorithms import multilogreg
    from systemml_b200.lib_matrix import Expr as E
    N, D, K = int(1_000_000 * scale), 5000, 20
    X = g.rand(N, D, seed=3)
    V, P = g.rand(D, K, seed=4), g.rand(N, K, seed=5)
    oNK, oDK = g.empty(N, K), g.empty(D, K)
    ops = []
    a, b = _timeit(g, lambda: g.call("sysml_b200_matmult_dense", _p(X), _p(V), _p(oNK), N, D, D, K, 0, 0), reps)
    ops.append(_op("X%*%V (5000x20)", a, b, hbm, dmma, nbytes=8.0 * (N * D + D * K + N * K), flops=2.0 * N * D * K))
    a, b = _timeit(g, lambda: g.call("sysml_b200_matmult_dense", _p(X), _p(P), _p(oDK), N, D, N, K, 1, 0), reps)
    ops.append(_op("t(X)%*%P (1Mx20)", a, b, hbm, dmma, nbytes=8.0 * (N * D + N * K + D * K), flops=2.0 * N * D * K))
    T = g.empty(D, D)
    a, b = _timeit(g, lambda: g.call("sysml_b200_tsmm_dense", _p(X), _p(T), N, D, 1, None), 1)
    ops.append(_op("tsmm t(X)%*%X", a, b, hbm, dmma, flops=float(N) * D * (D + 1)))
    T.free()
    Q = g.rand(N, K + 1, seed=6)
    oN = g.empty(N, 1)
    for name, op in (("rowSums (1M x 21)", 0), ("rowIndexMax (1M x 21)", 4)):
        a, b = _timeit(g, lambda: g.call("sysml_b200_agg_unary_dense", _p(Q), N, K + 1, op, 1, _p(oN)), reps + 2)
        ops.append(_op(name, a, b, hbm, dmma, nbytes=8.0 * N * (K + 1) + 8.0 * N))
    for o in (P, oNK, oDK, Q, oN):
        o.free()
    # labels: argmax of X %*% W over K+1 classes; MultiLogReg.dml icpt=0 reg=1 tol=1e-6 moi=5
    W = g.rand(D, K + 1, seed=7)
    sc = g.empty(N, K + 1)
    g.call("sysml_b200_matmult_dense", _p(X), _p(W), _p(sc), N, D, D, K + 1, 0, 0)
    lab = L.unaryAggregate(g, "uarimax", sc, "uarimax")
    ids = g.dense(np.arange(1, K + 1, dtype=np.float64).reshape(1, K))
    Yk = L.fusedCellwise(g, "spoof", [lab, ids], E.input(0).binary(E.input(1), "=="), rows=N, cols=K)
    ylast = L.matrixScalarArithmetic(g, "==", lab, float(K + 1), "==")
    for o in (W, sc, V):
        o.free()
    multilogreg(g, X, Yk, ylast, reg=1.0, tol=1e-6, maxiter=1)
    g.sync()
    t0 = time.perf_counter()
    log = []
    B, outer, obj = multilogreg(g, X, Yk, ylast, reg=1.0, tol=1e-6, maxiter=5, log=log)
    g.sync()
    dt = time.perf_counter() - t0
    res = {"shape": [N, D], "classes": K + 1, "ops": ops,
           "algorithm": {"name": "MultiLogReg (MultiLogReg.dml, icpt=0 reg=1 tol=1e-6 moi=5)", "outer_iterations": outer,
                         "inner_cg_iterations": [l[1] for l in log], "seconds": round(dt, 4), "objective": obj}}
    for o in (X, Yk, ylast, lab, ids):
        o.free()
    if cpu:
        from oracle import oracle as O
        from oracle.algorithms_cpu import cpu_multilogreg
        ns = min(N, 20_000)
        Xh = O.rand_uniform(ns * D, seed=3).reshape(ns, D)
        Wh = O.rand_uniform(D * (K + 1), seed=7).reshape(D, K + 1)
        labh = np.argmax(Xh @ Wh, axis=1)
        Yh = (labh[:, None] == np.arange(K)[None, :]).astype(np.float64)
        yl = (labh == K).astype(np.float64).reshape(ns, 1)
        t0 = time.perf_counter()
        _, it_c, _ = cpu_multilogreg(Xh, Yh, yl, 1.0, 1e-6, 5, 0)
        dc = time.perf_counter() - t0
        res["algorithm"]["cpu_twin"] = {"kind": "numpy twin of the same script (oracle/algorithms_cpu.py; BLAS matmults)", "sample_rows": ns,
                                        "outer_iterations": it_c, "seconds_on_sample": round(dc, 3),
                                        "seconds_scaled_to_full_rows": round(dc * N / ns, 2)}
    return res


def config_c5(g, hbm, dmma, scale=1.0, cpu=True, reps=5):
    from systemml_b200 import LibMatrixB200 as L
    from systemml_b200.algorithms import kmeans_lloyd
    from systemml_b200.lib_matrix import Expr as E
    n, m, k = int(10_000_000 * scale), 100, 50
    # 50 Gaussian-ish blobs: one-hot(assignment) %*% centres + U(-1,1) noise
    U = g.rand(n, k, seed=5)
    asg = L.unaryAggregate(g, "uarimax", U, "uarimax")
    ids = g.dense(np.arange(1, k + 1, dtype=np.float64).reshape(1, k))
    H = L.fusedCellwise(g, "spoof", [asg, ids], E.input(0).binary(E.input(1), "=="), rows=n, cols=k)
    U.free()
    centers = g.rand(k, m, seed=6, lo=-10.0, hi=10.0)
    X = g.empty(n, m)
    g.call("sysml_b200_matmult_dense", _p(H), _p(centers), _p(X), n, k, k, m, 0, 0)
    noise = g.rand(n, m, seed=7)
    g.call("sysml_b200_axpy", _p(X), _p(noise), 1.0, _p(X), n, m, 0)
    for o in (noise, asg, ids):
        o.free()
    C0 = g.rand(k, m, seed=8)                                # initial centroids: the blob centres, each displaced by U(-1,1)
    g.call("sysml_b200_binary_mm", _p(C0), _p(centers), _p(C0), k, m, 0, L.BINARY_OPS["+"])
    D, oKM = g.empty(n, k), g.empty(k, m)
    ops = []
    a, b = _timeit(g, lambda: g.call("sysml_b200_matmult_dense", _p(X), _p(C0), _p(D), n, m, k, m, 0, 1), reps)
    ops.append(_op("X%*%t(C) (100x50)", a, b, hbm, dmma, nbytes=8.0 * (n * m + k * m + n * k), flops=2.0 * n * m * k))
    a, b = _timeit(g, lambda: g.call("sysml_b200_matmult_dense", _p(H), _p(X), _p(oKM), n, k, n, m, 1, 0), reps)
    ops.append(_op("t(P)%*%X (50 x 10M x 100)", a, b, hbm, dmma, nbytes=8.0 * (n * k + n * m + k * m), flops=2.0 * n * m * k))
    oN, oK = g.empty(n, 1), g.empty(1, k)
    for name, op, d, o in (("rowMins (10M x 50)", 3, 1, oN), ("rowIndexMax (10M x 50)", 4, 1, oN), ("colSums (10M x 50)", 0, 2, oK), ("rowSums (10M x 50)", 0, 1, oN)):
        a, b = _timeit(g, lambda: g.call("sysml_b200_agg_unary_dense", _p(D), n, k, op, d, _p(o)), reps)
        ops.append(_op(name, a, b, hbm, dmma, nbytes=8.0 * n * k + 8.0 * (n if d == 1 else k)))
    for o in (H, D, oKM, oN, oK):
        o.free()
    # the whole Lloyd step as one kernel (csrc/kmeans.cu)
    c2d, sums, cnt, smin = g.empty(k, 1), g.empty(k, m), g.empty(k, 1), g.empty(1, 1)
    g.call("sysml_b200_agg_unary_dense", _p(C0), k, m, 1, 1, _p(c2d))
    a, b = _timeit(g, lambda: g.call("sysml_b200_kmeans_step", _p(X), _p(C0), _p(c2d), n, m, k, _p(sums), _p(cnt), _p(smin)), reps)
    ops.append(_op("one-pass Lloyd step (distances + rowMins + tie rule + t(P)%*%X + colSums + sum)", a, b, hbm, dmma, nbytes=8.0 * n * m, flops=2.0 * n * m * k))
    ops[-1]["note"] = "flops: the distance product only (2nmk); the centroid sums are n*m additions, not a second product"
    for o in (c2d, sums, cnt, smin):
        o.free()
    kmeans_lloyd(g, X, C0, max_iter=1)
    g.sync()
    t0 = time.perf_counter()
    # eps = 0: the run does not stop at convergence, so both plans execute exactly max_iter = 10 Lloyd iterations
    Cf, iters, wcss, term = kmeans_lloyd(g, X, C0, max_iter=10, eps=0.0)
    g.sync()
    dt = time.perf_counter() - t0
    t0 = time.perf_counter()
    _, iters_c, _, _ = kmeans_lloyd(g, X, C0, max_iter=10, eps=0.0, fused="chains")
    g.sync()
    dt_chains = time.perf_counter() - t0
    res = {"shape": [n, m], "k": k, "ops": ops,
           "algorithm": {"name": "Kmeans (Kmeans.dml Lloyd iterations of one run, eps=0 max_iter=10: exactly 10 iterations)", "iterations": iters, "termination_code": term,
                         "seconds": round(dt, 4), "ms_per_iteration": round(1000.0 * dt / max(iters, 1), 3), "wcss": wcss,
                         "plan": "one-pass step kernel", "ms_per_iteration_as_matmults_and_fused_chains": round(1000.0 * dt_chains / max(iters_c, 1), 3)}}
    if cpu:
        from oracle.algorithms_cpu import cpu_kmeans_lloyd
        ns = min(n, 200_000)
        Xh = g.d2h(X.ptr, (ns, m))
        C0h = C0.to_host()
        t0 = time.perf_counter()
        _, it_c, _, _ = cpu_kmeans_lloyd(Xh, C0h, 10, 0.0)
        dc = time.perf_counter() - t0
        res["algorithm"]["cpu_twin"] = {"kind": "numpy twin of the same script (oracle/algorithms_cpu.py; BLAS matmults)", "sample_rows": ns,
                                        "iterations": it_c, "seconds_on_sample": round(dc, 3), "ms_per_iteration_scaled_to_full_rows": round(1000.0 * dc / max(it_c, 1) * n / ns, 1)}
    for o in (X, C0, centers):
        o.free()
    return res


def run_configs(g, hbm_peak, dmma_peak, which=("C3", "C4", "C5"), scale=1.0, cpu=True):
    out = {"peaks": {"hbm_gbs": hbm_peak, "fp64_dmma_tflops": dmma_peak,
                     "note": "frac_hbm against MEASURED_PEAKS.json hbm_gbs; frac_dmma against profiles/fp64_peak.json (this repo's own DMMA probe, not measured in this run); ops near the ridge report both"},
           "timing": "CUDA events on the context's stream, mean of the stated reps after one warm-up call; inputs far larger than the 126 MB L2"}
    fns = {"C3": config_c3, "C4": config_c4, "C5": config_c5}
    for c in which:
        t0 = time.perf_counter()
        try:
            out[c] = fns[c](g, hbm_peak, dmma_peak, scale=scale, cpu=cpu)
        except Exception as e:  # noqa: BLE001 -- a config that does not fit must not take the headline line down with it
            out[c] = {"error": "%s: %s" % (type(e).__name__, e)}
        out[c]["wall_s"] = round(time.perf_counter() - t0, 1)
    return out


if __name__ == "__main__":
    from systemml_b200 import GPUContext
    which = tuple(sys.argv[1].split(",")) if len(sys.argv) > 1 else ("C3", "C4", "C5")
    scale = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
    hbm, dmma = 6650.0, 37.0
    try:
        hbm = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
        dmma = float(json.load(open(os.path.join(ROOT, "profiles", "fp64_peak.json")))["dmma_peak_tflops"])
    except Exception:
        pass
    g = GPUContext(0)
    print(json.dumps(run_configs(g, hbm, dmma, which, scale)))
    g.destroy()
