#!/usr/bin/env python
"""End-to-end workload timings on one GPU: LinearRegCG (scripts/perftest/runLinearRegCG.sh arguments:
icpt=0 reg=0.01 tol=1e-4) on C1 (100k x 1000) and C2-sized (10M x 1000) synthetic data, and the GLM
Poisson/log Newton-CG on 10M x 1000, and L2SVM (scripts/perftest/runL2SVM.sh style: icpt=0 reg=1 tol=0.001) on C3
(CSR 50M x 10k, ~Poisson(10) non-zeros per row) with the n-vector chains fused and operator by operator, MultiLogReg on
C4 (1M x 5000, 21 classes) and the Lloyd iterations of Kmeans on C5 (10M x 100, k = 50).
Wall-clock with a final device sync; prints JSON lines.  ROWS / SVM_ROWS shrink the inputs."""
import ctypes as C
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from systemml_b200 import GPUContext, LibMatrixB200 as L  # noqa: E402
from systemml_b200.algorithms import glm_poisson_log, kmeans_lloyd, l2svm, linreg_cg, multilogreg  # noqa: E402

g = GPUContext(0)
p_ = lambda o: C.c_void_p(o.ptr)


def make_regression(n, m):
    X = g.rand(n, m, seed=0)
    wtrue = g.rand(m, 1, seed=1)
    y = g.empty(n, 1)
    g.call("sysml_b200_matmult_dense", p_(X), p_(wtrue), p_(y), n, m, m, 1, 0, 0)       # y = X w
    noise = g.rand(n, 1, seed=2)
    g.call("sysml_b200_axpy", p_(y), p_(noise), 0.1, p_(y), n, 1, 0)                     # + 0.1 * U(-1,1)
    noise.free()
    return X, y, wtrue


for n, m, tag in ((100_000, 1000, "C1"), (int(os.environ.get("ROWS", 10_000_000)), 1000, "C2")):
    X, y, wtrue = make_regression(n, m)
    linreg_cg(g, X, y, reg=0.01, tol=1e-4, maxi=2)  # warm-up (workspace, module load)
    g.sync()
    t0 = time.perf_counter()
    log = []
    beta, iters = linreg_cg(g, X, y, reg=0.01, tol=1e-4, log=log)
    g.sync()
    dt = time.perf_counter() - t0
    err = float(abs(beta - wtrue.to_host().reshape(-1)).max())
    print(json.dumps({"workload": "LinearRegCG icpt=0 reg=0.01 tol=1e-4", "config": tag, "shape": [n, m], "iterations": iters,
                      "seconds": round(dt, 4), "iters_per_sec": round(iters / dt, 2), "final_residual_ratio": log[-1][1],
                      "max_abs_err_vs_true_w": err}), flush=True)
    if tag == "C2":
        # GLM Poisson/log on the same X: y ~ Poisson(exp(X b)), b ~ U(-0.05, 0.05) (approximated by its mean on device)
        b = g.rand(m, 1, seed=5, lo=-0.05, hi=0.05)
        eta = g.empty(n, 1)
        g.call("sysml_b200_matmult_dense", p_(X), p_(b), p_(eta), n, m, m, 1, 0, 0)
        g.call("sysml_b200_unary", p_(eta), p_(y), n, L.UNARY_OPS["exp"])  # y = E[Poisson] = exp(X b): noiseless counts
        g.sync()
        t0 = time.perf_counter()
        beta, inner = glm_poisson_log(g, X, y, reg=0.0, outer_iters=5, inner_iters=20, inner_tol=1e-6)
        g.sync()
        dt = time.perf_counter() - t0
        print(json.dumps({"workload": "GLM Poisson/log Newton-CG", "config": tag, "shape": [n, m], "outer_iterations": 5,
                          "inner_cg_iterations": inner, "seconds": round(dt, 4), "mmchain_calls_per_sec": round(sum(inner) / dt, 2),
                          "max_abs_err_vs_true_b": float(abs(beta - b.to_host().reshape(-1)).max())}), flush=True)
    X.free()
    y.free()

# ---- C3: L2SVM on a CSR matrix ----
import numpy as np  # noqa: E402
from csr_gen import poisson_rows_csr  # noqa: E402

n, m = int(os.environ.get("SVM_ROWS", 50_000_000)), 10_000
t0 = time.perf_counter()
rp, ci, va = poisson_rows_csr(n, m, 10, 2)
Xs = g.csr(rp, ci, va, (n, m))
nnz = int(rp[-1])
del rp, ci, va
wtrue = g.rand(m, 1, seed=1)
score = g.empty(n, 1)
g.call("sysml_b200_matmult_csr_dense", C.c_void_p(Xs.csr.rowPtr), C.c_void_p(Xs.csr.colInd), C.c_void_p(Xs.csr.val), p_(wtrue), p_(score), n, m, m, 1, 0)
Y = g.empty(n, 1)
g.call("sysml_b200_binary_ms", p_(score), 0.0, p_(Y), n, L.BINARY_OPS[">"], 0)      # labels: 2*(score > 0) - 1
g.call("sysml_b200_binary_ms", p_(Y), 2.0, p_(Y), n, L.BINARY_OPS["*"], 0)
g.call("sysml_b200_binary_ms", p_(Y), 1.0, p_(Y), n, L.BINARY_OPS["-"], 0)
g.sync()
gen_s = time.perf_counter() - t0
l2svm(g, Xs, Y, reg=1.0, tol=0.001, maxiter=1)  # warm-up: compiles the fused chains
for fused in (True, False):
    g.sync()
    t0 = time.perf_counter()
    log = []
    w, iters, obj = l2svm(g, Xs, Y, reg=1.0, tol=0.001, maxiter=20, fused=fused, log=log)
    g.sync()
    dt = time.perf_counter() - t0
    acc = None
    print(json.dumps({"workload": "L2SVM icpt=0 reg=1 tol=0.001 maxiter=20", "config": "C3", "shape": [n, m], "nnz": nnz,
                      "n_vector_chains": "fused" if fused else "one launch per operator", "outer_iterations": iters,
                      "seconds": round(dt, 4), "outer_iters_per_sec": round(iters / dt, 2), "objective": obj,
                      "data_generation_seconds": round(gen_s, 1)}), flush=True)
Xs.free()
Y.free()

# ---- C4: MultiLogReg 1M x 5000, K = 20 non-baseline classes ----
from systemml_b200.lib_matrix import Expr as E  # noqa: E402

N, Dm, K = int(os.environ.get("MLR_ROWS", 1_000_000)), 5000, 20
X = g.rand(N, Dm, seed=3)
W = g.rand(Dm, K + 1, seed=4)
sc = g.empty(N, K + 1)
g.call("sysml_b200_matmult_dense", p_(X), p_(W), p_(sc), N, Dm, Dm, K + 1, 0, 0)
lab = L.unaryAggregate(g, "uarimax", sc, "uarimax")                       # labels 1..K+1
ids = g.dense(np.arange(1, K + 1, dtype=np.float64).reshape(1, K))
Yk = L.fusedCellwise(g, "spoof", [lab, ids], E.input(0).binary(E.input(1), "=="), rows=N, cols=K)   # one-hot, first K classes
ylast = L.matrixScalarArithmetic(g, "==", lab, float(K + 1), "==")
for o in (W, sc):
    o.free()
multilogreg(g, X, Yk, ylast, reg=1.0, tol=1e-6, maxiter=1)  # warm-up (fused kernels, workspace)
g.sync()
t0 = time.perf_counter()
log = []
B, outer, obj = multilogreg(g, X, Yk, ylast, reg=1.0, tol=1e-6, maxiter=5, log=log)
g.sync()
dt = time.perf_counter() - t0
inner = [l[1] for l in log]
print(json.dumps({"workload": "MultiLogReg icpt=0 reg=1 tol=1e-6 moi=5", "config": "C4", "shape": [N, Dm], "classes": K + 1,
                  "outer_iterations": outer, "inner_cg_iterations": inner, "seconds": round(dt, 4),
                  "inner_iters_per_sec": round(sum(inner) / dt, 2), "objective": obj}), flush=True)
for o in (X, Yk, ylast, lab, ids):
    o.free()

# ---- C5: Kmeans 10M x 100, k = 50 (blobs: one-hot(assignment) %*% centers + noise) ----
n, m, k = int(os.environ.get("KM_ROWS", 10_000_000)), 100, 50
U = g.rand(n, k, seed=5)
asg = L.unaryAggregate(g, "uarimax", U, "uarimax")
ids = g.dense(np.arange(1, k + 1, dtype=np.float64).reshape(1, k))
H = L.fusedCellwise(g, "spoof", [asg, ids], E.input(0).binary(E.input(1), "=="), rows=n, cols=k)
U.free()
centers = g.rand(k, m, seed=6, lo=-10.0, hi=10.0)
X = g.empty(n, m)
g.call("sysml_b200_matmult_dense", p_(H), p_(centers), p_(X), n, k, k, m, 0, 0)
noise = g.rand(n, m, seed=7)
g.call("sysml_b200_axpy", p_(X), p_(noise), 1.0, p_(X), n, m, 0)
for o in (H, noise, asg, ids):
    o.free()
C0 = g.rand(k, m, seed=8)                                # initial centroids: the blob centres, each displaced by U(-1,1)
g.call("sysml_b200_binary_mm", p_(C0), p_(centers), p_(C0), k, m, 0, L.BINARY_OPS["+"])
kmeans_lloyd(g, X, C0, max_iter=1)  # warm-up
for fused in (True, False):
    g.sync()
    t0 = time.perf_counter()
    Cf, iters, wcss, term = kmeans_lloyd(g, X, C0, max_iter=20, eps=1e-6, fused=fused)
    g.sync()
    dt = time.perf_counter() - t0
    print(json.dumps({"workload": "Kmeans (Lloyd iterations of one run) eps=1e-6 max_iter=20", "config": "C5", "shape": [n, m], "k": k,
                      "cellwise_chains": "fused" if fused else "one launch per operator", "iterations": iters,
                      "termination_code": term, "seconds": round(dt, 4), "iters_per_sec": round(iters / dt, 2), "wcss": wcss}), flush=True)
g.destroy()
