#!/usr/bin/env python
"""End-to-end workload timings on one GPU: LinearRegCG (scripts/perftest/runLinearRegCG.sh arguments:
icpt=0 reg=0.01 tol=1e-4) on C1 (100k x 1000) and C2-sized (10M x 1000) synthetic data, and the GLM
Poisson/log Newton-CG on 10M x 1000, and L2SVM (scripts/perftest/runL2SVM.sh style: icpt=0 reg=1 tol=0.001) on C3
(CSR 50M x 10k, ~Poisson(10) non-zeros per row) with the n-vector chains fused and operator by operator.
Wall-clock with a final device sync; prints JSON lines.  ROWS / SVM_ROWS shrink the inputs."""
import ctypes as C
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from systemml_b200 import GPUContext, LibMatrixB200 as L  # noqa: E402
from systemml_b200.algorithms import glm_poisson_log, l2svm, linreg_cg  # noqa: E402

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
g.destroy()
