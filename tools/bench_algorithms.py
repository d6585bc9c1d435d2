#!/usr/bin/env python
"""End-to-end workload timings on one GPU: LinearRegCG (scripts/perftest/runLinearRegCG.sh arguments:
icpt=0 reg=0.01 tol=1e-4) on C1 (100k x 1000) and C2-sized (10M x 1000) synthetic data, and the GLM
Poisson/log Newton-CG on 10M x 1000.  Wall-clock with a final device sync; prints JSON lines."""
import ctypes as C
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from systemml_b200 import GPUContext, LibMatrixB200 as L  # noqa: E402
from systemml_b200.algorithms import glm_poisson_log, linreg_cg  # noqa: E402

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
g.destroy()
