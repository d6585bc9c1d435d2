#!/usr/bin/env python
"""Runs ONE op a few times (for ncu captures): python tools/run_op.py tsmm 2000000 1000 | mmchain n m | matmult M K N ta | spmv rows cols per [mode] | psell rows cols per | csrcell rows cols per kind | agg n m op dir"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from systemml_b200 import GPUContext  # noqa: E402

op = sys.argv[1]
a = [int(x) for x in sys.argv[2:]]
g = GPUContext(0)
p_ = lambda o: C.c_void_p(o.ptr)
if op == "tsmm":
    n, m = a
    X, o = g.rand(n, m, seed=1), g.empty(m, m)
    fn = lambda: g.call("sysml_b200_tsmm_dense", p_(X), p_(o), n, m, 1, None)
elif op == "mmchain":
    n, m = a
    X, v, w, o = g.rand(n, m, seed=1), g.rand(m, 1, seed=2), g.rand(n, 1, seed=3, lo=0.0, hi=1.0), g.empty(m, 1)
    fn = lambda: g.call("sysml_b200_mmchain_dense", p_(X), p_(v), p_(w), p_(o), n, m, 1)
elif op == "mmchain_in_big":  # the same product with X placed inside one allocation of `big_gb` GB (does the size of the ALLOCATION matter?)
    n, m, big_gb, off_gb = a
    from systemml_b200.gpu_context import GPUObject
    base = g.allocate(big_gb << 30)
    X = GPUObject(g, n, m)
    X.ptr, X.owns = base + (off_gb << 30), False
    g.call("sysml_b200_rand_uniform", C.c_void_p(X.ptr), n * m, 1, 0, -1.0, 1.0)
    v, w, o = g.rand(m, 1, seed=2), g.rand(n, 1, seed=3, lo=0.0, hi=1.0), g.empty(m, 1)
    fn = lambda: g.call("sysml_b200_mmchain_dense", p_(X), p_(v), p_(w), p_(o), n, m, 1)
elif op == "matmult":
    M, K, N, ta = a
    if ta:
        A, B, o = g.rand(M, K, seed=1), g.rand(M, N, seed=2), g.empty(K, N)
        fn = lambda: g.call("sysml_b200_matmult_dense", p_(A), p_(B), p_(o), M, K, M, N, 1, 0)
    else:
        A, B, o = g.rand(M, K, seed=1), g.rand(K, N, seed=2), g.empty(M, N)
        fn = lambda: g.call("sysml_b200_matmult_dense", p_(A), p_(B), p_(o), M, K, K, N, 0, 0)
elif op == "spmv":
    import numpy as np
    rows, cols, per = a[:3]
    mode = a[3] if len(a) > 3 else 0  # 0 X%*%v, 1 t(X)%*%y, 2 mmchain XtwXv
    from csr_gen import poisson_rows_csr
    rp, ci, va = poisson_rows_csr(rows, cols, per, 2)
    S = g.csr(rp, ci, va, (rows, cols))
    v, o = g.rand(cols, 1, seed=2), g.empty(rows, 1)
    y, oc = g.rand(rows, 1, seed=3), g.empty(cols, 1)
    csr3 = (C.c_void_p(S.csr.rowPtr), C.c_void_p(S.csr.colInd), C.c_void_p(S.csr.val))
    if mode == 0:
        fn = lambda: g.call("sysml_b200_matmult_csr_dense", *csr3, p_(v), p_(o), rows, cols, cols, 1, 0)
    elif mode == 1:
        fn = lambda: g.call("sysml_b200_matmult_csr_dense", *csr3, p_(y), p_(oc), rows, cols, rows, 1, 1)
    else:
        fn = lambda: g.call("sysml_b200_mmchain_csr", *csr3, p_(v), p_(y), p_(oc), rows, cols, 1)
elif op == "psell":
    rows, cols, per = a[:3]
    from systemml_b200.gpu_context import PSELLPointer
    import time
    S = g.rand_csr(rows, cols, per, seed=2)
    g.sync()
    t0 = time.perf_counter()
    pc = PSELLPointer(g, S.csr, rows, cols)
    g.sync()
    print("psell build s", round(time.perf_counter() - t0, 3), "nnz", S.csr.nnz, "slots", pc.slots, "groups", pc.groups, "padding",
          round(pc.slots / max(S.csr.nnz, 1) - 1, 4))
    y, oc = g.rand(rows, 1, seed=3), g.empty(cols, 1)
    fn = lambda: pc.vm(y.ptr, oc.ptr)
elif op == "csrcell":  # sparse-safe cellwise with a CSR result: kind 0 X*y (column vector), 1 X*0.5, 2 X*X
    rows, cols, per, kind = a[:4]
    S = g.rand_csr(rows, cols, per, seed=2)
    y = g.rand(rows, 1, seed=3)
    csr3 = (C.c_void_p(S.csr.rowPtr), C.c_void_p(S.csr.colInd), C.c_void_p(S.csr.val))
    rp, ci, va = g.allocate((rows + 1) * 4), g.allocate(S.csr.nnz * 4), g.allocate(S.csr.nnz * 8)
    nnz = C.c_int64()
    outs = lambda fill: (C.c_void_p(rp), C.c_void_p(ci) if fill else None, C.c_void_p(va) if fill else None, C.byref(nnz))

    def fn():
        for fill in (False, True):
            g.timer_start()
            if kind == 0:
                g.call("sysml_b200_csr_dense_op", *csr3, rows, cols, p_(y), 1, 2, 0, *outs(fill))
            elif kind == 1:
                g.call("sysml_b200_csr_scalar_op", *csr3, rows, cols, 0.5, 2, 0, *outs(fill))
            else:
                g.call("sysml_b200_csr_csr_op", *csr3, *csr3, rows, cols, 2, *outs(fill))
            ms = g.timer_stop_ms()
            # symbolic: reads 12 B per cell + 4 B per row (+ 8 B per row for y); numeric: the same reads + 12 B per cell written
            rd = S.csr.nnz * 12 * (2 if kind == 2 else 1) + rows * 4 * (2 if kind == 2 else 1) + (rows * 8 if kind == 0 else 0)
            by = rd + (nnz.value * 12 + rows * 4 if fill else rows * 4)
            print("csrcell kind", kind, "numeric" if fill else "symbolic", "ms", round(ms, 4), "GB/s", round(by / ms / 1e6, 1), "nnz", S.csr.nnz, "->", nnz.value)
elif op == "kmeans":
    n, m, k = a
    X, Cm = g.rand(n, m, seed=1, lo=-10.0, hi=10.0), g.rand(k, m, seed=2, lo=-10.0, hi=10.0)
    c2, sums, cnt, smin = g.empty(k, 1), g.empty(k, m), g.empty(k, 1), g.empty(1, 1)
    g.call("sysml_b200_agg_unary_dense", p_(Cm), k, m, 1, 1, p_(c2))
    fn = lambda: g.call("sysml_b200_kmeans_step", p_(X), p_(Cm), p_(c2), n, m, k, p_(sums), p_(cnt), p_(smin))
elif op == "agg":
    n, m, aop, d = a
    X = g.rand(n, m, seed=1)
    o = g.empty(max(n, m), 1)
    fn = lambda: g.call("sysml_b200_agg_unary_dense", p_(X), n, m, aop, d, p_(o))
else:
    raise SystemExit("unknown op")
for _ in range(5 if op.startswith("mmchain") else 3):
    g.timer_start()
    fn()
    print(op, a, "ms", round(g.timer_stop_ms(), 4))
g.destroy()
