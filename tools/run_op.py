#!/usr/bin/env python
"""Runs ONE op a few times (for ncu captures): python tools/run_op.py tsmm 2000000 1000 | mmchain n m | matmult M K N ta"""
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
    rows, cols, per = a
    rng = np.random.default_rng(2)
    ci = (np.arange(per, dtype=np.int64)[None, :] * (cols // per) + rng.integers(0, cols // per, size=(rows, per))).astype(np.int32).reshape(-1)
    va = rng.uniform(-1, 1, rows * per)
    rp = (np.arange(rows + 1, dtype=np.int64) * per).astype(np.int32)
    S = g.csr(rp, ci, va, (rows, cols))
    v, o = g.rand(cols, 1, seed=2), g.empty(rows, 1)
    fn = lambda: g.call("sysml_b200_matmult_csr_dense", C.c_void_p(S.csr.rowPtr), C.c_void_p(S.csr.colInd), C.c_void_p(S.csr.val), p_(v), p_(o), rows, cols, cols, 1, 0)
elif op == "agg":
    n, m, aop, d = a
    X = g.rand(n, m, seed=1)
    o = g.empty(max(n, m), 1)
    fn = lambda: g.call("sysml_b200_agg_unary_dense", p_(X), n, m, aop, d, p_(o))
else:
    raise SystemExit("unknown op")
for _ in range(3):
    g.timer_start()
    fn()
    print(op, a, "ms", round(g.timer_stop_ms(), 4))
g.destroy()
