#!/usr/bin/env python
"""X %*% v on the synthetic C3-shaped CSR matrix (for ncu captures): python tools/run_spmv.py rows cols per"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from systemml_b200 import GPUContext  # noqa: E402

rows, cols, per = (int(x) for x in sys.argv[1:4])
g = GPUContext(0)
S = g.rand_csr(rows, cols, per, seed=2)
v, o = g.rand(cols, 1, seed=2), g.empty(rows, 1)
csr3 = (C.c_void_p(S.csr.rowPtr), C.c_void_p(S.csr.colInd), C.c_void_p(S.csr.val))
for _ in range(3):
    g.timer_start()
    g.call("sysml_b200_matmult_csr_dense", *csr3, C.c_void_p(v.ptr), C.c_void_p(o.ptr), rows, cols, cols, 1, 0)
    print("spmv ms", round(g.timer_stop_ms(), 4))
g.destroy()
