"""Per-operator timing of one Lloyd iteration of Kmeans on C5 (10M x 100, k = 50), fused plan."""
import sys, ctypes as C
sys.path.insert(0, '/root/repo')
import numpy as np
from systemml_b200 import GPUContext, LibMatrixB200 as L
from systemml_b200.lib_matrix import Expr as E
g = GPUContext(0)
p_ = lambda o: C.c_void_p(o.ptr)
n, m, k = 10_000_000, 100, 50
X = g.rand(n, m, seed=1); Cc = g.rand(k, m, seed=2)
XC, P = g.empty(n, k), g.empty(n, k)
c2, minD, rs, den, Cn = g.empty(1, k), g.empty(n, 1), g.empty(n, 1), g.empty(k, 1), g.empty(k, m)
d_e = E.const(0) * E.input(0) + E.input(1)
le = d_e.binary(E.input(2), "<=")
ops = [
 ("X%*%t(C)", lambda: g.call("sysml_b200_matmult_dense", p_(X), p_(Cc), p_(XC), n, m, k, m, 0, 1)),
 ("rowSums(C^2)", lambda: g.call("sysml_b200_agg_unary_dense", p_(Cc), k, m, 1, 1, p_(c2))),
 ("fused rowMins", lambda: L.fusedCellwise(g, "s", [XC, c2], d_e, (-2.0,), "rowMins", out=minD)),
 ("fused rowSums(le)", lambda: L.fusedCellwise(g, "s", [XC, c2, minD], le, (-2.0,), "rowSums", out=rs)),
 ("fused P (le / rs)", lambda: L.fusedCellwise(g, "s", [XC, c2, minD, rs], le / E.input(3), (-2.0,), out=P)),
 ("fused P (le * inv)", lambda: L.fusedCellwise(g, "s", [XC, c2, minD, rs], le * E.input(3), (-2.0,), out=P)),
 ("colSums(P)", lambda: g.call("sysml_b200_agg_unary_dense", p_(P), n, k, 0, 2, p_(den))),
 ("t(P)%*%X", lambda: g.call("sysml_b200_matmult_dense", p_(P), p_(X), p_(Cn), n, k, n, m, 1, 0)),
]
for name, fn in ops:
    fn(); g.sync()
    ts = []
    for _ in range(3):
        g.timer_start(); fn(); ts.append(g.timer_stop_ms())
    print("%-20s %.3f ms" % (name, min(ts)))
