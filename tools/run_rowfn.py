import sys, ctypes as C
sys.path.insert(0, ".")
from systemml_b200 import GPUContext, LibMatrixB200 as L
g = GPUContext(0)
n, m = 4_000_000, 1000
X = g.rand(n, m, seed=1)
y = g.rand(n, 1, seed=5, lo=0.0, hi=3.0)
E = L.Expr
v = g.rand(m, 1, seed=2, lo=-0.01, hi=0.01)
mu_e = E.input(0).unary("exp")
out, side = L.mmchainRowFn(g, "mmchain", X, v, [E.input(1) - mu_e, mu_e], w=y)
for _ in range(2):
    L.mmchainRowFn(g, "mmchain", X, v, [E.input(1) - mu_e, mu_e], w=y, out=out, side=side)
g.sync()
