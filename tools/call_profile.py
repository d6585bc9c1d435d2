#!/usr/bin/env python
"""Where an algorithm's wall time goes, per native entry point: wraps GPUContext.call so that every call is bracketed by a
stream synchronise and timed on the host, then prints the total per entry point next to the untouched end-to-end time.
(The synchronises serialise host and device, so the instrumented run is slower than the real one: the point is the
split, and the gap between the sum of the calls and the untouched run = host-side work hidden behind the device.)

  python tools/call_profile.py l2svm [scale] [column_copy]     C3: CSR 50M x 10k * scale
  python tools/call_profile.py multilogreg [scale]             C4: 1M x 5000 * scale, K = 20
"""
import collections
import ctypes as C
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

from systemml_b200 import GPUContext, LibMatrixB200 as L  # noqa: E402

_p = lambda o: C.c_void_p(o.ptr)


def profiled(g, fn):
    g.sync()
    t0 = time.perf_counter()
    r = fn()
    g.sync()
    plain = time.perf_counter() - t0
    acc, cnt = collections.defaultdict(float), collections.defaultdict(int)
    orig = g.call

    def call(name, *a):
        if name == "sysml_b200_sync":
            return orig(name, *a)
        orig("sysml_b200_sync")
        t = time.perf_counter()
        orig(name, *a)
        orig("sysml_b200_sync")
        acc[name] += time.perf_counter() - t
        cnt[name] += 1

    g.call = call
    t0 = time.perf_counter()
    fn()
    g.sync()
    instr = time.perf_counter() - t0
    g.call = orig
    print("end to end, untouched: %.4f s; instrumented: %.4f s; sum of native calls: %.4f s" % (plain, instr, sum(acc.values())))
    for k, v in sorted(acc.items(), key=lambda kv: -kv[1]):
        print("  %-40s %5d calls  %9.3f ms  (%.3f ms each)" % (k, cnt[k], v * 1e3, v * 1e3 / cnt[k]))
    return r


def main():
    what = sys.argv[1]
    scale = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
    g = GPUContext(0)
    if what == "l2svm":
        from systemml_b200.algorithms import l2svm
        cc = len(sys.argv) > 3 and sys.argv[3] not in ("0", "false")
        n, m = int(50_000_000 * scale), 10_000
        X = g.rand_csr(n, m, 10, seed=2)
        csr = (C.c_void_p(X.csr.rowPtr), C.c_void_p(X.csr.colInd), C.c_void_p(X.csr.val))
        v, o_r, Y = g.rand(m, 1, seed=2), g.empty(n, 1), g.empty(n, 1)
        g.call("sysml_b200_matmult_csr_dense", *csr, _p(v), _p(o_r), n, m, m, 1, 0)
        g.call("sysml_b200_binary_ms", _p(o_r), 0.0, _p(Y), n, L.BINARY_OPS[">"], 0)
        g.call("sysml_b200_binary_ms", _p(Y), 2.0, _p(Y), n, L.BINARY_OPS["*"], 0)
        g.call("sysml_b200_binary_ms", _p(Y), 1.0, _p(Y), n, L.BINARY_OPS["-"], 0)
        o_r.free()
        l2svm(g, X, Y, reg=1.0, tol=0.001, maxiter=1, column_copy=cc)
        _, it, obj = profiled(g, lambda: l2svm(g, X, Y, reg=1.0, tol=0.001, maxiter=20, column_copy=cc))
        print("outer iterations", it, "objective", obj)
    elif what == "multilogreg":
        from systemml_b200.algorithms import multilogreg
        E = L.Expr
        N, D, K = int(1_000_000 * scale), 5000, 20
        X, W, sc = g.rand(N, D, seed=3), g.rand(D, K + 1, seed=7), g.empty(N, K + 1)
        g.call("sysml_b200_matmult_dense", _p(X), _p(W), _p(sc), N, D, D, K + 1, 0, 0)
        lab = L.unaryAggregate(g, "uarimax", sc, "uarimax")
        ids = g.dense(np.arange(1, K + 1, dtype=np.float64).reshape(1, K))
        Yk = L.fusedCellwise(g, "spoof", [lab, ids], E.input(0).binary(E.input(1), "=="), rows=N, cols=K)
        ylast = L.matrixScalarArithmetic(g, "==", lab, float(K + 1), "==")
        multilogreg(g, X, Yk, ylast, reg=1.0, tol=1e-6, maxiter=1)
        _, it, obj = profiled(g, lambda: multilogreg(g, X, Yk, ylast, reg=1.0, tol=1e-6, maxiter=5))
        print("outer iterations", it, "objective", obj)
    g.destroy()


if __name__ == "__main__":
    main()
