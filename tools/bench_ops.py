#!/usr/bin/env python
"""Per-op timing on one GPU through the C ABI (CUDA events on the context's stream): the shapes of
BASELINE configs C1-C5.  Prints one JSON object per op with achieved GB/s or TFLOP/s and the roofline
fraction (HBM copy peak from MEASURED_PEAKS.json, FP64 DMMA peak from profiles/fp64_peak.json).

    python tools/bench_ops.py [--only tsmm,mmchain,...] [--scale 1.0] [--reps 5]
"""
import argparse
import ctypes as C
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from systemml_b200 import GPUContext, GPUObject, LibMatrixB200 as L  # noqa: E402


def peaks():
    hbm, dmma = 6650.0, 37.0
    try:
        hbm = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    try:
        dmma = float(json.load(open(os.path.join(ROOT, "profiles", "fp64_peak.json")))["dmma_peak_tflops"])
    except Exception:
        pass
    return hbm, dmma


def timeit(g, fn, reps):
    fn()
    g.sync()
    best, tot = 1e30, 0.0
    for _ in range(reps):
        g.timer_start()
        fn()
        ms = g.timer_stop_ms()
        best = min(best, ms)
        tot += ms
    return best, tot / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="")
    ap.add_argument("--scale", type=float, default=1.0, help="scale the row counts (1.0 = full BASELINE sizes)")
    ap.add_argument("--reps", type=int, default=5)
    args = ap.parse_args()
    only = set(x for x in args.only.split(",") if x)
    hbm, dmma = peaks()
    g = GPUContext(0)
    S = args.scale
    out = []

    def want(name):
        return not only or any(name.startswith(o) for o in only)

    def report(name, shape, ms_best, ms_mean, nbytes=None, flops=None, bound="hbm"):
        r = {"op": name, "shape": shape, "ms_best": round(ms_best, 4), "ms_mean": round(ms_mean, 4)}
        if nbytes is not None:
            r["GBs"] = round(nbytes / ms_mean / 1e6, 1)
            r["frac_hbm"] = round(nbytes / ms_mean / 1e6 / hbm, 3)
        if flops is not None:
            r["TFLOPs"] = round(flops / ms_mean / 1e9, 2)
            r["frac_dmma"] = round(flops / ms_mean / 1e9 / dmma, 3)
        r["bound"] = bound
        print(json.dumps(r), flush=True)
        out.append(r)

    p_ = lambda o: C.c_void_p(o.ptr)

    # ---------------- mmchain / MV / VM ----------------
    for (n, m, tag) in ((100_000, 1000, "C1"), (int(10_000_000 * S), 1000, "C2"), (int(10_000_000 * S), 100, "C5shape")):
        if not want("mmchain") and not want("mv") and not want("vm"):
            break
        X = g.rand(n, m, seed=1)
        v, w, y = g.rand(m, 1, seed=2), g.rand(n, 1, seed=3, lo=0.0, hi=1.0), g.rand(n, 1, seed=4)
        o_m, o_n = g.empty(m, 1), g.empty(n, 1)
        if want("mmchain"):
            for ct, nm in ((0, "XtXv"), (1, "XtwXv")):
                b, a = timeit(g, lambda: g.call("sysml_b200_mmchain_dense", p_(X), p_(v), p_(w), p_(o_m), n, m, ct), args.reps)
                report("mmchain_%s_%s" % (nm, tag), [n, m], b, a, nbytes=8.0 * n * m + 16 * m + (8 * n if ct else 0), flops=None)
        if want("mv"):
            b, a = timeit(g, lambda: g.call("sysml_b200_matmult_dense", p_(X), p_(v), p_(o_n), n, m, m, 1, 0, 0), args.reps)
            report("mv_Xv_%s" % tag, [n, m], b, a, nbytes=8.0 * n * m + 8 * m + 8 * n)
        if want("vm"):
            b, a = timeit(g, lambda: g.call("sysml_b200_matmult_dense", p_(X), p_(y), p_(o_m), n, m, n, 1, 1, 0), args.reps)
            report("vm_tXy_%s" % tag, [n, m], b, a, nbytes=8.0 * n * m + 8 * m + 8 * n)
        for o in (X, v, w, y, o_m, o_n):
            o.free()

    # ---------------- tsmm ----------------
    if want("tsmm"):
        for (n, m, tag) in ((int(10_000_000 * S), 1000, "C2shape"), (int(1_000_000 * S), 5000, "C4"), (int(2_000_000 * S), 1024, "pow2"),
                            (int(2_000_000 * S), 1000, "even_2M"), (int(2_000_000 * S), 999, "odd_2Mx999"), (int(1_000_000 * S), 5001, "odd_1Mx5001"),
                            (int(2_000_001 * S), 1000, "oddrows_2000001x1000")):
            X = g.rand(n, m, seed=1)
            o = g.empty(m, m)
            b, a = timeit(g, lambda: g.call("sysml_b200_tsmm_dense", p_(X), p_(o), n, m, 1, None), max(2, args.reps // 2))
            report("tsmm_left_%s" % tag, [n, m], b, a, flops=float(n) * m * (m + 1), bound="dmma")
            X.free()
            o.free()

    # ---------------- dense matmult (C4 / C5 skinny right-hand sides) ----------------
    if want("matmult"):
        for (M, K, N, ta, tag) in ((int(1_000_000 * S), 5000, 20, 0, "C4_XV"), (int(1_000_000 * S), 5000, 20, 1, "C4_tXP"),
                                   (int(10_000_000 * S), 100, 50, 0, "C5_XtC"), (int(10_000_000 * S), 100, 50, 1, "C5_tXP_100x50"),
                                   (int(10_000_000 * S), 50, 100, 1, "C5_tPX_50x100_kmeans"),
                                   (8192, 8192, 8192, 0, "square8k")):
            if ta:  # t(A) %*% B with A: M x K stored, B: M x N  -> K x N
                A, B = g.rand(M, K, seed=1), g.rand(M, N, seed=2)
                o = g.empty(K, N)
                fn = lambda: g.call("sysml_b200_matmult_dense", p_(A), p_(B), p_(o), M, K, M, N, 1, 0)
                nbytes = 8.0 * (M * K + M * N + K * N)
            else:
                A, B = g.rand(M, K, seed=1), g.rand(K, N, seed=2)
                o = g.empty(M, N)
                fn = lambda: g.call("sysml_b200_matmult_dense", p_(A), p_(B), p_(o), M, K, K, N, 0, 0)
                nbytes = 8.0 * (M * K + K * N + M * N)
            b, a = timeit(g, fn, max(2, args.reps // 2))
            report("matmult_%s" % tag, [M, K, N, "tA" if ta else ""], b, a, nbytes=nbytes, flops=2.0 * M * K * N, bound="hbm|dmma")
            for x in (A, B, o):
                x.free()

    # ---------------- aggregates + cellwise (C5: 10M x 50) ----------------
    if want("agg") or want("cell"):
        n, m = int(10_000_000 * S), 50
        X = g.rand(n, m, seed=1)
        o_n, o_m, o_1 = g.empty(n, 1), g.empty(1, m), g.empty(1, 1)
        if want("agg"):
            for name, op, d, o in (("sum", 0, 0, o_1), ("rowSums", 0, 1, o_n), ("colSums", 0, 2, o_m), ("rowMins", 3, 1, o_n),
                                   ("rowIndexMax", 4, 1, o_n), ("colMaxs", 2, 2, o_m), ("sumsq", 1, 0, o_1)):
                b, a = timeit(g, lambda: g.call("sysml_b200_agg_unary_dense", p_(X), n, m, op, d, p_(o)), args.reps)
                report("agg_%s" % name, [n, m], b, a, nbytes=8.0 * n * m + 8 * (1 if d == 0 else (n if d == 1 else m)))
        if want("cell"):
            Y, Z = g.rand(n, m, seed=2), g.empty(n, m)
            cv, rv = g.rand(n, 1, seed=3), g.rand(1, m, seed=4)
            for name, fn, nb in (
                ("mul_mm", lambda: g.call("sysml_b200_binary_mm", p_(X), p_(Y), p_(Z), n, m, 0, 2), 24.0 * n * m),
                ("div_colvec", lambda: g.call("sysml_b200_binary_mm", p_(X), p_(cv), p_(Z), n, m, 1, 3), 16.0 * n * m + 8 * n),
                ("plus_rowvec", lambda: g.call("sysml_b200_binary_mm", p_(X), p_(rv), p_(Z), n, m, 2, 0), 16.0 * n * m + 8 * m),
                ("le_colvec", lambda: g.call("sysml_b200_binary_mm", p_(X), p_(cv), p_(Z), n, m, 1, 6), 16.0 * n * m + 8 * n),
                ("scalar_mul", lambda: g.call("sysml_b200_binary_ms", p_(X), -2.0, p_(Z), n * m, 2, 0), 16.0 * n * m),
                ("exp", lambda: g.call("sysml_b200_unary", p_(X), p_(Z), n * m, 0), 16.0 * n * m),
                ("axpy", lambda: g.call("sysml_b200_axpy", p_(X), p_(Y), 0.5, p_(Z), n, m, 0), 24.0 * n * m),
            ):
                b, a = timeit(g, fn, args.reps)
                report("cell_%s" % name, [n, m], b, a, nbytes=nb)
            for x in (Y, Z, cv, rv):
                x.free()
        for x in (X, o_n, o_m, o_1):
            x.free()

    # ---------------- fused cellwise chains vs the same chain one launch per operator ----------------
    if want("fused"):
        from systemml_b200.lib_matrix import Expr as E
        n = int(50_000_000 * S)  # C3's row count: the n-vectors of the L2SVM inner loop
        Y, Xw, Xd = g.rand(n, 1, seed=2), g.rand(n, 1, seed=3), g.rand(n, 1, seed=4)
        t1, t2, t3 = g.empty(n, 1), g.empty(n, 1), g.empty(n, 1)
        o1 = g.empty(1, 1)
        e_out = E.const(0) - E.input(0) * E.input(1)
        e_sv = e_out.binary(E.const(1), ">")
        gexpr = e_out * e_sv * E.input(0) * E.input(2)  # sum(out * sv * Y * Xd), l2-svm.dml

        def fused_fn(objs, kinds, expr, consts, rows, cols, agg, outobj):
            ptrs = (C.c_void_p * len(objs))(*[o.ptr for o in objs])
            kd = (C.c_int * len(objs))(*kinds)
            cs = (C.c_double * max(len(consts), 1))(*consts)
            pg = (C.c_int32 * len(expr.tokens))(*expr.tokens)
            return lambda: g.call("sysml_b200_fused_cellwise", ptrs, kd, len(objs), cs, len(consts), pg, len(expr.tokens), rows, cols, agg, p_(outobj))
        fused = fused_fn([Y, Xw, Xd], [0, 0, 0], gexpr, (1.0, 0.0), n, 1, 1, o1)

        def unfused():
            g.call("sysml_b200_binary_mm", p_(Y), p_(Xw), p_(t1), n, 1, 0, 15)      # 1 - Y*Xw
            g.call("sysml_b200_binary_ms", p_(t1), 0.0, p_(t2), n, 7, 0)             # out > 0
            g.call("sysml_b200_binary_mm", p_(t1), p_(t2), p_(t3), n, 1, 0, 2)       # out * sv
            g.call("sysml_b200_binary_mm", p_(t3), p_(Y), p_(t3), n, 1, 0, 2)        # * Y
            g.call("sysml_b200_binary_mm", p_(t3), p_(Xd), p_(t3), n, 1, 0, 2)       # * Xd
            g.call("sysml_b200_agg_unary_dense", p_(t3), n, 1, 0, 0, p_(o1))         # sum
        b, a = timeit(g, fused, args.reps)
        report("fused_l2svm_sum_chain", [n, 1], b, a, nbytes=24.0 * n)
        b, a = timeit(g, unfused, args.reps)
        report("unfused_l2svm_sum_chain_6_launches", [n, 1], b, a, nbytes=24.0 * n)
        for x in (Y, Xw, Xd, t1, t2, t3, o1):
            x.free()
        # Kmeans: P = ((-2 * XC + rowvec) <= colvec), 10M x 50
        n, m = int(10_000_000 * S), 50
        XC, Z, T = g.rand(n, m, seed=5), g.empty(n, m), g.empty(n, m)
        rv, cv = g.rand(1, m, seed=6), g.rand(n, 1, seed=7)
        kexpr = (E.const(0) * E.input(0) + E.input(1)).binary(E.input(2), "<=")
        fused = fused_fn([XC, rv, cv], [0, 2, 1], kexpr, (-2.0,), n, m, 0, Z)

        def unfused2():
            g.call("sysml_b200_binary_ms", p_(XC), -2.0, p_(T), n * m, 2, 1)
            g.call("sysml_b200_binary_mm", p_(T), p_(rv), p_(T), n, m, 2, 0)
            g.call("sysml_b200_binary_mm", p_(T), p_(cv), p_(Z), n, m, 1, 6)
        b, a = timeit(g, fused, args.reps)
        report("fused_kmeans_assign_chain", [n, m], b, a, nbytes=16.0 * n * m + 8 * n + 8 * m)
        b, a = timeit(g, unfused2, args.reps)
        report("unfused_kmeans_assign_chain_3_launches", [n, m], b, a, nbytes=16.0 * n * m + 8 * n + 8 * m)
        # row-wise endings on the same narrow matrix: rowSums(XC * Z') and rowMins(-2*XC + rv) in one pass each
        rs = g.empty(n, 1)
        g.call("sysml_b200_rand_uniform", p_(T), n * m, 9, 0, -1.0, 1.0)
        f1 = fused_fn([XC, T], [0, 0], E.input(0) * E.input(1), (), n, m, 2, rs)
        b, a = timeit(g, f1, args.reps)
        report("fused_rowSums_of_product", [n, m], b, a, nbytes=16.0 * n * m + 8 * n)
        f2 = fused_fn([XC, rv], [0, 2], E.const(0) * E.input(0) + E.input(1), (-2.0,), n, m, 5, rs)
        b, a = timeit(g, f2, args.reps)
        report("fused_rowMins_of_distance_chain", [n, m], b, a, nbytes=8.0 * n * m + 8 * n + 8 * m)
        rs.free()
        for x in (XC, Z, T, rv, cv):
            x.free()

    # ---------------- CSR (C3 scaled: 20M x 10k, ~Poisson(10) nnz/row) ----------------
    if want("csr"):
        rows, cols, per = int(20_000_000 * S), 10_000, 10
        from csr_gen import poisson_rows_csr
        rp, ci, va = poisson_rows_csr(rows, cols, per, 2)  # row lengths ~Poisson(10), as C3's Bernoulli cells give
        Xs = g.csr(rp, ci, va, (rows, cols))
        v, y, w = g.rand(cols, 1, seed=2), g.rand(rows, 1, seed=3), g.rand(rows, 1, seed=4, lo=0.0, hi=1.0)
        o_r, o_c = g.empty(rows, 1), g.empty(cols, 1)
        nnz = int(rp[-1])
        spb = 12.0 * nnz + 4 * (rows + 1) + 8 * cols + 8 * rows
        args_csr = (p_(GPUObject.wrap(g, Xs.csr.rowPtr, 1, 1)), p_(GPUObject.wrap(g, Xs.csr.colInd, 1, 1)), p_(GPUObject.wrap(g, Xs.csr.val, 1, 1)))
        b, a = timeit(g, lambda: g.call("sysml_b200_matmult_csr_dense", *args_csr, p_(v), p_(o_r), rows, cols, cols, 1, 0), args.reps)
        report("csr_spmv", [rows, cols, nnz], b, a, nbytes=spb)
        b, a = timeit(g, lambda: g.call("sysml_b200_matmult_csr_dense", *args_csr, p_(y), p_(o_c), rows, cols, rows, 1, 1), args.reps)
        report("csr_tXy", [rows, cols, nnz], b, a, nbytes=spb)
        b, a = timeit(g, lambda: g.call("sysml_b200_mmchain_csr", *args_csr, p_(v), p_(w), p_(o_c), rows, cols, 1), args.reps)
        report("csr_mmchain_XtwXv", [rows, cols, nnz], b, a, nbytes=spb + 8 * rows + 8 * cols)
        b, a = timeit(g, lambda: g.call("sysml_b200_agg_unary_csr", *args_csr, rows, cols, 0, 1, p_(o_r)), args.reps)
        report("csr_rowSums", [rows, cols, nnz], b, a, nbytes=12.0 * nnz + 4 * (rows + 1) + 8 * rows)

    g.destroy()


if __name__ == "__main__":
    main()
