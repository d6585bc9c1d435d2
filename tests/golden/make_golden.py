"""Generates tests/golden/*.json: small inputs with results computed in EXACT rational arithmetic
(python fractions), independent of both the oracle and the CUDA path.  Inputs are dyadic rationals so
they are exactly representable as doubles.

    python tests/golden/make_golden.py

The reference itself (Java) cannot run in this image, so these are not outputs of the reference; they
pin the math the reference implements (R-style definitions in src/test/scripts/functions/binary/matrix/
MapMultChain*.R, TransposeSelfMatrixMultiplication*.R) independently of summation order.
"""
import json
import os
import random
from fractions import Fraction as F

HERE = os.path.dirname(os.path.abspath(__file__))


def mat(rows, cols, rnd, sparsity=0.8):
    return [[F(rnd.randint(-4096, 4096), 1024) if rnd.random() < sparsity else F(0) for _ in range(cols)] for _ in range(rows)]


def fl(M):
    return [[float(x) for x in r] for r in M] if isinstance(M[0], list) else [float(x) for x in M]


def mm(A, B):
    return [[sum(A[i][k] * B[k][j] for k in range(len(B))) for j in range(len(B[0]))] for i in range(len(A))]


def tr(A):
    return [list(r) for r in zip(*A)]


def main():
    rnd = random.Random(20240607)
    # mmchain
    X = mat(53, 19, rnd)
    v = [F(rnd.randint(-1024, 1024), 512) for _ in range(19)]
    w = [F(rnd.randint(0, 2048), 1024) for _ in range(53)]
    Xv = [sum(X[i][j] * v[j] for j in range(19)) for i in range(53)]
    def xt(t):
        return [sum(X[i][j] * t[i] for i in range(53)) for j in range(19)]
    json.dump({"X": fl(X), "v": fl(v), "w": fl(w), "XtXv": fl(xt(Xv)), "XtwXv": fl(xt([w[i] * Xv[i] for i in range(53)])),
               "XtXvy": fl(xt([Xv[i] - w[i] for i in range(53)]))}, open(os.path.join(HERE, "mmchain_small.json"), "w"))
    # tsmm
    X = mat(41, 13, rnd, 0.6)
    L = mm(tr(X), X)
    R = mm(X, tr(X))
    json.dump({"X": fl(X), "left": fl(L), "right": fl(R), "left_nnz": sum(1 for r in L for x in r if x != 0)},
              open(os.path.join(HERE, "tsmm_small.json"), "w"))
    # matmult
    A = mat(37, 29, rnd, 0.5)
    B = mat(29, 11, rnd, 0.9)
    json.dump({"X": fl(A), "B": fl(B), "C": fl(mm(A, B))}, open(os.path.join(HERE, "matmult_small.json"), "w"))
    # aggregates
    X = mat(23, 17, rnd, 0.7)
    def imax(r):
        m = max(r); return max(j for j, x in enumerate(r) if x == m) + 1
    def imin(r):
        m = min(r); return max(j for j, x in enumerate(r) if x == m) + 1
    json.dump({"X": fl(X), "sum": float(sum(sum(r) for r in X)), "rowsums": [float(sum(r)) for r in X],
               "colsums": [float(sum(c)) for c in tr(X)], "sumsq": float(sum(x * x for r in X for x in r)),
               "rowindexmax": [imax(r) for r in X], "rowindexmin": [imin(r) for r in X]},
              open(os.path.join(HERE, "agg_small.json"), "w"))


if __name__ == "__main__":
    main()
