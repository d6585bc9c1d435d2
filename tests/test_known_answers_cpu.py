"""Known-answer tables for the semantics that only the reference's SOURCE pins (no golden vectors exist for them): every
expected value below was derived by hand from the cited lines, not produced by running the oracle.  They hold the oracle
(oracle/sysml_oracle.cpp) to a second, independent reading of those lines; the GPU tests then hold the kernels to the oracle.
Paths: J/ = /root/reference/src/main/java/org/apache/sysml/runtime/."""
import math

import numpy as np
import pytest

from oracle import oracle as O

NAN, INF = float("nan"), float("inf")
EPS53 = 2.0 ** -53


def same(a, b):
    """equal as doubles, NaN == NaN, and the sign of a zero matters"""
    if a != a or b != b:
        return a != a and b != b
    return a == b and math.copysign(1.0, a) == math.copysign(1.0, b)


# ---- %/% : IntegerDivide.executeIntDiv (J/functionobjects/IntegerDivide.java:62-73) = floor(fl(x/y) + 2^-53) unless the
# quotient is NaN or infinite; UtilFunctions.toLong (J/util/UtilFunctions.java:327-329), DOUBLE_EPS = 2^-53 (:44) ----
INTDIV = [
    (7.0, 2.0, 3.0),            # 3.5 + 2^-53 = 3.5 (absorbed) -> floor 3
    (-7.0, 2.0, -4.0),          # -3.5 -> floor -4 (floor, not truncation)
    (7.0, -2.0, -4.0),
    (6.0, 3.0, 2.0),
    (1.0, 0.0, INF),            # infinite quotient is returned as it is
    (-1.0, 0.0, -INF),
    (0.0, 0.0, NAN),
    (NAN, 2.0, NAN),
    (0.3, 0.1, 2.0),            # fl(0.3/0.1) = 2.9999999999999996 = 3 - 2^-51; + 2^-53 rounds back to it -> floor 2
    (1.0 - EPS53, 1.0, 1.0),    # quotient 1 - 2^-53; + 2^-53 = 1 exactly -> floor 1: the visible effect of DOUBLE_EPS
    (-(1.0 - EPS53), 1.0, -1.0),  # -(1 - 2^-53) + 2^-53 = -1 + 2^-52 -> floor -1
    (5.0, INF, 0.0),            # quotient +0 -> floor(2^-53) = 0
    (-5.0, INF, 0.0),           # quotient -0 -> floor(-0 + 2^-53) = 0 (a long has no negative zero)
]


@pytest.mark.parametrize("x,y,exp", INTDIV)
def test_integer_divide(x, y, exp):
    got = O.binary_scalar(x, y, 18)
    assert same(got, exp), (x, y, got, exp)


# ---- %% : Modulus.execute (J/functionobjects/Modulus.java:49-53) = NaN when y == 0, else x - (x %/% y) * y ----
MODULUS = [
    (7.0, 2.0, 1.0),
    (-7.0, 2.0, 1.0),           # -7 - (-4)*2
    (7.0, -2.0, -1.0),          # 7 - (-4)*(-2)
    (5.5, 0.0, NAN),
    (5.5, -0.0, NAN),
    (0.3, 0.1, 0.3 - 2.0 * 0.1),  # 0.09999999999999998: follows from 0.3 %/% 0.1 = 2
    (6.0, 3.0, 0.0),
    (INF, 2.0, NAN),            # inf - inf*2 = NaN
]


@pytest.mark.parametrize("x,y,exp", MODULUS)
def test_modulus(x, y, exp):
    got = O.binary_scalar(x, y, 17)
    assert same(got, exp), (x, y, got, exp)


# ---- round : Builtin ROUND = (double) Math.round(x) (J/functionobjects/Builtin.java:314); java.lang.Math.round (JDK 8):
# the closest long, ties towards +infinity; NaN -> 0; saturation at the long range.  NOT floor(fl(x + 0.5)). ----
ROUND = [
    (2.5, 3.0), (-2.5, -2.0), (0.5, 1.0), (-0.5, 0.0), (3.5, 4.0), (-3.5, -3.0),
    (0.49999999999999994, 0.0),          # fl(x + 0.5) = 1.0, but the closest long is 0
    (-0.5000000000000001, -1.0),
    (-0.2, 0.0), (-0.0, 0.0), (0.0, 0.0),  # a long has no negative zero
    (4503599627370497.0, 4503599627370497.0),   # 2^52 + 1: fl(x + 0.5) would round to 2^52 + 2
    (-4503599627370497.0, -4503599627370497.0),
    (NAN, 0.0), (INF, 9223372036854775808.0), (-INF, -9223372036854775808.0), (1e300, 9223372036854775808.0),
    (2.4999999999999996, 2.0),
]


@pytest.mark.parametrize("x,exp", ROUND)
def test_round(x, exp):
    got = float(O.java_round(np.array([x]))[0])
    assert same(got, exp), (x, got, exp)


# ---- sign : Math.signum (J/functionobjects/Builtin.java:311) ----
@pytest.mark.parametrize("x,exp", [(3.0, 1.0), (-0.1, -1.0), (0.0, 0.0), (-0.0, -0.0), (NAN, NAN), (-INF, -1.0)])
def test_signum(x, exp):
    assert same(float(O.java_signum(np.array([x]))[0]), exp)


# ---- min / max : Builtin.execute with Math.min / Math.max (J/functionobjects/Builtin.java:356-361): NaN propagates,
# -0.0 < +0.0 ----
@pytest.mark.parametrize("x,y,mx,mn", [(1.0, 2.0, 2.0, 1.0), (NAN, 2.0, NAN, NAN), (2.0, NAN, NAN, NAN), (-0.0, 0.0, 0.0, -0.0),
                                        (0.0, -0.0, 0.0, -0.0), (-INF, INF, INF, -INF)])
def test_min_max(x, y, mx, mn):
    assert same(O.binary_scalar(x, y, 12), mx) and same(O.binary_scalar(x, y, 11), mn)


# ---- 1-* , -nz , && , || (J/functionobjects/{Minus1Multiply.java:41-43, MinusNz.java:42-44, And.java:46-48, Or.java:46-48}) ----
@pytest.mark.parametrize("op,x,y,exp", [(15, 0.5, 0.5, 0.75), (15, 3.0, 2.0, -5.0), (16, 0.0, 5.0, 0.0), (16, 2.0, 5.0, -3.0), (16, -0.0, 5.0, 0.0),
                                         (13, 2.0, -1.0, 1.0), (13, 2.0, 0.0, 0.0), (13, NAN, 1.0, 1.0), (14, 0.0, 0.0, 0.0), (14, 0.0, -3.0, 1.0)])
def test_other_value_functions(op, x, y, exp):
    assert same(O.binary_scalar(x, y, op), exp), (op, x, y)


# ---- rowIndexMax / rowIndexMin, dense: d_uarimxx / d_uarimin + indexmax / indexmin (J/matrix/data/LibMatrixAgg.java:1902-1927,
# 3081-3099): 1-based index, `>=` / `<=` so the LAST extremum wins, a NaN never wins, an all-NaN row gives index 0 (maxindex
# stays -1, +1) ----
def test_row_index_max_min_dense():
    X = np.array([[1.0, 3.0, 3.0, 2.0],       # max 3 at columns 2 and 3 -> 3 ; min 1 at column 1 -> 1
                  [5.0, 5.0, 5.0, 5.0],       # all equal -> last column, 4, for both
                  [NAN, 7.0, NAN, 6.0],       # NaN never wins -> max at 2, min at 4
                  [NAN, NAN, NAN, NAN],       # nothing wins -> 0
                  [-INF, -INF, -INF, -INF],   # equals the initial value -INF: `>=` accepts every column -> 4; min: -INF <= +INF -> 4
                  [-0.0, 0.0, -0.0, 0.0]])    # -0.0 >= 0.0 is true for doubles (comparison, not Math.max) -> 4 for both
    assert O.agg_unary_dense(X, O.MAXINDEX, O.ROW).reshape(-1).tolist() == [3.0, 4.0, 2.0, 0.0, 4.0, 4.0]
    assert O.agg_unary_dense(X, O.MININDEX, O.ROW).reshape(-1).tolist() == [1.0, 4.0, 4.0, 0.0, 4.0, 4.0]


# ---- rowIndexMax / rowIndexMin, sparse: s_uarimxx / s_uarimin (LibMatrixAgg.java:2536-2614): the extremum over the STORED
# values first; only if an implicit zero STRICTLY beats it (builtin.execute(0, best) == 1, Builtin.java:374-381) the result
# is the LAST column that holds an implicit zero (:2550-2557); an empty row gives index n = number of columns (:2559-2563) ----
def test_row_index_max_min_sparse():
    cols = 5
    rows = [
        {1: -2.0, 3: -5.0},            # max: stored best -2 at column 2, but 0 > -2: last implicit zero is column 5 -> 5
                                        # min: stored best -5 at column 4; 0 < -5 is false -> 4
        {0: 4.0, 2: 9.0, 4: 9.0},      # max: 9 at columns 3 and 5, last wins -> 5 ; min: stored best 4 at column 1, but 0 < 4:
                                        #      last implicit zero is column 4 -> 4
        {},                             # empty row -> 5 for both
        {0: 1.0, 1: 2.0, 2: 3.0, 3: 4.0, 4: 5.0},   # full row: no implicit zero -> max 5, min 1
        {4: -1.0},                      # max: 0 > -1 -> last implicit zero = column 4 -> 4 ; min: -1 at column 5 -> 5
    ]
    rp, ci, va = [0], [], []
    for r in rows:
        for c in sorted(r):
            ci.append(c)
            va.append(r[c])
        rp.append(len(ci))
    rp, ci, va = np.array(rp, dtype=np.int64), np.array(ci, dtype=np.int32), np.array(va)
    assert O.agg_unary_csr(rp, ci, va, (len(rows), cols), O.MAXINDEX, O.ROW).reshape(-1).tolist() == [5.0, 5.0, 5.0, 5.0, 4.0]
    assert O.agg_unary_csr(rp, ci, va, (len(rows), cols), O.MININDEX, O.ROW).reshape(-1).tolist() == [4.0, 4.0, 5.0, 1.0, 5.0]


# ---- sum with Kahan compensation: KahanPlus.execute (J/functionobjects/KahanPlus.java:89-107) keeps what a plain sum loses ----
def test_kahan_sum_keeps_small_terms():
    # 1e16 + 1 + 1 + ... (1000 ones): a plain left-to-right double sum stays at 1e16 (1 is half an ulp of 1e16 and ties to
    # even); the compensated sum carries the ones in its correction term and returns 1e16 + 1000 exactly
    X = np.ones((1, 1001))
    X[0, 0] = 1e16
    assert float(np.cumsum(X)[-1]) == 1e16                       # what the loss looks like
    assert O.agg_unary_dense(X, O.SUM, O.ALL)[0] == 1e16 + 1000.0
    assert O.agg_unary_dense(X, O.SUM, O.ROW)[0] == 1e16 + 1000.0
    # +-Inf fast path (KahanPlus.java:89-102): the sum becomes infinite instead of NaN from inf - inf in the correction
    Y = np.array([[1.0, INF, 2.0, 3.0]])
    assert O.agg_unary_dense(Y, O.SUM, O.ALL)[0] == INF


# ---- variance: CM operator (J/functionobjects/CM.java:166-176), the sample variance m2 / (n - 1) ----
def test_variance_known_answers():
    X = np.array([[1.0, 2.0, 3.0, 4.0], [2.0, 2.0, 2.0, 2.0], [0.0, 0.0, 10.0, 10.0]])
    got = O.agg_unary_dense(X, O.VAR, O.ROW).reshape(-1)
    assert abs(got[0] - 5.0 / 3.0) < 1e-15 and got[1] == 0.0 and abs(got[2] - 100.0 / 3.0) < 1e-13
    col = O.agg_unary_dense(X, O.VAR, O.COL).reshape(-1)           # columns: (1,2,0), (2,2,0), (3,2,10), (4,2,10)
    exp = [1.0, 4.0 / 3.0, 19.0, 52.0 / 3.0]
    assert all(abs(a - b) < 1e-13 for a, b in zip(col, exp))
