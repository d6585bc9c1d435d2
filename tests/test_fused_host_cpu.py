"""Host side of the fused cellwise chains, no GPU: the expression builder's postfix linearisation (incl. shared
subexpressions through PICK / NIP) and the oracle's unfused evaluation of a program against plain numpy."""
import numpy as np
import pytest

from oracle import oracle as O
from systemml_b200.lib_matrix import BINARY_OPS, FUSED_AGG, UNARY_OPS, Expr as E


def tok(kind, arg):
    return (kind << 24) | arg


def test_postfix_linearisation():
    e = (E.const(0) - E.input(0) * E.input(1)).binary(E.const(1), ">")
    assert e.tokens == [tok(1, 0), tok(0, 0), tok(0, 1), tok(2, BINARY_OPS["*"]), tok(2, BINARY_OPS["-"]), tok(1, 1),
                        tok(2, BINARY_OPS[">"])]
    assert E.input(2).unary("exp").tokens == [tok(0, 2), tok(3, UNARY_OPS["exp"])]
    assert E([5, 6]).tokens == [5, 6]  # raw programs pass through untouched


def test_let_binding_uses_pick_and_nip():
    out = E.const(0) - E.input(0)
    e = E.let(out, lambda o: o * o.binary(E.const(1), ">"))
    t = e.tokens
    # value once, two PICKs of it at the right depths, NIP at the end
    assert t[:3] == [tok(1, 0), tok(0, 0), tok(2, BINARY_OPS["-"])]
    assert t[3] == tok(4, 0) and t[4] == tok(4, 1) and t[-1] == tok(5, 0)
    x = np.array([[0.25], [3.0], [1.0]])
    got = O.fused_cellwise([x], (1.0, 0.0), t, 3, 1, 0)
    o = 1.0 - x
    assert np.array_equal(got, o * (o > 0))
    # nested lets
    e2 = E.let(E.input(0) + E.const(0), lambda a: E.let(a * a, lambda b: b - a))
    got2 = O.fused_cellwise([x], (1.0,), e2.tokens, 3, 1, 0)
    assert np.array_equal(got2, (x + 1.0) * (x + 1.0) - (x + 1.0))


@pytest.mark.parametrize("agg", ["sum", "rowSums", "colSums", "min", "rowMins", "colMins", "max", "rowMaxs", "colMaxs"])
def test_oracle_program_evaluation_against_numpy(agg):
    rng = np.random.default_rng(3)
    rows, cols = 37, 11
    A, cv, rv = rng.uniform(-1, 1, (rows, cols)), rng.uniform(0.5, 2, (rows, 1)), rng.uniform(-1, 1, (1, cols))
    e = (E.input(0) / E.input(1) + E.input(2)) * E.const(0)
    full = (A / cv + rv) * 3.0
    got = O.fused_cellwise([A, cv, rv], (3.0,), e.tokens, rows, cols, FUSED_AGG[agg]).reshape(-1)
    fn = {"sum": np.sum, "min": np.min, "max": np.max}[agg.replace("rowS", "s").replace("colS", "s").replace("rowM", "m").replace("colM", "m")
                                                         .replace("sums", "sum").replace("mins", "min").replace("maxs", "max")]
    axis = None if agg in ("sum", "min", "max") else (1 if agg.startswith("row") else 0)
    exp = np.atleast_1d(fn(full, axis=axis)).reshape(-1)
    assert np.allclose(got, exp, rtol=1e-13, atol=1e-13)
    if "sum" not in agg.lower():
        assert np.array_equal(got, exp)  # min / max are exact


def test_csr_generator_invariants():
    import sys, os
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
    from csr_gen import poisson_rows_csr
    rp, ci, va = poisson_rows_csr(5000, 300, 10, 1)
    assert rp[0] == 0 and rp[-1] == len(ci) == len(va)
    for r in range(0, 5000, 97):
        seg = ci[rp[r]:rp[r + 1]]
        assert np.all(np.diff(seg) > 0) and (len(seg) == 0 or (seg[0] >= 0 and seg[-1] < 300))
    assert abs(np.diff(rp).mean() - 10) < 0.3
