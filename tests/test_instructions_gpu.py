"""Instruction-level parity: SystemML GPU instruction strings executed by the C++ host runtime
(GPUInstructionParser / *GPUInstruction / ExecutionContext / GPUObject / GPUMemoryManager mirror) against
the CPU oracle -- the same shape as the reference's gpu/ tests (run on CPU, run on GPU, compare)."""
import numpy as np
import pytest

from oracle import oracle as O
from systemml_b200 import native
from systemml_b200.execution_context import ExecutionContext, inst, literal, mat, scalar_var
from tests.util import assert_equal_matrices

pytestmark = pytest.mark.gpu


@pytest.fixture()
def ec():
    e = ExecutionContext(0)
    yield e
    e.destroy()


def test_matmult_tsmm_mmchain_instructions(ec):
    X = O.java_test_matrix(3471, 567, 0, 1, 0.7, 7)
    v = O.java_test_matrix(567, 1, 0, 1, 0.7, 3)
    w = O.java_test_matrix(3471, 1, 0, 1, 0.7, 10)
    ec.setMatrix("X", X)
    ec.setMatrix("v", v)
    ec.setMatrix("w", w)
    ec.execute(inst("ba+*", mat("X"), mat("v"), mat("Xv"), "false", "false"))
    assert_equal_matrices(O.matmult_dense(X, v, 4)[0], ec.getMatrix("Xv"))
    ec.execute(inst("ba+*", mat("X"), mat("Xv"), mat("g"), "true", "false"))  # t(X) %*% (X v), transposes folded
    assert_equal_matrices(O.mmchain_dense(X, v, None, 0, 4), ec.getMatrix("g"))
    for ct, fields in (("XtXv", (mat("X"), mat("v"), mat("q"))), ("XtwXv", (mat("X"), mat("v"), mat("w"), mat("q"))),
                       ("XtXvy", (mat("X"), mat("v"), mat("w"), mat("q")))):
        ec.execute(inst("mmchain", *fields, ct, 16))
        assert ec.getMatrixDims("q")[:2] == (567, 1)
        assert_equal_matrices(O.mmchain_dense(X, v, None if ct == "XtXv" else w, O.lib and {"XtXv": 0, "XtwXv": 1, "XtXvy": 2}[ct], 4), ec.getMatrix("q"))
    ec.execute(inst("tsmm", mat("X"), mat("T"), "LEFT"))
    T, nnz = O.tsmm_dense(X, True, 8)
    assert_equal_matrices(T, ec.getMatrix("T"))
    assert ec.getMatrixDims("T") == (567, 567, nnz)
    st = ec.stats()
    assert st["instructions"] == 6
    assert st["h2d"] == 3, "X, v, w are uploaded once and stay resident (GPUObject cache)"


def test_aggregate_and_cellwise_instructions(ec):
    X = O.java_test_matrix(1000, 50, -1, 1, 0.9, 42)
    Y = O.java_test_matrix(1000, 50, -1, 1, 0.9, 43)
    cv = O.java_test_matrix(1000, 1, 0.5, 1.5, 1.0, 44)
    ec.setMatrix("X", X)
    ec.setMatrix("Y", Y)
    ec.setMatrix("cv", cv)
    ec.execute(inst("uak+", mat("X"), scalar_var("s")))
    assert abs(ec.getScalar("s") - O.agg_unary_dense(X, O.SUM, O.ALL)[0]) < 1e-10
    for opc, op, d in (("uark+", O.SUM, O.ROW), ("uack+", O.SUM, O.COL), ("uarmin", O.MIN, O.ROW), ("uarimax", O.MAXINDEX, O.ROW),
                       ("uacmax", O.MAX, O.COL), ("uarsqk+", O.SUMSQ, O.ROW)):
        ec.execute(inst(opc, mat("X"), mat("r")))
        exp = O.agg_unary_dense(X, op, d)
        got = ec.getMatrix("r")
        assert got.shape == ((1000, 1) if d == O.ROW else (1, 50))
        assert_equal_matrices(exp, got, what=opc)
    for opc, code in (("+", 0), ("*", 2), ("/", 3), ("<=", 6), ("max", 12), ("1-*", 15), ("%%", 17)):
        ec.execute(inst(opc, mat("X"), mat("Y"), mat("Z")))
        assert np.array_equal(O.binary_mm(X, Y, code), ec.getMatrix("Z"), equal_nan=True), opc
    ec.execute(inst("/", mat("X"), mat("cv"), mat("Z")))  # matrix / column vector (P / rowSums(P))
    assert np.array_equal(O.binary_mm(X, cv, 3), ec.getMatrix("Z"))
    ec.execute(inst("*", mat("X"), literal(-2.0), mat("Z")))
    assert np.array_equal(-2.0 * X, ec.getMatrix("Z"))
    ec.execute(inst("-", literal(1.0), mat("X"), mat("Z")))
    assert np.array_equal(1.0 - X, ec.getMatrix("Z"))
    ec.setScalar("alpha", 0.25)
    ec.execute(inst("+*", mat("X"), scalar_var("alpha"), mat("Y"), mat("Z")))
    assert np.array_equal(O.axpy(X, Y, 0.25), ec.getMatrix("Z"))
    ec.execute(inst("-*", mat("X"), literal(0.5), mat("Y"), mat("Z")))
    assert np.array_equal(O.axpy(X, Y, -0.5), ec.getMatrix("Z"))
    ec.execute(inst("exp", mat("X"), mat("E")))
    assert_equal_matrices(np.exp(X), ec.getMatrix("E"), rel=1e-14)
    ec.execute(inst("r'", mat("X"), mat("Xt")))
    assert np.array_equal(X.T, ec.getMatrix("Xt"))


def test_sparse_inputs(ec):
    X = O.java_test_matrix(2000, 300, 0, 1, 0.05, 7)
    v = O.java_test_matrix(300, 1, 0, 1, 0.7, 3)
    rp, ci, va = O.dense_to_csr(X)
    ec.setMatrixCSR("X", rp, ci, va, X.shape)
    ec.setMatrix("v", v)
    ec.execute(inst("ba+*", mat("X"), mat("v"), mat("Xv"), "false", "false"))
    assert_equal_matrices(O.matmult_csr_dense(rp, ci, va, X.shape, v, 1)[0], ec.getMatrix("Xv"))
    ec.execute(inst("mmchain", mat("X"), mat("v"), mat("q"), "XtXv"))
    assert_equal_matrices(O.mmchain_csr(rp, ci, va, X.shape, v, None, 0, 1), ec.getMatrix("q"))
    ec.execute(inst("uark+", mat("X"), mat("r")))
    assert_equal_matrices(O.agg_unary_csr(rp, ci, va, X.shape, O.SUM, O.ROW), ec.getMatrix("r"))
    assert np.array_equal(ec.getMatrix("X"), X)


def test_linregcg_iterations_as_instructions(ec):
    """The inner loop of scripts/algorithms/LinearRegCG.dml:174-200 written as the instructions the compiler
    emits (mmchain + dot products as ba+* + axpy), scalars on the host as in SystemML, vs a numpy/oracle run."""
    n, m, lam = 20_000, 200, 0.01
    X = O.rand_uniform(n * m, seed=0).reshape(n, m)
    y = X @ O.rand_uniform(m, seed=1) + 0.01 * O.rand_uniform(n, seed=2)
    ec.setMatrix("X", X)
    ec.setMatrix("y", y)
    ec.execute(inst("ba+*", mat("X"), mat("y"), mat("Xty"), "true", "false"))
    r = -ec.getMatrix("Xty")
    ec.setMatrix("r", r)
    ec.setMatrix("p", -r)
    ec.setMatrix("beta", np.zeros((m, 1)))
    ec.execute(inst("ba+*", mat("r"), mat("r"), mat("nr2m"), "true", "false"))
    nr2 = float(ec.getMatrix("nr2m")[0, 0])
    # oracle-side state
    r_o, p_o, b_o, nr2_o = -(X.T @ y).reshape(m, 1), (X.T @ y).reshape(m, 1), np.zeros((m, 1)), float((X.T @ y) @ (X.T @ y))
    for _ in range(5):
        ec.execute(inst("mmchain", mat("X"), mat("p"), mat("q"), "XtXv", 16))
        ec.execute(inst("+*", mat("q"), literal(lam), mat("p"), mat("q2")))
        ec.execute(inst("ba+*", mat("p"), mat("q2"), mat("pq"), "true", "false"))
        a = nr2 / float(ec.getMatrix("pq")[0, 0])
        ec.execute(inst("+*", mat("beta"), literal(a), mat("p"), mat("beta2")))
        ec.execute(inst("+*", mat("r"), literal(a), mat("q2"), mat("r2")))
        ec.execute(inst("ba+*", mat("r2"), mat("r2"), mat("nr2m"), "true", "false"))
        nr2n = float(ec.getMatrix("nr2m")[0, 0])
        ec.execute(inst("*", mat("p"), literal(nr2n / nr2), mat("bp")))
        ec.execute(inst("-", mat("bp"), mat("r2"), mat("p")))
        ec.execute(inst("+", mat("beta2"), literal(0.0), mat("beta")))
        ec.execute(inst("+", mat("r2"), literal(0.0), mat("r")))
        nr2 = nr2n
        q_o = O.mmchain_dense(X, p_o, None, 0, 4).reshape(m, 1) + lam * p_o
        a_o = nr2_o / float((p_o.T @ q_o)[0, 0])
        b_o = b_o + a_o * p_o
        r_o = r_o + a_o * q_o
        nr2n_o = float((r_o.T @ r_o)[0, 0])
        p_o = -r_o + (nr2n_o / nr2_o) * p_o
        nr2_o = nr2n_o
    assert_equal_matrices(b_o, ec.getMatrix("beta"), rel=1e-9, what="beta after 5 CG iterations")
    st = ec.stats()
    assert st["freelist_reuse"] > 0, "rmvar-ed pointers of equal size must be reused (GPUMemoryManager step 1)"


def test_memory_manager_free_list_and_eviction():
    # 3 matrices of 8 MB under a 20 MB budget: the third acquisition must evict the least recently used one
    with ExecutionContext(0, memory_limit=20 << 20) as ec:
        A = O.rand_uniform(1000 * 1000, seed=1).reshape(1000, 1000)
        for nm in ("A", "B", "C"):
            ec.setMatrix(nm, A + (ord(nm) - 65))
        ec.execute(inst("uak+", mat("A"), scalar_var("sa")))
        ec.execute(inst("uak+", mat("B"), scalar_var("sb")))
        assert ec.stats()["evictions"] == 0
        ec.execute(inst("uak+", mat("C"), scalar_var("sc")))
        st = ec.stats()
        assert st["evictions"] >= 1 and st["allocated_bytes"] <= 20 << 20
        assert abs(ec.getScalar("sc") - ec.getScalar("sa") - 2e6) < 1e-3
        # an evicted DIRTY matrix is written back to the host first
        ec.execute(inst("*", mat("C"), literal(2.0), mat("D")))     # D dirty on device (evicts A or B)
        ec.execute(inst("uak+", mat("A"), scalar_var("sa2")))        # forces more evictions
        ec.execute(inst("uak+", mat("B"), scalar_var("sb2")))
        assert np.array_equal(ec.getMatrix("D"), 2.0 * (A + 2))
        assert ec.getScalar("sa2") == ec.getScalar("sa")
        # a request larger than the budget fails with the reference's message
        with pytest.raises(native.NativeError) as ei:
            ec.setMatrix("big", np.zeros((2000, 2000)))
            ec.execute(inst("uak+", mat("big"), scalar_var("x")))
        assert ei.value.code == native.E_OOM and "not enough memory on device" in str(ei.value)
    with ExecutionContext(0, eager_cuda_free=True) as ec2:
        ec2.setMatrix("A", np.ones((100, 100)))
        ec2.execute(inst("*", mat("A"), literal(2.0), mat("B")))
        ec2.removeVariable("B")
        st = ec2.stats()
        assert st["cuda_free"] >= 1 and st["freelist_bytes"] == 0


def test_error_paths(ec):
    ec.setMatrix("A", np.ones((3, 4)))
    ec.setMatrix("B", np.ones((5, 2)))
    with pytest.raises(native.NativeError) as ei:
        ec.execute(inst("ba+*", mat("A"), mat("B"), mat("C"), "false", "false"))
    assert ei.value.code == native.E_DIM
    with pytest.raises(native.NativeError) as ei:
        ec.execute(inst("ba+*", mat("A"), mat("nope"), mat("C"), "false", "false"))
    assert "does not exist" in str(ei.value)
    with pytest.raises(native.NativeError):
        ec.execute(inst("+", mat("A"), mat("B"), mat("C")))
    # locks are released on failure: the inputs can still be used and removed
    ec.execute(inst("uak+", mat("A"), scalar_var("s")))
    assert ec.getScalar("s") == 12.0
    ec.removeVariable("A")


def test_borrowed_device_matrix(gctx, ec):
    """A matrix that already lives on the device (e.g. produced by another library) is bound, used and left alone."""
    X = gctx.rand(5000, 300, seed=3)
    Xh = X.to_host()
    ec.setMatrixDevice("X", X.ptr, 5000, 300)
    v = O.rand_uniform(300, seed=4).reshape(300, 1)
    ec.setMatrix("v", v)
    ec.execute(inst("mmchain", mat("X"), mat("v"), mat("q"), "XtXv"))
    assert_equal_matrices(O.mmchain_dense(Xh, v, None, 0, 4), ec.getMatrix("q"))
    assert np.array_equal(ec.getMatrix("X"), Xh)
    ec.removeVariable("X")                      # must NOT free the caller's block
    assert np.array_equal(X.to_host(), Xh)
    assert ec.stats()["h2d"] == 1               # only v was uploaded
    X.free()
