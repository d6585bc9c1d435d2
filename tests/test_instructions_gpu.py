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


def test_matmult_all_transpose_flags_for_every_sparsity_pair(ec):
    """LibMatrixCuMatMult.matmult handles the four (leftT, rightT) combinations for dense/sparse operands alike
    (LibMatrixCuMatMult.java:139-229); so must ba+*.  Also: the output carries its exact nnz (LibMatrixMult.java:112)."""
    rng = {"A": O.java_test_matrix(120, 90, -1, 1, 0.08, 11), "At": O.java_test_matrix(90, 120, -1, 1, 0.08, 12),
           "B": O.java_test_matrix(90, 70, -1, 1, 0.1, 13), "Bt": O.java_test_matrix(70, 90, -1, 1, 0.1, 14)}
    for a_sparse in (False, True):
        for b_sparse in (False, True):
            for lt in (False, True):
                for rt in (False, True):
                    A, B = rng["At" if lt else "A"], rng["Bt" if rt else "B"]
                    for nm, M, sp in (("A", A, a_sparse), ("B", B, b_sparse)):
                        if sp:
                            ec.setMatrixCSR(nm, *O.dense_to_csr(M), M.shape)
                        else:
                            ec.setMatrix(nm, M)
                    ec.execute(inst("ba+*", mat("A"), mat("B"), mat("C"), str(lt).lower(), str(rt).lower()))
                    exp = (A.T if lt else A).astype(np.longdouble) @ (B.T if rt else B).astype(np.longdouble)
                    got = ec.getMatrix("C")
                    what = "ba+* sparse=(%s,%s) lt=%s rt=%s" % (a_sparse, b_sparse, lt, rt)
                    assert got.shape == (120, 70), what
                    assert np.max(np.abs(got - exp.astype(np.float64))) < 1e-13, what
                    assert ec.getMatrixDims("C")[2] == int(np.count_nonzero(got)), what


def test_sparse_operands_of_cellwise_unary_axpy_scalar_are_densified(ec):
    """LibMatrixCUDA.getDensePointer converts a CSR operand with sparseToDense and proceeds
    (LibMatrixCUDA.java:655-664, 1382-1383); the results equal the dense-input results bit for bit."""
    X = O.java_test_matrix(300, 40, -1, 1, 0.1, 21)
    Y = O.java_test_matrix(300, 40, -1, 1, 0.2, 22)
    csr = lambda M: O.dense_to_csr(M) + (M.shape,)

    def fresh():
        ec.setMatrixCSR("X", *csr(X))
        ec.setMatrixCSR("Y", *csr(Y))
    fresh()
    ec.execute(inst("+", mat("X"), mat("Y"), mat("Z")))
    assert np.array_equal(ec.getMatrix("Z"), X + Y)
    fresh()
    ec.execute(inst("*", mat("X"), literal(3.0), mat("Z")))
    assert np.array_equal(ec.getMatrix("Z"), X * 3.0)
    fresh()
    ec.execute(inst("+*", mat("X"), literal(0.5), mat("Y"), mat("Z")))
    assert np.array_equal(ec.getMatrix("Z"), O.axpy(X, Y, 0.5))
    fresh()
    ec.execute(inst("abs", mat("X"), mat("Z")))
    assert np.array_equal(ec.getMatrix("Z"), np.abs(X))
    fresh()
    ec.execute(inst("r'", mat("X"), mat("Z")))
    assert np.array_equal(ec.getMatrix("Z"), X.T)
    assert np.array_equal(ec.getMatrix("X"), X), "the converted input still holds the same values"
    # mmchain with sparse v / w
    v = O.java_test_matrix(40, 1, 0, 1, 0.5, 23)
    w = O.java_test_matrix(300, 1, 0, 1, 0.5, 24)
    ec.setMatrix("Xd", X)
    ec.setMatrixCSR("v", *csr(v))
    ec.setMatrixCSR("w", *csr(w))
    ec.execute(inst("mmchain", mat("Xd"), mat("v"), mat("w"), mat("q"), "XtwXv"))
    assert_equal_matrices(O.mmchain_dense(X, v, w, 1, 1), ec.getMatrix("q"))


def test_exam_sparsity_of_a_matmult_result(ec):
    """MatrixBlock.examSparsity after matrixMult (LibMatrixMult.java:113): a result with few non-zeros becomes CSR."""
    A = np.zeros((200, 300))
    A[5, 7], A[17, 3], A[199, 299] = 2.0, -1.5, 4.0
    B = np.eye(300)
    ec.setMatrix("A", A)
    ec.setMatrix("B", B)
    ec.execute(inst("ba+*", mat("A"), mat("B"), mat("C"), "false", "false"))
    assert ec.getMatrixDims("C") == (200, 300, 3)
    assert ec.examSparsity("C") is True
    assert np.array_equal(ec.getMatrix("C"), A)
    D = O.rand_uniform(200 * 300, seed=5).reshape(200, 300)
    ec.setMatrix("D", D)
    ec.execute(inst("ba+*", mat("D"), mat("B"), mat("E"), "false", "false"))
    assert ec.examSparsity("E") is False and ec.getMatrixDims("E")[2] == 60000


@pytest.mark.parametrize("policy,order,victim", [("lru", "ABS", "A"), ("mru", "ASB", "B"), ("lfu", "AASSB", "B"), ("min_evict", "ABS", "S")])
def test_eviction_policies(policy, order, victim):
    """sysml.gpu.eviction.policy (DMLScript.EvictionPolicy; GPUComparator, GPUMemoryManager.java:531-571; timestamps per
    GPUObject.updateReleaseLocks, GPUObject.java:650-667): which unlocked matrix leaves the device when the next one does
    not fit.  A, B: 8 MB, S: 4 MB, budget 22 MB, request 4 MB.  LRU: the oldest release; MRU: the latest; LFU: the fewest
    releases; MIN_EVICT: the smallest block that is at least as large as the request."""
    big, small = (1000, 1000), (500, 1000)
    with ExecutionContext(0, memory_limit=22 << 20) as ec:
        ec.setEvictionPolicy(policy)
        ec.setMatrix("A", np.full(big, 1.0))
        ec.setMatrix("B", np.full(big, 2.0))
        ec.setMatrix("S", np.full(small, 3.0))
        ec.setMatrix("N", np.full(small, 4.0))
        for nm in order:
            ec.execute(inst("uak+", mat(nm), scalar_var("s")))
        assert ec.stats()["evictions"] == 0
        ec.execute(inst("uak+", mat("N"), scalar_var("s")))   # needs 4 MB: exactly one eviction
        assert ec.stats()["evictions"] == 1
        for nm in [x for x in "ABS" if x != victim]:          # the survivors are still resident: no upload
            h0 = ec.stats()["h2d"]
            ec.execute(inst("uamax", mat(nm), scalar_var("t")))
            assert ec.stats()["h2d"] == h0, "%s should still be resident under %s" % (nm, policy)
        h0 = ec.stats()["h2d"]
        ec.execute(inst("uamax", mat(victim), scalar_var("t")))
        assert ec.stats()["h2d"] == h0 + 1, "%s should have been evicted under %s" % (victim, policy)
        assert ec.getScalar("t") == {"A": 1.0, "B": 2.0, "S": 3.0}[victim]
    with ExecutionContext(0) as e3:
        with pytest.raises(native.NativeError):
            e3.setEvictionPolicy("random")


def test_csr_upload_under_memory_pressure_does_not_evict_itself():
    """A CSR upload takes three allocations; under pressure the second or third one evicts, and the half-built matrix
    must not be the victim (the reference publishes the CSRPointer only when complete, CSRPointer.java:271-287)."""
    X = O.java_test_matrix(4000, 500, -1, 1, 0.25, 31)     # nnz ~ 500k: val 4 MB, colind 2 MB, rowptr 16 KB
    rp, ci, va = O.dense_to_csr(X)
    v = O.rand_uniform(500, seed=2).reshape(500, 1)
    with ExecutionContext(0, memory_limit=11 << 20) as ec:
        ec.setMatrix("F", np.ones((1000, 1000)))           # 8 MB resident and unlocked
        ec.execute(inst("uak+", mat("F"), scalar_var("s")))
        ec.setMatrixCSR("X", rp, ci, va, X.shape)
        ec.setMatrix("v", v)
        ec.execute(inst("ba+*", mat("X"), mat("v"), mat("Xv"), "false", "false"))
        assert ec.stats()["evictions"] >= 1
        assert_equal_matrices(O.matmult_csr_dense(rp, ci, va, X.shape, v, 1)[0], ec.getMatrix("Xv"))
        assert np.array_equal(ec.getMatrix("X"), X)


def test_outer_vector_vector_instruction(ec):
    """`GPU°<°cv°rv°out` with an n x 1 and a 1 x m operand produces the n x m outer comparison (Kmeans-predict / table-style
    idioms; BinaryAccessType.OUTER_VECTOR_VECTOR, LibMatrixBincell.java:150-163)."""
    cv = O.java_test_matrix(200, 1, -1, 1, 1.0, 1)
    rv = O.java_test_matrix(1, 70, -1, 1, 1.0, 2)
    ec.setMatrix("cv", cv)
    ec.setMatrix("rv", rv)
    ec.execute(inst("<", mat("cv"), mat("rv"), mat("Z")))
    assert ec.getMatrixDims("Z")[:2] == (200, 70)
    assert np.array_equal(ec.getMatrix("Z"), (cv < rv).astype(np.float64))
    ec.execute(inst("*", mat("cv"), mat("rv"), mat("Z")))
    assert np.array_equal(ec.getMatrix("Z"), cv * rv)


def test_scratch_is_a_block_of_the_memory_manager():
    """The kernels' scratch is allocated through GPUMemoryManager like every matrix (so `memory_limit` covers it): an
    instruction whose scratch request exceeds the block fails without side effects (SYSML_E_WORKSPACE), the block is
    re-allocated at the requested size and the instruction runs again."""
    X = O.rand_uniform(30_000 * 600, seed=3).reshape(30_000, 600)
    with ExecutionContext(0, memory_limit=400 << 20) as ec:
        base = ec.stats()["allocated_bytes"]
        assert base >= 1 << 20, "the initial scratch block is on the manager's books"
        ec.setMatrix("X", X)
        ec.execute(inst("tsmm", mat("X"), mat("T"), "LEFT"))            # split-K partials: tens of MB of scratch
        T, nnz = O.tsmm_dense(X, True, 8)
        assert_equal_matrices(T, ec.getMatrix("T"))
        st = ec.stats()
        assert st["allocated_bytes"] > base + X.nbytes, "scratch grew through the manager"
        assert st["instructions"] == 1, "the repeated attempt is not counted twice"
        assert st["allocated_bytes"] <= 400 << 20


def test_caller_owned_workspace_through_the_abi(gctx):
    """sysml_b200_ctx_set_workspace: with a caller-owned block the library never allocates scratch; a request that does not
    fit returns SYSML_E_WORKSPACE and ctx_workspace_needed says how much the operation wanted."""
    import ctypes as C
    from systemml_b200 import GPUContext
    with GPUContext(0) as g:
        n, m = 50_000, 1000
        X, v, out, ref = g.rand(n, m, seed=1), g.rand(m, 1, seed=2), g.empty(m, 1), g.empty(m, 1)
        g.call("sysml_b200_mmchain_dense", C.c_void_p(X.ptr), C.c_void_p(v.ptr), None, C.c_void_p(ref.ptr), n, m, 0)
        small = g.allocate(4096)
        g.call("sysml_b200_ctx_set_workspace", C.c_void_p(small), 4096)
        free0, _ = g.getAvailableMemory()
        rc = g._lib.sysml_b200_mmchain_dense(g.handle, C.c_void_p(X.ptr), C.c_void_p(v.ptr), None, C.c_void_p(out.ptr), n, m, 0)
        assert rc == native.E_WORKSPACE
        need = g._lib.sysml_b200_ctx_workspace_needed(g.handle)
        assert need > 4096
        assert g.getAvailableMemory()[0] == free0, "nothing was allocated behind the caller's back"
        big = g.allocate(need)
        g.call("sysml_b200_ctx_set_workspace", C.c_void_p(big), need)
        g.call("sysml_b200_mmchain_dense", C.c_void_p(X.ptr), C.c_void_p(v.ptr), None, C.c_void_p(out.ptr), n, m, 0)
        assert np.array_equal(out.to_host(), ref.to_host())
        assert g._lib.sysml_b200_ctx_workspace_bytes(g.handle) == need
        g.call("sysml_b200_ctx_set_workspace", None, 0)                  # back to the library-owned default
        g.call("sysml_b200_mmchain_dense", C.c_void_p(X.ptr), C.c_void_p(v.ptr), None, C.c_void_p(out.ptr), n, m, 0)
        assert np.array_equal(out.to_host(), ref.to_host())
        for p in (small, big):
            g.cudaFreeHelper(p)
        for o in (X, v, out, ref):
            o.free()


def test_sparse_operands_stay_sparse_under_sparse_safe_cellwise_instructions(ec):
    """Cellwise instructions on a CSR operand: where the CPU back-end's rules call the operator sparse-safe (ScalarOperator /
    BinaryOperator / UnaryOperator constructors, isSparseSafeDivide, estimateSparsityOnBinary) the device block stays CSR --
    result cells, nnz and the format are checked -- and everywhere else the operand is converted to dense as
    LibMatrixCUDA does, with the same cells as the dense operator."""
    rows, cols = 4000, 300
    A = O.java_test_matrix(rows, cols, -2, 2, 0.02, 61)
    B = O.java_test_matrix(rows, cols, -2, 2, 0.03, 62)
    cv = O.java_test_matrix(rows, 1, -1, 1, 0.6, 63)
    rv = O.java_test_matrix(1, cols, 0.5, 1.5, 1.0, 64)
    D = O.java_test_matrix(rows, cols, 0.5, 2, 1.0, 65)

    def put():
        for n, M in (("A", A), ("B", B)):
            rp, ci, va = O.dense_to_csr(M)
            ec.setMatrixCSR(n, rp, ci, va, M.shape)
        ec.setMatrix("cv", cv)
        ec.setMatrix("rv", rv)
        ec.setMatrix("D", D)

    put()
    sparse_cases = [
        (inst("*", mat("A"), literal(0.5), mat("Z")), O.binary_ms(A, 0.5, 2)),
        (inst("*", literal(-3), mat("A"), mat("Z")), O.binary_ms(A, -3.0, 2, True)),
        (inst("/", mat("A"), literal(4), mat("Z")), O.binary_ms(A, 4.0, 3)),
        (inst("^2", mat("A"), literal(2), mat("Z")), O.binary_ms(A, 2.0, 4)),
        (inst(">", mat("A"), literal(0), mat("Z")), O.binary_ms(A, 0.0, 7)),
        (inst("!=", mat("A"), literal(0), mat("Z")), O.binary_ms(A, 0.0, 10)),
        (inst("abs", mat("A"), mat("Z")), np.abs(A)),
        (inst("sign", mat("A"), mat("Z")), O.java_signum(A)),
        (inst("round", mat("A"), mat("Z")), O.java_round(A)),
        (inst("*", mat("A"), mat("B"), mat("Z")), O.binary_mm(A, B, 2)),
        (inst("+", mat("A"), mat("B"), mat("Z")), O.binary_mm(A, B, 0)),
        (inst("-", mat("A"), mat("B"), mat("Z")), O.binary_mm(A, B, 1)),
        (inst("*", mat("A"), mat("cv"), mat("Z")), O.binary_mm(A, cv, 2)),
        (inst("*", mat("A"), mat("rv"), mat("Z")), O.binary_mm(A, rv, 2)),
        (inst("/", mat("A"), mat("rv"), mat("Z")), O.binary_mm(A, rv, 3)),
        (inst("*", mat("A"), mat("D"), mat("Z")), O.binary_mm(A, D, 2)),
        (inst("*", mat("D"), mat("A"), mat("Z")), O.binary_mm(D, A, 2)),
        (inst("/", mat("A"), mat("D"), mat("Z")), O.binary_mm(A, D, 3)),
    ]
    for instruction, exp in sparse_cases:
        ec.execute(instruction)
        assert ec.isSparseOnDevice("Z") is True, instruction
        assert ec.isSparseOnDevice("A") is True, instruction
        assert ec.getMatrixDims("Z") == (rows, cols, int(np.count_nonzero(exp))), instruction
        got = ec.getMatrix("Z")
        if instruction.split("°")[1] in ("^2",):
            assert_equal_matrices(exp + 0.0, got, rel=1e-13)
        else:
            assert np.array_equal(got, exp + 0.0), instruction
    dense_cases = [
        (inst("+", mat("A"), literal(1), mat("Z")), O.binary_ms(A, 1.0, 0)),          # 0 + 1
        (inst("^", mat("A"), literal(-1), mat("Z")), None),                           # 0 ^ -1 = Inf: dense here (see host_runtime.cpp)
        (inst("<", mat("A"), literal(0.5), mat("Z")), O.binary_ms(A, 0.5, 5)),
        (inst("exp", mat("A"), mat("Z")), None),
        (inst("+", mat("A"), mat("cv"), mat("Z")), O.binary_mm(A, cv, 0)),
        (inst("/", mat("A"), mat("cv"), mat("Z")), O.binary_mm(A, cv, 3)),            # cv holds zeros: 0/0
        (inst("==", mat("A"), mat("B"), mat("Z")), O.binary_mm(A, B, 9)),
        (inst("+", mat("A"), mat("D"), mat("Z")), O.binary_mm(A, D, 0)),              # the estimate says dense
    ]
    for instruction, exp in dense_cases:
        put()
        ec.execute(instruction)
        assert ec.isSparseOnDevice("Z") is False, instruction
        if exp is not None:
            assert np.array_equal(ec.getMatrix("Z"), exp, equal_nan=True), instruction
    # a sparse result feeds the next instruction like any other variable (CSR x dense matmult, aggregate)
    put()
    v = O.java_test_matrix(cols, 1, -1, 1, 1.0, 66)
    ec.setMatrix("v", v)
    ec.execute(inst("*", mat("A"), mat("cv"), mat("Z")))
    ec.execute(inst("ba+*", mat("Z"), mat("v"), mat("q"), "false", "false"))
    assert_equal_matrices(O.matmult_dense(A * cv, v, 4)[0], ec.getMatrix("q"))
    ec.execute(inst("uak+", mat("Z"), scalar_var("s")))
    assert abs(ec.getScalar("s") - O.agg_unary_dense(A * cv, O.SUM, O.ALL)[0]) < 1e-9


def test_ultra_sparse_matmult_instruction_gives_a_sparse_result(ec):
    """ba+* with an ultra-sparse left operand (LibMatrixMult.isUltraSparseMatrixMult, LibMatrixMult.java:3761-3777) keeps the
    result sparse on the device, as matrixMultUltraSparse does on the CPU; a merely sparse left operand goes through the
    general kernel into a dense block. Same cells either way."""
    m, k, n = 200_000, 400, 50
    P = np.zeros((m, k))
    P[[5, 77_000, 150_001], [3, 399, 17]] = [1.0, 1.0, -2.5]          # three cells in 80M: a row selection
    X = O.java_test_matrix(k, n, -1, 1, 0.9, 4)
    rp, ci, va = O.dense_to_csr(P)
    ec.setMatrixCSR("P", rp, ci, va, P.shape)
    ec.setMatrix("X", X)
    ec.execute(inst("ba+*", mat("P"), mat("X"), mat("Z"), "false", "false"))
    assert ec.isSparseOnDevice("Z") is True
    exp = P @ X
    assert ec.getMatrixDims("Z") == (m, n, int(np.count_nonzero(exp)))
    assert np.array_equal(ec.getMatrix("Z"), exp)
    S = O.java_test_matrix(2000, k, -1, 1, 0.05, 6)
    rp, ci, va = O.dense_to_csr(S)
    ec.setMatrixCSR("S", rp, ci, va, S.shape)
    ec.execute(inst("ba+*", mat("S"), mat("X"), mat("Z2"), "false", "false"))
    assert ec.isSparseOnDevice("Z2") is False
    assert_equal_matrices(O.matmult_dense(S, X, 4)[0], ec.getMatrix("Z2"))
