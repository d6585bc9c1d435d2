"""ctypes loader for the CPU oracle (oracle/sysml_oracle.cpp).

TEST INFRASTRUCTURE: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs import this module.  It is the checker, never the product.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libsysml_oracle.so")

i64 = C.c_int64
dp = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")
ip = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")
lp = np.ctypeslib.ndpointer(dtype=np.int64, flags="C_CONTIGUOUS")

XtXv, XtwXv, XtXvy = 0, 1, 2
# aggregate opcodes / directions (same numbering as include/sysml_b200.h)
SUM, SUMSQ, MAX, MIN, MAXINDEX, MININDEX, MEAN, PROD, VAR = range(9)
ALL, ROW, COL = 0, 1, 2


def build(force=False):
    src = os.path.join(_HERE, "sysml_oracle.cpp")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "all"], stdout=subprocess.DEVNULL)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = C.CDLL(_SO)
        L.orc_num_threads.restype = C.c_int
        L.orc_csr_scalar_op.restype = i64
        L.orc_csr_scalar_op.argtypes = [lp, C.c_void_p, C.c_void_p, i64, C.c_double, C.c_int, C.c_int, lp, C.c_void_p, C.c_void_p]
        L.orc_csr_dense_op.restype = i64
        L.orc_csr_dense_op.argtypes = [lp, C.c_void_p, C.c_void_p, i64, i64, dp, C.c_int, C.c_int, lp, C.c_void_p, C.c_void_p]
        L.orc_matmult_ultra_sparse_left.restype = i64
        L.orc_matmult_ultra_sparse_left.argtypes = [lp, C.c_void_p, C.c_void_p, dp, i64, i64, lp, C.c_void_p, C.c_void_p]
        L.orc_csr_csr_op.restype = i64
        L.orc_csr_csr_op.argtypes = [lp, C.c_void_p, C.c_void_p, lp, C.c_void_p, C.c_void_p, i64, C.c_int, lp, C.c_void_p, C.c_void_p]
        L.orc_set_num_threads.argtypes = [C.c_int]
        L.orc_java_test_matrix.argtypes = [C.c_int, C.c_int, C.c_double, C.c_double, C.c_double, i64, dp]
        L.orc_java_next_doubles.argtypes = [i64, C.c_int, dp]
        L.orc_balanced_blocks.argtypes = [C.c_int, C.c_int, C.c_int, ip]
        L.orc_balanced_blocks.restype = C.c_int
        L.orc_mmchain_dense.argtypes = [dp, dp, C.c_void_p, dp, i64, i64, C.c_int, C.c_int]
        L.orc_mmchain_csr.argtypes = [lp, ip, dp, dp, C.c_void_p, dp, i64, i64, C.c_int, C.c_int]
        L.orc_tsmm_dense.argtypes = [dp, dp, i64, i64, C.c_int, C.c_int]
        L.orc_tsmm_dense.restype = i64
        L.orc_tsmm_csr_left.argtypes = [lp, ip, dp, dp, i64, i64, C.c_int]
        L.orc_tsmm_csr_left.restype = i64
        L.orc_matmult_dense.argtypes = [dp, dp, dp, i64, i64, i64, C.c_int]
        L.orc_matmult_dense.restype = i64
        L.orc_matmult_csr_dense.argtypes = [lp, ip, dp, dp, dp, i64, i64, i64, C.c_int]
        L.orc_matmult_csr_dense.restype = i64
        L.orc_agg_unary_dense.argtypes = [dp, i64, i64, C.c_int, C.c_int, dp, C.c_void_p, C.c_int]
        L.orc_agg_unary_csr.argtypes = [lp, ip, dp, i64, i64, C.c_int, C.c_int, dp, C.c_void_p, C.c_int]
        L.orc_binary_scalar.argtypes = [C.c_double, C.c_double, C.c_int]
        L.orc_binary_scalar.restype = C.c_double
        L.orc_binary_mm.argtypes = [dp, dp, dp, i64, i64, C.c_int, C.c_int]
        L.orc_binary_ms.argtypes = [dp, C.c_double, dp, i64, C.c_int, C.c_int]
        L.orc_axpy.argtypes = [dp, dp, C.c_double, dp, i64, i64, C.c_int]
        L.orc_unary_round.argtypes = [dp, dp, i64]
        L.orc_unary_signum.argtypes = [dp, dp, i64]
        L.orc_count_nnz.argtypes = [dp, i64]
        L.orc_count_nnz.restype = i64
        L.orc_dense_to_csr.argtypes = [dp, i64, i64, lp, C.c_void_p, C.c_void_p]
        L.orc_dense_to_csr.restype = i64
        L.orc_rand_uniform.argtypes = [dp, i64, C.c_uint64, i64, C.c_double, C.c_double]
        L.orc_rand_csr.argtypes = [i64, i64, C.c_double, C.c_uint64, i64, lp, C.c_void_p, C.c_void_p]
        L.orc_rand_csr.restype = i64
        _lib = L
    return _lib


def num_threads():
    return lib().orc_num_threads()


def host_cpus():
    """(logical CPUs this process may run on, distinct physical cores among them).  Independent of
    OMP_NUM_THREADS, which torchrun sets to 1 for every rank."""
    cpus = sorted(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else list(range(os.cpu_count() or 1))
    cores = set()
    for c in cpus:
        try:
            base = "/sys/devices/system/cpu/cpu%d/topology/" % c
            cores.add((open(base + "physical_package_id").read().strip(), open(base + "core_id").read().strip()))
        except OSError:
            cores.add(("?", str(c)))
    return len(cpus), len(cores)


def use_all_host_threads():
    """Sets the OpenMP team size to every CPU in this process's affinity mask and returns it."""
    k, _ = host_cpus()
    lib().orc_set_num_threads(k)
    return k


def _vp(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def java_test_matrix(rows, cols, lo, hi, sparsity, seed):
    """TestUtils.generateTestMatrix (src/test/java/org/apache/sysml/test/utils/TestUtils.java:1285-1297)."""
    out = np.zeros((rows, cols), dtype=np.float64)
    lib().orc_java_test_matrix(rows, cols, lo, hi, sparsity, seed, out)
    return out


def java_next_doubles(seed, n):
    out = np.zeros(n, dtype=np.float64)
    lib().orc_java_next_doubles(seed, n, out)
    return out


def balanced_blocks(length, k, const_k):
    out = np.zeros(max(8 * k + k, 1), dtype=np.int32)
    n = lib().orc_balanced_blocks(length, k, int(const_k), out)
    return out[:n].tolist()


def rand_uniform(length, seed, offset=0, lo=-1.0, hi=1.0):
    out = np.empty(length, dtype=np.float64)
    lib().orc_rand_uniform(out, length, seed, offset, lo, hi)
    return out


def rand_csr(rows, cols, mean, seed, row_offset=0):
    """CPU twin of sysml_b200_rand_csr: (rowptr int64, colind int32, val float64), bit-identical to the device generator."""
    rp = np.zeros(rows + 1, dtype=np.int64)
    nnz = lib().orc_rand_csr(rows, cols, float(mean), seed, row_offset, rp, None, None)
    ci = np.empty(max(nnz, 1), dtype=np.int32)
    va = np.empty(max(nnz, 1), dtype=np.float64)
    lib().orc_rand_csr(rows, cols, float(mean), seed, row_offset, rp, ci.ctypes.data_as(C.c_void_p), va.ctypes.data_as(C.c_void_p))
    return rp, ci[:nnz], va[:nnz]


def mmchain_dense(X, v, w, ct, k=1):
    n, m = X.shape
    out = np.zeros(m, dtype=np.float64)
    v = np.ascontiguousarray(v.reshape(-1))
    w = None if w is None else np.ascontiguousarray(w.reshape(-1))
    lib().orc_mmchain_dense(X, v, _vp(w), out, n, m, ct, k)
    return out


def mmchain_csr(rp, ci, va, shape, v, w, ct, k=1):
    n, m = shape
    out = np.zeros(m, dtype=np.float64)
    v = np.ascontiguousarray(v.reshape(-1))
    w = None if w is None else np.ascontiguousarray(w.reshape(-1))
    lib().orc_mmchain_csr(rp, ci, va, v, _vp(w), out, n, m, ct, k)
    return out


def tsmm_dense(X, left=True, k=1):
    rows, cols = X.shape
    dim = cols if left else rows
    out = np.zeros((dim, dim), dtype=np.float64)
    nnz = lib().orc_tsmm_dense(X, out, rows, cols, int(left), k)
    return out, nnz


def tsmm_csr_left(rp, ci, va, shape, k=1):
    rows, cols = shape
    out = np.zeros((cols, cols), dtype=np.float64)
    nnz = lib().orc_tsmm_csr_left(rp, ci, va, out, rows, cols, k)
    return out, nnz


def matmult_dense(A, B, k=1):
    m, cd = A.shape
    cd2, n = B.shape
    assert cd == cd2
    Cm = np.zeros((m, n), dtype=np.float64)
    nnz = lib().orc_matmult_dense(A, B, Cm, m, cd, n, k)
    return Cm, nnz


def matmult_csr_dense(rp, ci, va, shape, B, k=1):
    m, cd = shape
    cd2, n = B.shape
    assert cd == cd2
    Cm = np.zeros((m, n), dtype=np.float64)
    nnz = lib().orc_matmult_csr_dense(rp, ci, va, B, Cm, m, cd, n, k)
    return Cm, nnz


def _outlen(rows, cols, d):
    return 1 if d == ALL else (rows if d == ROW else cols)


def agg_unary_dense(X, op, d, k=1, with_corr=False):
    rows, cols = X.shape
    out = np.zeros(_outlen(rows, cols, d), dtype=np.float64)
    corr = np.zeros_like(out)
    lib().orc_agg_unary_dense(X, rows, cols, op, d, out, _vp(corr), k)
    return (out, corr) if with_corr else out


def agg_unary_csr(rp, ci, va, shape, op, d, k=1, with_corr=False):
    rows, cols = shape
    out = np.zeros(_outlen(rows, cols, d), dtype=np.float64)
    corr = np.zeros_like(out)
    lib().orc_agg_unary_csr(rp, ci, va, rows, cols, op, d, out, _vp(corr), k)
    return (out, corr) if with_corr else out


def binary_scalar(x, y, op):
    return lib().orc_binary_scalar(x, y, op)


def binary_mm(A, B, op):
    rows, cols = A.shape
    if B.shape == A.shape:
        bc = 0
    elif cols == 1 and B.shape[0] == 1 and B.shape[1] > 1:   # OUTER_VECTOR_VECTOR (LibMatrixBincell.java:150-163)
        out = np.empty((rows, B.shape[1]), dtype=np.float64)
        lib().orc_binary_mm(np.ascontiguousarray(A), np.ascontiguousarray(B), out, rows, B.shape[1], 3, op)
        return out
    elif B.shape == (rows, 1):
        bc = 1
    elif B.shape == (1, cols):
        bc = 2
    else:
        raise ValueError("shape mismatch")
    out = np.empty_like(A)
    lib().orc_binary_mm(A, np.ascontiguousarray(B), out, rows, cols, bc, op)
    return out


def binary_ms(A, s, op, scalar_left=False):
    out = np.empty_like(A)
    lib().orc_binary_ms(A, float(s), out, A.size, op, int(scalar_left))
    return out


def axpy(A, B, alpha):
    rows, cols = A.shape
    bc = 0 if B.shape == A.shape else (1 if B.shape == (rows, 1) else 2)
    out = np.empty_like(A)
    lib().orc_axpy(A, np.ascontiguousarray(B), float(alpha), out, rows, cols, bc)
    return out


def java_round(x):
    """Builtin ROUND = Math.round(double) as a double (functionobjects/Builtin.java:314): ties towards +infinity, exact,
    NaN -> 0, saturating at +-2^63."""
    a = np.ascontiguousarray(np.asarray(x, dtype=np.float64))
    out = np.empty_like(a)
    lib().orc_unary_round(a.reshape(-1), out.reshape(-1), a.size)
    return out if out.ndim else float(out)


def java_signum(x):
    """Builtin SIGN = Math.signum (functionobjects/Builtin.java:311): zeros (either sign) and NaN pass through."""
    a = np.ascontiguousarray(np.asarray(x, dtype=np.float64))
    out = np.empty_like(a)
    lib().orc_unary_signum(a.reshape(-1), out.reshape(-1), a.size)
    return out if out.ndim else float(out)


_UNARY_NP = {
    0: np.exp, 1: np.sqrt, 2: java_round, 3: np.abs, 4: np.log, 5: np.floor, 6: np.ceil, 7: java_signum,
    8: lambda x: 1.0 / (1.0 + np.exp(-x)), 9: np.sin, 10: np.cos, 11: np.tan, 12: np.tanh,
}


def fused_cellwise(inputs, consts, tokens, rows, cols, agg=0):
    """The UNFUSED reference evaluation of a postfix cellwise program (tokens as SYSML_FUSED_TOKEN): one
    LibMatrixBincell op per binary token (bincellOp, LibMatrixBincell.java:87-339), numpy for the unary builtins,
    then LibMatrixAgg sum / rowSums / colSums (agg 1/2/3), min (4/5/6) or max (7/8/9).  Operands broadcast like the reference: rows x 1 and
    1 x cols vectors against rows x cols."""
    stack = []

    def full(x):
        if np.isscalar(x):
            return np.full((rows, cols), float(x))
        return np.ascontiguousarray(np.broadcast_to(x, (rows, cols)))

    with np.errstate(all="ignore"):
        for t in tokens:
            kind, arg = (t >> 24) & 0xFF, t & 0xFFFFFF
            if kind == 0:
                stack.append(np.asarray(inputs[arg], dtype=np.float64))
            elif kind == 1:
                stack.append(float(consts[arg]))
            elif kind == 2:
                y = stack.pop()
                x = stack.pop()
                if np.isscalar(x) and np.isscalar(y):
                    stack.append(binary_scalar(x, y, arg))
                elif np.isscalar(y):
                    stack.append(binary_ms(np.ascontiguousarray(x), y, arg, False))
                elif np.isscalar(x):
                    stack.append(binary_ms(np.ascontiguousarray(y), x, arg, True))
                elif x.shape == y.shape or y.shape in ((x.shape[0], 1), (1, x.shape[1])):
                    stack.append(binary_mm(np.ascontiguousarray(x), y, arg))
                else:  # vector on the left or two different vectors: materialise both (outer / reversed broadcast)
                    stack.append(binary_mm(full(x), full(y), arg))
            elif kind == 3:
                x = stack.pop()
                stack.append(_UNARY_NP[arg](np.asarray(x, dtype=np.float64)) if not np.isscalar(x) else float(_UNARY_NP[arg](np.float64(x))))
            elif kind == 4:
                stack.append(stack[-1 - arg])
            elif kind == 5:
                del stack[-2]
            else:
                raise ValueError("bad token")
    assert len(stack) == 1
    res = full(stack[0])
    if agg == 0:
        return res
    op = SUM if agg <= 3 else (MIN if agg <= 6 else MAX)
    return agg_unary_dense(res, op, {1: ALL, 2: ROW, 3: COL}[(agg - 1) % 3 + 1], 1)


def count_nnz(A):
    return lib().orc_count_nnz(np.ascontiguousarray(A).reshape(-1), A.size)


def _csr_two_call(fn, rows, *args):
    crp = np.zeros(rows + 1, dtype=np.int64)
    nnz = fn(*args, crp, None, None)
    ci = np.zeros(max(nnz, 1), dtype=np.int32)
    va = np.zeros(max(nnz, 1), dtype=np.float64)
    fn(*args, crp, _vp(ci), _vp(va))
    return crp, ci[:nnz].copy(), va[:nnz].copy()


def csr_scalar_op(rp, ci, va, rows, s, op, scalar_left=False):
    """LibMatrixBincell.safeBinaryScalar SPARSE <- SPARSE (LibMatrixBincell.java:952-995)."""
    return _csr_two_call(lib().orc_csr_scalar_op, rows, rp, _vp(ci), _vp(va), rows, float(s), op, 1 if scalar_left else 0)


def csr_unary_op(rp, ci, va, rows, un_op):
    """MatrixBlock.sparseUnaryOperations SPARSE <- SPARSE (MatrixBlock.java:2640-2672)."""
    return _csr_two_call(lib().orc_csr_scalar_op, rows, rp, _vp(ci), _vp(va), rows, 0.0, un_op, 2)


def csr_dense_op(rp, ci, va, rows, cols, B, op):
    """safeBinaryMVSparse (LibMatrixBincell.java:338-440) for a vector B, safeBinaryMMSparseDenseSkip (:765-789) for a matrix."""
    bc = 0 if B.shape == (rows, cols) else (1 if B.shape == (rows, 1) else 2)
    return _csr_two_call(lib().orc_csr_dense_op, rows, rp, _vp(ci), _vp(va), rows, cols, np.ascontiguousarray(B), bc, op)


def matmult_ultra_sparse_left(a, B):
    """LibMatrixMult.matrixMultUltraSparse with an ultra-sparse left operand a = (rp, ci, va) and a dense B: CSR result."""
    rows, n = len(a[0]) - 1, B.shape[1]
    return _csr_two_call(lib().orc_matmult_ultra_sparse_left, rows, a[0], _vp(a[1]), _vp(a[2]), np.ascontiguousarray(B), rows, n)


def csr_csr_op(a, b, rows, op):
    """safeBinaryMMSparseSparse / mergeForSparseBinary (LibMatrixBincell.java:597-660, 1279-1340). a, b = (rp, ci, va)."""
    return _csr_two_call(lib().orc_csr_csr_op, rows, a[0], _vp(a[1]), _vp(a[2]), b[0], _vp(b[1]), _vp(b[2]), rows, op)


def dense_to_csr(A):
    rows, cols = A.shape
    rp = np.zeros(rows + 1, dtype=np.int64)
    nnz = lib().orc_dense_to_csr(A, rows, cols, rp, None, None)
    ci = np.zeros(max(nnz, 1), dtype=np.int32)
    va = np.zeros(max(nnz, 1), dtype=np.float64)
    lib().orc_dense_to_csr(A, rows, cols, rp, _vp(ci), _vp(va))
    return rp, ci[:nnz].copy() if nnz else ci[:0].copy(), va[:nnz].copy() if nnz else va[:0].copy()
