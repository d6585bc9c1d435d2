"""LibMatrixB200 -- host-side mirror of LibMatrixCUDA / LibMatrixCuMatMult over the C ABI.

Function names, argument order and error behaviour follow the reference's static methods
(J/runtime/matrix/data/LibMatrixCUDA.java, LibMatrixCuMatMult.java) so that a GPU instruction
class can call these in place of the JCuda-based ones:

  matmult                 <- LibMatrixCuMatMult.matmult            (LibMatrixCuMatMult.java:139-229)
  matmultTSMM             <- LibMatrixCUDA.matmultTSMM             (LibMatrixCUDA.java:477-524)
  mmchain                 <- (new) CP twin MatrixBlock.chainMatrixMultOperations (MatrixBlock.java:3491-3518)
  unaryAggregate          <- LibMatrixCUDA.unaryAggregate          (LibMatrixCUDA.java:573-894)
  matrixMatrixArithmetic  <- LibMatrixCUDA.matrixMatrixArithmetic  (LibMatrixCUDA.java:1245-1270)
  matrixScalarArithmetic  <- LibMatrixCUDA.matrixScalarArithmetic  (LibMatrixCUDA.java:1126-1195)
  axpy                    <- LibMatrixCUDA.axpy                    (LibMatrixCUDA.java:2295-2350)
  transpose               <- LibMatrixCUDA.transpose               (LibMatrixCUDA.java:1723-1729)

Outputs are freshly allocated GPUObjects (getDenseMatrixOutputForGPUInstruction, LibMatrixCUDA.java:2474);
scalar aggregates come back as Python floats (ec.setScalarOutput, LibMatrixCUDA.java:692-693).
"""
import ctypes as C

from . import native
from .gpu_context import GPUObject

XtXv, XtwXv, XtXvy = 0, 1, 2
CHAIN_TYPES = {"XtXv": XtXv, "XtwXv": XtwXv, "XtXvy": XtXvy}

AGG_SUM, AGG_SUMSQ, AGG_MAX, AGG_MIN, AGG_MAXINDEX, AGG_MININDEX, AGG_MEAN, AGG_PROD, AGG_VAR = range(9)
DIR_ALL, DIR_ROW, DIR_COL = 0, 1, 2

# opcode -> (agg op, direction): GPUInstructionParser.java:111-133 plus the uarimax/uarimin this library adds
AGG_OPCODES = {
    "uak+": (AGG_SUM, DIR_ALL), "uark+": (AGG_SUM, DIR_ROW), "uack+": (AGG_SUM, DIR_COL),
    "uasqk+": (AGG_SUMSQ, DIR_ALL), "uarsqk+": (AGG_SUMSQ, DIR_ROW), "uacsqk+": (AGG_SUMSQ, DIR_COL),
    "uamax": (AGG_MAX, DIR_ALL), "uarmax": (AGG_MAX, DIR_ROW), "uacmax": (AGG_MAX, DIR_COL),
    "uamin": (AGG_MIN, DIR_ALL), "uarmin": (AGG_MIN, DIR_ROW), "uacmin": (AGG_MIN, DIR_COL),
    "uarimax": (AGG_MAXINDEX, DIR_ROW), "uarimin": (AGG_MININDEX, DIR_ROW),
    "uamean": (AGG_MEAN, DIR_ALL), "uarmean": (AGG_MEAN, DIR_ROW), "uacmean": (AGG_MEAN, DIR_COL),
    "ua*": (AGG_PROD, DIR_ALL), "uar*": (AGG_PROD, DIR_ROW), "uac*": (AGG_PROD, DIR_COL),
    "uavar": (AGG_VAR, DIR_ALL), "uarvar": (AGG_VAR, DIR_ROW), "uacvar": (AGG_VAR, DIR_COL),
}

# LibMatrixCUDA.getBinaryOp numbering (LibMatrixCUDA.java:1551-1575)
BINARY_OPS = {
    "+": 0, "-": 1, "*": 2, "/": 3, "^": 4, "<": 5, "<=": 6, ">": 7, ">=": 8, "==": 9, "!=": 10,
    "min": 11, "max": 12, "&&": 13, "||": 14, "1-*": 15, "-nz": 16, "%%": 17, "%/%": 18,
}
UNARY_OPS = {"exp": 0, "sqrt": 1, "round": 2, "abs": 3, "log": 4, "floor": 5, "ceil": 6, "sign": 7,
             "sigmoid": 8, "sin": 9, "cos": 10, "tan": 11, "tanh": 12}
BCAST_NONE, BCAST_COL, BCAST_ROW, BCAST_OUTER = 0, 1, 2, 3


def _p(x):
    return C.c_void_p(x if x else None)


def _dim_error(msg):
    raise native.NativeError(native.E_DIM, msg)


# ------------------------------------------------------------------------------------------------
def mmchain(gCtx, instName, X, v, w, ctype):
    """t(X)%*%(X%*%v) | t(X)%*%(w*(X%*%v)) | t(X)%*%((X%*%v)-w); output m x 1 dense."""
    ct = CHAIN_TYPES[ctype] if isinstance(ctype, str) else int(ctype)
    n, m = X.rows, X.cols
    if v.rows * v.cols != m:
        _dim_error("mmchain: v has %d cells, X has %d columns" % (v.rows * v.cols, m))
    if ct != XtXv and (w is None or w.rows * w.cols != n):
        _dim_error("mmchain: w must have %d cells" % n)
    out = GPUObject.allocate_dense(gCtx, m, 1)
    wp = w.ptr if (w is not None and ct != XtXv) else 0
    if X.isSparse():
        pc = X.columnOrderedCopy()
        if pc is not None:
            tmp = GPUObject.allocate_dense(gCtx, n, 1)
            pc.mmchain(X.csr, v.ptr, wp, tmp.ptr, out.ptr, ct)
            tmp.free()
        else:
            gCtx.call("sysml_b200_mmchain_csr", _p(X.csr.rowPtr), _p(X.csr.colInd), _p(X.csr.val), _p(v.ptr), _p(wp),
                      _p(out.ptr), n, m, ct)
    else:
        gCtx.call("sysml_b200_mmchain_dense", _p(X.ptr), _p(v.ptr), _p(wp), _p(out.ptr), n, m, ct)
    return out


def mmchain_hostvec(gCtx, X, h_v, h_w, ctype, d_w=None):
    """Host-vector entry: X resident on the device, v/w/out on the host (NativeHelper call shape)."""
    import numpy as np
    ct = CHAIN_TYPES[ctype] if isinstance(ctype, str) else int(ctype)
    h_v = np.ascontiguousarray(h_v, dtype=np.float64).reshape(-1)
    out = np.empty(X.cols, dtype=np.float64)
    hw = None
    if h_w is not None and d_w is None:
        h_w = np.ascontiguousarray(h_w, dtype=np.float64).reshape(-1)
        hw = h_w.ctypes.data_as(C.c_void_p)
    gCtx.call("sysml_b200_mmchain_dense_hostvec", _p(X.ptr), h_v.ctypes.data_as(C.c_void_p), hw,
              _p(d_w.ptr if d_w is not None else 0), out.ctypes.data_as(C.c_void_p), X.rows, X.cols, ct)
    return out


def matmultTSMM(gCtx, instName, left, isLeftTransposed=True, with_nnz=False):
    """t(X)%*%X (isLeftTransposed) or X%*%t(X)."""
    dim = left.cols if isLeftTransposed else left.rows
    out = GPUObject.allocate_dense(gCtx, dim, dim)
    nnz = C.c_int64(-1)
    pn = C.byref(nnz) if with_nnz else None
    if left.isSparse():
        gCtx.call("sysml_b200_tsmm_csr_left" if isLeftTransposed else "sysml_b200_tsmm_csr_right", _p(left.csr.rowPtr),
                  _p(left.csr.colInd), _p(left.csr.val), _p(out.ptr), left.rows, left.cols, pn)
    else:
        gCtx.call("sysml_b200_tsmm_dense", _p(left.ptr), _p(out.ptr), left.rows, left.cols,
                  1 if isLeftTransposed else 0, pn)
    return (out, nnz.value) if with_nnz else out


def matmult(gCtx, instName, left, right, isLeftTransposed=False, isRightTransposed=False):
    """C = op(A) %*% op(B) (dense x dense, CSR x dense)."""
    m = left.cols if isLeftTransposed else left.rows
    k1 = left.rows if isLeftTransposed else left.cols
    k2 = right.cols if isRightTransposed else right.rows
    n = right.rows if isRightTransposed else right.cols
    if k1 != k2:
        _dim_error("matmult: incompatible dimensions %d vs %d" % (k1, k2))
    out = GPUObject.allocate_dense(gCtx, m, n)
    if right.isSparse():
        if isRightTransposed or (left.isSparse() and isLeftTransposed):
            raise native.NativeError(native.E_UNSUPPORTED, "transposed sparse operand in this combination")
        if left.isSparse():
            gCtx.call("sysml_b200_matmult_csr_csr", _p(left.csr.rowPtr), _p(left.csr.colInd), _p(left.csr.val),
                      _p(right.csr.rowPtr), _p(right.csr.colInd), _p(right.csr.val), _p(out.ptr), left.rows, left.cols,
                      right.rows, right.cols)
        else:
            gCtx.call("sysml_b200_matmult_dense_csr", _p(left.ptr), _p(right.csr.rowPtr), _p(right.csr.colInd),
                      _p(right.csr.val), _p(out.ptr), left.rows, left.cols, right.rows, right.cols, int(isLeftTransposed))
        return out
    if left.isSparse():
        if isRightTransposed:
            raise native.NativeError(native.E_UNSUPPORTED, "CSR x t(dense)")
        pc = left.columnOrderedCopy() if (isLeftTransposed and n == 1) else None
        if pc is not None:   # t(X) %*% y as a gather over the column-ordered copy
            pc.vm(right.ptr, out.ptr)
            return out
        gCtx.call("sysml_b200_matmult_csr_dense", _p(left.csr.rowPtr), _p(left.csr.colInd), _p(left.csr.val),
                  _p(right.ptr), _p(out.ptr), left.rows, left.cols, right.rows, right.cols,
                  int(isLeftTransposed))
    else:
        gCtx.call("sysml_b200_matmult_dense", _p(left.ptr), _p(right.ptr), _p(out.ptr), left.rows, left.cols,
                  right.rows, right.cols, int(isLeftTransposed), int(isRightTransposed))
    return out


def unaryAggregate(gCtx, instName, in1, opcode):
    """opcode: 'uak+', 'uark+', 'uack+', 'uarimax', ... ; scalar for ALL, GPUObject otherwise."""
    op, d = AGG_OPCODES[opcode] if isinstance(opcode, str) else opcode
    rows, cols = in1.rows, in1.cols
    orows, ocols = (1, 1) if d == DIR_ALL else ((rows, 1) if d == DIR_ROW else (1, cols))
    out = GPUObject.allocate_dense(gCtx, orows, ocols)
    if in1.isSparse():
        gCtx.call("sysml_b200_agg_unary_csr", _p(in1.csr.rowPtr), _p(in1.csr.colInd), _p(in1.csr.val), rows, cols,
                  op, d, _p(out.ptr))
    else:
        gCtx.call("sysml_b200_agg_unary_dense", _p(in1.ptr), rows, cols, op, d, _p(out.ptr))
    if d == DIR_ALL:
        val = float(out.to_host()[0, 0])
        out.free()
        return val
    return out


def _bcast(A, B):
    if (B.rows, B.cols) == (A.rows, A.cols):
        return BCAST_NONE
    if B.cols == 1 and B.rows == A.rows:
        return BCAST_COL
    if B.rows == 1 and B.cols == A.cols:
        return BCAST_ROW
    _dim_error("cellwise: %dx%d vs %dx%d" % (A.rows, A.cols, B.rows, B.cols))


def matrixMatrixArithmetic(gCtx, instName, in1, in2, op):
    """C = A op B with column/row-vector broadcast of B (also the relational ops); a column vector against a row
    vector gives the outer form (BinaryAccessType.OUTER_VECTOR_VECTOR, LibMatrixBincell.java:150-163)."""
    opc = BINARY_OPS[op] if isinstance(op, str) else int(op)
    if in1.cols == 1 and in2.rows == 1 and in2.cols > 1 and in1.rows > 1:
        out = GPUObject.allocate_dense(gCtx, in1.rows, in2.cols)
        gCtx.call("sysml_b200_binary_mm", _p(in1.ptr), _p(in2.ptr), _p(out.ptr), in1.rows, in2.cols, BCAST_OUTER, opc)
        return out
    out = GPUObject.allocate_dense(gCtx, in1.rows, in1.cols)
    gCtx.call("sysml_b200_binary_mm", _p(in1.ptr), _p(in2.ptr), _p(out.ptr), in1.rows, in1.cols, _bcast(in1, in2), opc)
    return out


matrixMatrixRelational = matrixMatrixArithmetic


def matrixScalarArithmetic(gCtx, instName, in1, scalar, op, scalarLeft=False):
    opc = BINARY_OPS[op] if isinstance(op, str) else int(op)
    out = GPUObject.allocate_dense(gCtx, in1.rows, in1.cols)
    gCtx.call("sysml_b200_binary_ms", _p(in1.ptr), float(scalar), _p(out.ptr), in1.rows * in1.cols, opc, int(scalarLeft))
    return out


def axpy(gCtx, instName, in1, in2, constant):
    """out = in1 + constant*in2 ('+*'; pass -constant for '-*'), in2 may be a row/col vector."""
    out = GPUObject.allocate_dense(gCtx, in1.rows, in1.cols)
    gCtx.call("sysml_b200_axpy", _p(in1.ptr), _p(in2.ptr), float(constant), _p(out.ptr), in1.rows, in1.cols, _bcast(in1, in2))
    return out


def unaryOp(gCtx, instName, in1, op):
    opc = UNARY_OPS[op] if isinstance(op, str) else int(op)
    out = GPUObject.allocate_dense(gCtx, in1.rows, in1.cols)
    gCtx.call("sysml_b200_unary", _p(in1.ptr), _p(out.ptr), in1.rows * in1.cols, opc)
    return out


FUSED_AGG = {None: 0, "none": 0, "sum": 1, "rowSums": 2, "colSums": 3, "min": 4, "rowMins": 5, "colMins": 6,
             "max": 7, "rowMaxs": 8, "colMaxs": 9}


class Expr:
    """Expression builder for fusedCellwise:  (Expr.const(0) - Expr.input(0) * Expr.input(1)).binary(Expr.const(1), ">")
    -> postfix tokens of include/sysml_b200.h (SYSML_FUSED_TOKEN).  Expr.let(value, lambda v: body(v)) evaluates
    `value` once and lets `body` use it several times (PICK / NIP tokens)."""

    def __init__(self, node):
        # a raw token list is accepted as-is (tests of malformed programs)
        self.node = ("raw", list(node)) if isinstance(node, (list, tuple)) and (not node or isinstance(node[0], int)) else node

    @staticmethod
    def input(i):
        return Expr(("in", i))

    @staticmethod
    def const(i):
        return Expr(("const", i))

    @staticmethod
    def let(value, body):
        ref = Expr(("ref", object()))
        return Expr(("let", ref.node[1], value, body(ref)))

    def binary(self, other, op):
        opc = BINARY_OPS[op] if isinstance(op, str) else int(op)
        return Expr(("bin", opc, self, other))

    def unary(self, op):
        opc = UNARY_OPS[op] if isinstance(op, str) else int(op)
        return Expr(("un", opc, self))

    def __add__(self, o):
        return self.binary(o, "+")

    def __sub__(self, o):
        return self.binary(o, "-")

    def __mul__(self, o):
        return self.binary(o, "*")

    def __truediv__(self, o):
        return self.binary(o, "/")

    def _lin(self, depth, env, out):
        n = self.node
        k = n[0]
        if k == "raw":
            out.extend(n[1])
        elif k == "in":
            out.append((0 << 24) | n[1])
        elif k == "const":
            out.append((1 << 24) | n[1])
        elif k == "bin":
            n[2]._lin(depth, env, out)
            n[3]._lin(depth + 1, env, out)
            out.append((2 << 24) | n[1])
        elif k == "un":
            n[2]._lin(depth, env, out)
            out.append((3 << 24) | n[1])
        elif k == "ref":
            out.append((4 << 24) | (depth - 1 - env[n[1]]))
        elif k == "let":
            n[2]._lin(depth, env, out)                      # the shared value sits at stack index `depth`
            n[3]._lin(depth + 1, {**env, n[1]: depth}, out)
            out.append(5 << 24)                             # drop it from under the result

    @property
    def tokens(self):
        out = []
        self._lin(0, {}, out)
        return out


def fusedCellwise(gCtx, instName, inputs, expr, consts=(), agg=None, rows=None, cols=None, out=None, read_back=True):
    """One-pass evaluation of a cellwise chain over `inputs` (GPUObjects: rows x cols matrices, rows x 1 column
    vectors or 1 x cols row vectors), optionally reduced by sum / rowSums / colSums -- the GPU twin of the code
    generator's cellwise template (runtime/codegen/SpoofCellwise.java).  Returns a GPUObject, or a float for
    agg="sum" (read back like LibMatrixCUDA.unaryAggregate does for scalars, LibMatrixCUDA.java:692-693).
    `out` reuses an existing buffer of the right size instead of allocating the result."""
    if rows is None:
        rows = max(i.rows for i in inputs)
    if cols is None:
        cols = max(i.cols for i in inputs)
    kinds = []
    for i in inputs:
        if (i.rows, i.cols) == (rows, cols):
            kinds.append(0)
        elif (i.rows, i.cols) == (rows, 1):
            kinds.append(1)
        elif (i.rows, i.cols) == (1, cols):
            kinds.append(2)
        else:
            raise native.NativeError(native.E_DIM, "fusedCellwise: operand %dx%d does not broadcast to %dx%d" % (i.rows, i.cols, rows, cols))
    tokens = expr.tokens if isinstance(expr, Expr) else list(expr)
    a = FUSED_AGG[agg] if not isinstance(agg, int) else agg
    orow, ocol = {0: (rows, cols), 1: (1, 1), 2: (rows, 1), 3: (1, cols)}[a if a <= 3 else (a - 1) % 3 + 1]
    given = out is not None
    if out is None:
        out = GPUObject.allocate_dense(gCtx, orow, ocol)
    elif out.rows * out.cols != orow * ocol:
        raise native.NativeError(native.E_DIM, "fusedCellwise: out holds %dx%d, result is %dx%d" % (out.rows, out.cols, orow, ocol))
    n = len(inputs)
    ptrs = (C.c_void_p * max(n, 1))(*[i.ptr for i in inputs])
    kd = (C.c_int * max(n, 1))(*kinds)
    cs = (C.c_double * max(len(consts), 1))(*[float(c) for c in consts])
    pg = (C.c_int32 * len(tokens))(*tokens)
    gCtx.call("sysml_b200_fused_cellwise", ptrs, kd, n, cs, len(consts), pg, len(tokens), rows, cols, a, _p(out.ptr))
    if a in (1, 4, 7) and read_back:
        v = float(gCtx.d2h(out.ptr, (1,))[0])
        if not given:
            out.free()
        return v
    return out


def mmchainRowFn(gCtx, instName, X, v, exprs, w=None, y=None, consts=(), out=None, side=None):
    """out = t(X) %*% f(X %*% v, w, y) in ONE pass over a dense X, f an expression over Expr.input(0) = (X %*% v)[i],
    input(1) = w[i], input(2) = y[i]; a second expression, if given, is written per row to `side` (n x 1).
    Generalises MapMultChain's XtXv / XtwXv / XtXvy (MapMultChain.java:31-32) to any row function -- e.g. GLM's outer
    step. Returns (out, side)."""
    n, m = X.rows, X.cols
    if X.isSparse():
        raise native.NativeError(native.E_UNSUPPORTED, "mmchainRowFn: dense X only")
    if v.rows * v.cols != m:
        _dim_error("mmchainRowFn: v has %d cells, X has %d columns" % (v.rows * v.cols, m))
    for o in (w, y):
        if o is not None and o.rows * o.cols != n:
            _dim_error("mmchainRowFn: the row vectors must have %d cells" % n)
    exprs = list(exprs) if isinstance(exprs, (list, tuple)) else [exprs]
    tokens = []
    for e in exprs:
        tokens += e.tokens if isinstance(e, Expr) else list(e)
    if out is None:
        out = GPUObject.allocate_dense(gCtx, m, 1)
    if len(exprs) == 2 and side is None:
        side = GPUObject.allocate_dense(gCtx, n, 1)
    cs = (C.c_double * max(len(consts), 1))(*[float(c) for c in consts])
    pg = (C.c_int32 * len(tokens))(*tokens)
    gCtx.call("sysml_b200_mmchain_rowfn_dense", _p(X.ptr), _p(v.ptr), _p(w.ptr) if w is not None else None,
              _p(y.ptr) if y is not None else None, pg, len(tokens), cs, len(consts), len(exprs),
              _p(side.ptr) if side is not None else None, _p(out.ptr), n, m)
    return out, side


def fusedMultiAggregate(gCtx, instName, inputs, exprs, consts=(), rows=None, cols=None, out=None, read_back=True):
    """sum(expr_j(inputs)) for every expression in ONE pass over the operands -- the code generator's multi-aggregate
    template (runtime/codegen/SpoofMultiAggregate.java). Returns the sums as a list of floats (one read-back).
    `out` may name a device buffer of at least len(exprs) doubles to reuse."""
    if rows is None:
        rows = max(i.rows for i in inputs)
    if cols is None:
        cols = max(i.cols for i in inputs)
    kinds = []
    for i in inputs:
        if (i.rows, i.cols) == (rows, cols):
            kinds.append(0)
        elif (i.rows, i.cols) == (rows, 1):
            kinds.append(1)
        elif (i.rows, i.cols) == (1, cols):
            kinds.append(2)
        else:
            raise native.NativeError(native.E_DIM, "fusedMultiAggregate: operand %dx%d does not broadcast to %dx%d" % (i.rows, i.cols, rows, cols))
    tokens = []
    for e in exprs:      # self-contained postfix runs: each leaves its value above the previous ones
        tokens += e.tokens if isinstance(e, Expr) else list(e)
    k = len(exprs)
    given = out is not None
    if out is None:
        out = GPUObject.allocate_dense(gCtx, k, 1)
    n = len(inputs)
    ptrs = (C.c_void_p * max(n, 1))(*[i.ptr for i in inputs])
    kd = (C.c_int * max(n, 1))(*kinds)
    cs = (C.c_double * max(len(consts), 1))(*[float(c) for c in consts])
    pg = (C.c_int32 * len(tokens))(*tokens)
    try:
        gCtx.call("sysml_b200_fused_multi_agg", ptrs, kd, n, cs, len(consts), pg, len(tokens), rows, cols, k, _p(out.ptr))
        if not read_back:       # the caller combines the sums on the device first (a sharded operand)
            return None
        return [float(v) for v in gCtx.d2h(out.ptr, (k,))]
    finally:
        if not given:
            out.free()


def transpose(gCtx, instName, in1):
    out = GPUObject.allocate_dense(gCtx, in1.cols, in1.rows)
    gCtx.call("sysml_b200_transpose", _p(in1.ptr), _p(out.ptr), in1.rows, in1.cols)
    return out


def _adopt_csr(gCtx, rows, cols, rp, pci, pva, nnz):
    from .gpu_context import CSRPointer
    out = GPUObject(gCtx, rows, cols)
    out.csr = CSRPointer.adopt(gCtx, rows, rp, pci.value, pva.value, nnz.value)
    return out


def sparseTranspose(gCtx, instName, in1):
    """t(X) of a CSR operand, CSR result (CSRPointer.transpose / csr2csc, CSRPointer.java:463-481)."""
    from .gpu_context import CSRPointer
    out = GPUObject(gCtx, in1.cols, in1.rows)
    out.csr = CSRPointer(gCtx, in1.csr.nnz, in1.cols)
    gCtx.call("sysml_b200_csr_transpose", _p(in1.csr.rowPtr), _p(in1.csr.colInd), _p(in1.csr.val), in1.rows, in1.cols,
              _p(out.csr.rowPtr), _p(out.csr.colInd), _p(out.csr.val))
    return out


def sparseSparseAdd(gCtx, instName, in1, in2, alpha=1.0, beta=1.0):
    """alpha*A + beta*B for two CSR operands, CSR result (LibMatrixCUDA.dgeam's sparse branch, LibMatrixCUDA.java:1593-1701)."""
    if (in1.rows, in1.cols) != (in2.rows, in2.cols):
        _dim_error("sparse + sparse: %dx%d vs %dx%d" % (in1.rows, in1.cols, in2.rows, in2.cols))
    rp = gCtx.allocate((in1.rows + 1) * 4)
    pci, pva, nnz = C.c_void_p(), C.c_void_p(), C.c_int64()
    gCtx.call("sysml_b200_csr_add", _p(in1.csr.rowPtr), _p(in1.csr.colInd), _p(in1.csr.val), float(alpha), _p(in2.csr.rowPtr),
              _p(in2.csr.colInd), _p(in2.csr.val), float(beta), in1.rows, in1.cols, _p(rp), C.byref(pci), C.byref(pva), C.byref(nnz))
    return _adopt_csr(gCtx, in1.rows, in1.cols, rp, pci, pva, nnz)


def sparseSparseMatmult(gCtx, instName, in1, in2):
    """CSR %*% CSR with a CSR result (LibMatrixCuMatMult.sparseSparseMatMult, LibMatrixCuMatMult.java:150-195)."""
    if in1.cols != in2.rows:
        _dim_error("sparse %%*%% sparse: inner dimensions %d vs %d" % (in1.cols, in2.rows))
    rp = gCtx.allocate((in1.rows + 1) * 4)
    pci, pva, nnz = C.c_void_p(), C.c_void_p(), C.c_int64()
    gCtx.call("sysml_b200_matmult_csr_csr_sparse", _p(in1.csr.rowPtr), _p(in1.csr.colInd), _p(in1.csr.val), _p(in2.csr.rowPtr),
              _p(in2.csr.colInd), _p(in2.csr.val), in1.rows, in1.cols, in2.rows, in2.cols, _p(rp), C.byref(pci), C.byref(pva),
              C.byref(nnz))
    return _adopt_csr(gCtx, in1.rows, in2.cols, rp, pci, pva, nnz)


def _two_pass_csr(gCtx, rows, cols, fn, *operands):
    """The symbolic / numeric call pair of the CSR-result cellwise entry points (CSRPointer.allocateForDgeam's protocol,
    CSRPointer.java:227-299): sizes from the first call, arrays from this context's allocator, values from the second."""
    from .gpu_context import CSRPointer
    rp = gCtx.allocate((rows + 1) * 4)
    nnz = C.c_int64()
    try:
        gCtx.call(fn, *operands, _p(rp), None, None, C.byref(nnz))
    except Exception:
        gCtx.cudaFreeHelper(rp)
        raise
    out = GPUObject(gCtx, rows, cols)
    out.csr = CSRPointer.adopt(gCtx, rows, rp, gCtx.allocate(max(nnz.value, 1) * 4), gCtx.allocate(max(nnz.value, 1) * 8), nnz.value)
    gCtx.call(fn, *operands, _p(rp), _p(out.csr.colInd), _p(out.csr.val), C.byref(nnz))
    return out


def sparseScalarOp(gCtx, instName, in1, scalar, op, scalar_left=False):
    """CSR op scalar with a CSR result, for (operator, constant) pairs that map 0 to 0 (LibMatrixBincell.safeBinaryScalar,
    LibMatrixBincell.java:938-995); anything else raises (the dense form is the caller's fallback, as in the reference)."""
    return _two_pass_csr(gCtx, in1.rows, in1.cols, "sysml_b200_csr_scalar_op", _p(in1.csr.rowPtr), _p(in1.csr.colInd),
                         _p(in1.csr.val), in1.rows, in1.cols, float(scalar), int(op), 1 if scalar_left else 0)


def sparseUnaryOp(gCtx, instName, in1, un_op):
    """f(CSR) with a CSR result for f(0) == 0 (sqrt, abs, round, floor, ceil, sign, sin, tan, tanh)."""
    return _two_pass_csr(gCtx, in1.rows, in1.cols, "sysml_b200_csr_unary_op", _p(in1.csr.rowPtr), _p(in1.csr.colInd),
                         _p(in1.csr.val), in1.rows, in1.cols, int(un_op))


def sparseDenseOp(gCtx, instName, in1, in2, op, dense_left=False):
    """CSR op dense (matrix, column vector or row vector) evaluated at the CSR operand's stored cells, CSR result
    (LibMatrixBincell.safeBinaryMVSparse :338-440, safeBinaryMMSparseDenseSkip :760-790). The caller vouches that the operator
    is sparse-safe for this dense operand (Multiply, And; Divide by an operand without zeros, isSparseSafeDivide :185-189)."""
    if (in2.rows, in2.cols) == (in1.rows, in1.cols):
        bcast = BCAST_NONE
    elif in2.cols == 1 and in2.rows == in1.rows:
        bcast = BCAST_COL
    elif in2.rows == 1 and in2.cols == in1.cols:
        bcast = BCAST_ROW
    else:
        _dim_error("sparse op dense: %dx%d vs %dx%d" % (in1.rows, in1.cols, in2.rows, in2.cols))
    return _two_pass_csr(gCtx, in1.rows, in1.cols, "sysml_b200_csr_dense_op", _p(in1.csr.rowPtr), _p(in1.csr.colInd),
                         _p(in1.csr.val), in1.rows, in1.cols, _p(in2.ptr), bcast, int(op), 1 if dense_left else 0)


def ultraSparseMatmult(gCtx, instName, in1, in2):
    """CSR (ultra-sparse: a selection / permutation matrix, a handful of cells) %*% dense with a CSR result:
    LibMatrixMult.matrixMultUltraSparse, left branch (LibMatrixMult.java:1510-1566)."""
    if in1.cols != in2.rows:
        _dim_error("ultra-sparse %%*%% dense: inner dimensions %d vs %d" % (in1.cols, in2.rows))
    return _two_pass_csr(gCtx, in1.rows, in2.cols, "sysml_b200_matmult_ultra_sparse_left", _p(in1.csr.rowPtr), _p(in1.csr.colInd),
                         _p(in1.csr.val), _p(in2.ptr), in1.rows, in1.cols, in2.cols)


def sparseSparseOp(gCtx, instName, in1, in2, op):
    """CSR op CSR with a CSR result for operators with 0 op 0 == 0: inner join for Multiply, outer join otherwise
    (LibMatrixBincell.safeBinaryMMSparseSparse / mergeForSparseBinary, LibMatrixBincell.java:597-660, 1279-1340)."""
    if (in1.rows, in1.cols) != (in2.rows, in2.cols):
        _dim_error("sparse op sparse: %dx%d vs %dx%d" % (in1.rows, in1.cols, in2.rows, in2.cols))
    return _two_pass_csr(gCtx, in1.rows, in1.cols, "sysml_b200_csr_csr_op", _p(in1.csr.rowPtr), _p(in1.csr.colInd), _p(in1.csr.val),
                         _p(in2.csr.rowPtr), _p(in2.csr.colInd), _p(in2.csr.val), in1.rows, in1.cols, int(op))


def denseToSparse(gCtx, in1):
    """GPUObject.denseToSparse (GPUObject.java:199-245): two-phase nnz count + fill."""
    from .gpu_context import CSRPointer
    rows, cols = in1.rows, in1.cols
    tmp_rp = gCtx.allocate((rows + 1) * 4)
    nnz = C.c_int64()
    gCtx.call("sysml_b200_dense_to_csr", _p(in1.ptr), rows, cols, _p(tmp_rp), None, None, C.byref(nnz))
    out = GPUObject(gCtx, rows, cols)
    out.csr = CSRPointer(gCtx, nnz.value, rows)
    gCtx.call("sysml_b200_dense_to_csr", _p(in1.ptr), rows, cols, _p(out.csr.rowPtr), _p(out.csr.colInd),
              _p(out.csr.val), C.byref(nnz))
    gCtx.cudaFreeHelper(tmp_rp)
    return out


def sparseToDense(gCtx, in1):
    """GPUObject.sparseToDense (GPUObject.java:508-526)."""
    out = GPUObject.allocate_dense(gCtx, in1.rows, in1.cols)
    gCtx.call("sysml_b200_csr_to_dense", _p(in1.csr.rowPtr), _p(in1.csr.colInd), _p(in1.csr.val), in1.rows, in1.cols,
              _p(out.ptr))
    return out
