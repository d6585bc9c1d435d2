"""GPUInstructionParser mirror (csrc/host_runtime.cpp): instruction strings as SystemML's lops emit them
(lops/Lop.java:82-86; GPUInstructionParser.java:42-205).  Parsing needs no GPU."""
import pytest

from systemml_b200 import native
from systemml_b200.execution_context import inst, literal, mat, parse_instruction, scalar_var


def test_operand_and_field_layout():
    s = inst("ba+*", mat("A"), mat("B"), mat("C"), "false", "true")
    assert s == "GPU°ba+*°A·MATRIX·DOUBLE°B·MATRIX·DOUBLE°C·MATRIX·DOUBLE°false°true"
    assert parse_instruction(s) == "AggregateBinary ba+* A:MATRIX B:MATRIX C:MATRIX [false] [true]"


@pytest.mark.parametrize("ins,expect", [
    (inst("tsmm", mat("X"), mat("C"), "LEFT"), "MMTSJ tsmm X:MATRIX C:MATRIX [LEFT]"),
    (inst("mmchain", mat("X"), mat("v"), mat("q"), "XtXv", 16), "MMChain mmchain X:MATRIX v:MATRIX q:MATRIX [XtXv]"),
    (inst("mmchain", mat("X"), mat("v"), mat("w"), mat("q"), "XtwXv"), "MMChain mmchain X:MATRIX v:MATRIX w:MATRIX q:MATRIX [XtwXv]"),
    (inst("uark+", mat("X"), mat("r")), "AggregateUnary uark+ X:MATRIX r:MATRIX"),
    (inst("uak+", mat("X"), scalar_var("s")), "AggregateUnary uak+ X:MATRIX s:SCALAR"),
    (inst("uarimax", mat("X"), mat("r")), "AggregateUnary uarimax X:MATRIX r:MATRIX"),
    (inst("*", mat("A"), literal(2.5), mat("C")), "ArithmeticBinary * A:MATRIX 2.5:SCALAR:lit C:MATRIX"),
    (inst("+*", mat("A"), literal(0.5), mat("B"), mat("C")), "ArithmeticBinary +* A:MATRIX 0.5:SCALAR:lit B:MATRIX C:MATRIX"),
    (inst("<=", mat("A"), mat("B"), mat("C")), "RelationalBinary <= A:MATRIX B:MATRIX C:MATRIX"),
    (inst("max", mat("A"), mat("B"), mat("C")), "BuiltinBinary max A:MATRIX B:MATRIX C:MATRIX"),
    (inst("exp", mat("A"), mat("C")), "BuiltinUnary exp A:MATRIX C:MATRIX"),
    (inst("r'", mat("A"), mat("C")), "Reorg r' A:MATRIX C:MATRIX"),
    ("GPU;ba+*;A:MATRIX:DOUBLE;B:MATRIX:DOUBLE;C:MATRIX:DOUBLE;true;false", "AggregateBinary ba+* A:MATRIX B:MATRIX C:MATRIX [true] [false]"),
])
def test_parse(ins, expect):
    assert parse_instruction(ins) == expect


@pytest.mark.parametrize("ins,code,frag", [
    (inst("tsmm", mat("X"), mat("C"), "UP"), native.E_INVALID, "MMTSJType"),
    (inst("ba+*", mat("A"), mat("B"), mat("C"), "false"), native.E_INVALID, "checkNumFields"),
    ("CP°ba+*°A°B°C°false°false", native.E_INVALID, "Not a GPU instruction"),
    (inst("conv2d", mat("A"), mat("B")), native.E_UNSUPPORTED, "Invalid GPU Instruction Type"),
    (inst("mmchain", mat("X"), mat("v"), mat("q"), "XtwXv"), native.E_INVALID, "ChainType"),
    (inst("+*", mat("A"), mat("S"), mat("B"), mat("C")), native.E_INVALID, "scalar"),
    ("", native.E_INVALID, "empty"),
])
def test_parse_errors(ins, code, frag):
    with pytest.raises(native.NativeError) as ei:
        parse_instruction(ins)
    assert ei.value.code == code and frag in str(ei.value)


def test_eval_sparse_format_in_memory_known_answers():
    """MatrixBlock.evalSparseFormatInMemory (MatrixBlock.java:1001-1015) by hand:
    sparse iff nnz/rows/cols < 0.4 AND the MCSR estimate 84 + 8*rows + rlen*(76 + 12*cnnz) is below 44 + 8*rows*cols,
    with cnnz = max(4, ceil(sp*cols)), rlen = min(rows, ceil(sp*rows*cols)) (SparseBlockMCSR.java:101-116)."""
    from systemml_b200.execution_context import eval_sparse_format_in_memory as ev
    assert ev(1000, 1000, 1000) is True          # sp 0.001: 84 + 8000 + 1000*(76+48) = 132084 < 8000044
    assert ev(1000, 1000, 399_999) is True       # sp just under 0.4: 84 + 8000 + 1000*(76+4800) = 4884084 < 8000044
    assert ev(1000, 1000, 400_000) is False      # sp == 0.4 is not below the turn point
    assert ev(1000, 1, 100) is False             # column vector: 84 + 8000 + 100*(76+48) = 20484 > 8044
    assert ev(100, 4, 40) is False               # sp 0.1: 84 + 800 + 40*124 = 5844 > 3244
    assert ev(10, 1000, 100) is True             # sp 0.01: cnnz 10, rlen 10: 84 + 80 + 10*196 = 2124 < 80044
    assert ev(0, 10, 0) is False


# ---- which cellwise instructions keep a CSR operand CSR: hand-derived from the constructors of ScalarOperator
# (J/runtime/matrix/operators/ScalarOperator.java:53-59, isSparseSafeStatic :84-89), RightScalarOperator (:42-49),
# LeftScalarOperator (:41-46), BinaryOperator (:60-66), UnaryOperator (:38-47) -- not produced by running anything ----
SCALAR_RULES = [
    # opcode, constant, scalar on the left?, sparse-safe?
    ("*", 3.0, False, True), ("*", 0.0, True, True), ("*2", 2.0, False, True), ("^2", 2.0, False, True), ("&&", 1.0, False, True),
    ("-nz", 5.0, False, True),
    ("!=", 0.0, False, True), ("!=", 1.0, False, False), ("==", 2.0, False, True), ("==", 0.0, False, False),
    ("-", 0.0, False, True), ("-", 1.0, False, False), ("+", 0.0, False, False),          # Plus is not in the list, even with 0
    ("max", 0.0, False, True), ("max", -1.0, True, True), ("max", 0.5, False, False),
    ("min", 0.0, False, True), ("min", 2.0, True, True), ("min", -0.5, False, False),
    (">", 0.0, False, True), (">", 1.0, False, True), (">", -1.0, False, False),              # X > c is 0 at 0 iff c >= 0
    (">=", 1.0, False, True), (">=", 0.0, False, False),
    ("<", 0.0, False, True), ("<", -1.0, False, True), ("<", 1.0, False, False),
    ("<=", -1.0, False, True), ("<=", 0.0, False, False),
    ("/", 2.0, False, True), ("/", 0.0, False, False), ("/", 2.0, True, False),               # c / X is infinite at 0
    ("^", 2.0, False, True), ("^", 0.0, False, False),
    ("^", -1.0, False, False),    # the reference says true (RightScalarOperator.java:47, "Power && cst != 0"); 0^-1 is Inf: dense here
    (">", 0.0, True, True), (">", -1.0, True, True), (">", 1.0, True, False),                 # c > X at X = 0
    (">=", -1.0, True, True), (">=", 0.0, True, False),
    ("<", 0.0, True, True), ("<", 1.0, True, True), ("<", -1.0, True, False),
    ("<=", 1.0, True, True), ("<=", 0.0, True, False),
    ("%%", 3.0, False, False), ("%/%", 3.0, False, False), ("1-*", 0.0, False, False), ("||", 0.0, False, False),
]


@pytest.mark.parametrize("opcode,c,left,exp", SCALAR_RULES)
def test_scalar_sparse_safety_rule(opcode, c, left, exp):
    from systemml_b200 import native
    assert native.load().sysml_b200_cellwise_sparse_safe(opcode.encode(), 0, float(c), int(left)) == int(exp)


def test_binary_and_unary_sparse_safety_rules():
    from systemml_b200 import native
    lib = native.load()
    safe2 = {"+", "-", "*", "&&", "||"}          # BinaryOperator: Plus, Multiply, Minus, And, Or (+*/-* take the axpy path)
    for o in ("+", "-", "*", "/", "^", "<", "<=", ">", ">=", "==", "!=", "min", "max", "&&", "||", "1-*", "-nz", "%%", "%/%"):
        assert lib.sysml_b200_cellwise_sparse_safe(o.encode(), 1, 0.0, 0) == int(o in safe2), o
    safe1 = {"sin", "tan", "tanh", "round", "abs", "sqrt", "sign"}   # UnaryOperator: SIN TAN (SINH) TANH ROUND ABS SQRT (SPROP LOG_NZ) SIGN
    for o in ("exp", "log", "abs", "sqrt", "round", "floor", "ceil", "sin", "cos", "tan", "tanh", "sign", "sigmoid"):
        assert lib.sysml_b200_cellwise_sparse_safe(o.encode(), 2, 0.0, 0) == int(o in safe1), o
    assert lib.sysml_b200_cellwise_sparse_safe(b"*", 7, 0.0, 0) == -1
