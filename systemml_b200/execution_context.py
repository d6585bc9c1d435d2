"""ExecutionContext -- Python face of the C++ host runtime (csrc/host_runtime.cpp, include/sysml_b200_host.h):
SystemML GPU instruction strings in, matrices and scalars out.

Mirrors how the reference drives a GPU instruction: ProgramBlock.executeSingleInstruction
(J/runtime/controlprogram/ProgramBlock.java:241-255) -> GPUInstructionParser.parseSingleInstruction ->
<X>GPUInstruction.processInstruction(ec), with variables living in the ExecutionContext's symbol table.
"""
import ctypes as C

import numpy as np

from . import native

OD, VP = "°", "·"  # Lop.OPERAND_DELIMITOR / VALUETYPE_PREFIX (lops/Lop.java:83-86)


def mat(name):
    return "%s%sMATRIX%sDOUBLE" % (name, VP, VP)


def scalar_var(name):
    return "%s%sSCALAR%sDOUBLE" % (name, VP, VP)


def literal(value):
    return "%r%sSCALAR%sDOUBLE%strue" % (float(value), VP, VP, VP)


def inst(opcode, *fields):
    """Builds 'GPU°opcode°f1°f2...' exactly as the lops' getInstructions() do."""
    return OD.join(("GPU", opcode) + tuple(str(f) for f in fields))


class ExecutionContext:
    STAT_NAMES = ("instructions", "h2d", "d2h", "cuda_malloc", "freelist_reuse", "cuda_free", "evictions",
                  "allocated_bytes", "freelist_bytes")

    def __init__(self, device=0, mem_util_factor=0.9, eager_cuda_free=False, memory_limit=0):
        self._lib = native.load()
        h = C.c_void_p()
        rc = self._lib.sysml_b200_ec_create(device, mem_util_factor, int(eager_cuda_free), int(memory_limit), C.byref(h))
        if rc:
            raise native.NativeError(rc, self._lib.sysml_b200_ec_last_error(None).decode())
        self.handle = h

    def _check(self, rc):
        if rc:
            raise native.NativeError(rc, self._lib.sysml_b200_ec_last_error(self.handle).decode("utf-8", "replace"))

    def destroy(self):
        if self.handle:
            self._lib.sysml_b200_ec_destroy(self.handle)
            self.handle = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.destroy()

    # -- symbol table ------------------------------------------------------------------------------
    def setMatrix(self, name, value):
        a = np.ascontiguousarray(value, dtype=np.float64)
        if a.ndim == 1:
            a = a.reshape(-1, 1)
        self._check(self._lib.sysml_b200_ec_set_matrix_dense(self.handle, name.encode(), a.ctypes.data_as(C.c_void_p), a.shape[0], a.shape[1]))

    def setMatrixCSR(self, name, rowptr, colind, val, shape):
        rp = np.ascontiguousarray(rowptr, dtype=np.int32)
        ci = np.ascontiguousarray(colind, dtype=np.int32)
        va = np.ascontiguousarray(val, dtype=np.float64)
        self._check(self._lib.sysml_b200_ec_set_matrix_csr(self.handle, name.encode(), rp.ctypes.data_as(C.c_void_p),
                                                           ci.ctypes.data_as(C.c_void_p), va.ctypes.data_as(C.c_void_p), shape[0], shape[1]))

    def setMatrixDevice(self, name, dptr, rows, cols):
        """Bind a dense matrix that already lives on the device (borrowed, never freed here)."""
        self._check(self._lib.sysml_b200_ec_set_matrix_device(self.handle, name.encode(), C.c_void_p(dptr), rows, cols))

    def setScalar(self, name, value):
        self._check(self._lib.sysml_b200_ec_set_scalar(self.handle, name.encode(), float(value)))

    def removeVariable(self, name):
        self._check(self._lib.sysml_b200_ec_remove(self.handle, name.encode()))

    def getMatrixDims(self, name):
        r, c, z = C.c_int64(), C.c_int64(), C.c_int64()
        self._check(self._lib.sysml_b200_ec_get_matrix_dims(self.handle, name.encode(), C.byref(r), C.byref(c), C.byref(z)))
        return r.value, c.value, z.value

    def getMatrix(self, name):
        r, c, _ = self.getMatrixDims(name)
        out = np.empty((r, c), dtype=np.float64)
        self._check(self._lib.sysml_b200_ec_get_matrix_dense(self.handle, name.encode(), out.ctypes.data_as(C.c_void_p)))
        return out

    def getScalar(self, name):
        v = C.c_double()
        self._check(self._lib.sysml_b200_ec_get_scalar(self.handle, name.encode(), C.byref(v)))
        return v.value

    # -- execution ---------------------------------------------------------------------------------
    def execute(self, instruction):
        self._check(self._lib.sysml_b200_ec_execute(self.handle, instruction.encode("utf-8")))

    def setEvictionPolicy(self, policy):
        """sysml.gpu.eviction.policy: 'lru' | 'lfu' | 'min_evict' | 'mru' (DMLScript.EvictionPolicy)."""
        self._check(self._lib.sysml_b200_ec_set_eviction_policy(self.handle, policy.encode()))

    def examSparsity(self, name):
        """MatrixBlock.examSparsity on the host block of `name`; returns True when it is now held as CSR."""
        f = C.c_int()
        self._check(self._lib.sysml_b200_ec_exam_sparsity(self.handle, name.encode(), C.byref(f)))
        return bool(f.value)

    def isSparseOnDevice(self, name):
        """GPUObject.isSparse: True / False for a CSR / dense device block, None when the variable has none."""
        f = C.c_int()
        self._check(self._lib.sysml_b200_ec_is_sparse_on_device(self.handle, name.encode(), C.byref(f)))
        return None if f.value < 0 else bool(f.value)

    def stats(self):
        out = (C.c_int64 * 9)()
        self._check(self._lib.sysml_b200_ec_stats(self.handle, out))
        return dict(zip(self.STAT_NAMES, list(out)))


def eval_sparse_format_in_memory(rows, cols, nnz):
    """MatrixBlock.evalSparseFormatInMemory (MatrixBlock.java:1001-1015)."""
    return bool(native.load().sysml_b200_eval_sparse_format_in_memory(int(rows), int(cols), int(nnz)))


def parse_instruction(instruction):
    lib = native.load()
    buf = C.create_string_buffer(1024)
    rc = lib.sysml_b200_parse_instruction(instruction.encode("utf-8"), buf, 1024)
    if rc:
        raise native.NativeError(rc, lib.sysml_b200_ec_last_error(None).decode("utf-8", "replace"))
    return buf.value.decode()
