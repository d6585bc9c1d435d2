"""Host-side mirror of the reference's device-state types, over the C ABI.

  GPUContext  <- J/runtime/instructions/gpu/context/GPUContext.java:93-184 (+ GPUMemoryManager.java:157-322)
  GPUObject   <- .../context/GPUObject.java:52-99, 528-607, 755-946
  CSRPointer  <- .../context/CSRPointer.java:76-100

(J/ = src/main/java/org/apache/sysml/).  Names and argument meaning follow the reference so the
parity tests read like its own GPU tests.  Everything numeric happens in the native library.
"""
import ctypes as C
import os

import numpy as np

from . import native


class GPUContext:
    """One device + one stream.  Not thread-safe (GPUContext.java:181-184)."""

    def __init__(self, device=0, stream=None):
        self._lib = native.load()
        h = C.c_void_p()
        if stream is None:
            rc = self._lib.sysml_b200_ctx_create(device, C.byref(h))
        else:
            rc = self._lib.sysml_b200_ctx_create_on_stream(device, C.c_void_p(stream), C.byref(h))
        native.check(None, rc)
        self.handle = h
        self.device = device

    # -- lifecycle -------------------------------------------------------------------------------
    def destroy(self):
        if self.handle:
            self._lib.sysml_b200_ctx_destroy(self.handle)
            self.handle = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.destroy()

    def call(self, name, *args):
        rc = getattr(self._lib, name)(self.handle, *args)
        native.check(self.handle, rc)

    def sync(self):
        self.call("sysml_b200_sync")

    @property
    def stream(self):
        return self._lib.sysml_b200_ctx_stream(self.handle)

    @property
    def sm_count(self):
        return self._lib.sysml_b200_ctx_sm_count(self.handle)

    def launch_count(self):
        return self._lib.sysml_b200_launch_count(self.handle)

    def getAvailableMemory(self):
        """GPUContext.getAvailableMemory (GPUContext.java:218-226)."""
        f, t = C.c_int64(), C.c_int64()
        self.call("sysml_b200_mem_info", C.byref(f), C.byref(t))
        return f.value, t.value

    # -- memory (GPUMemoryManager.malloc / free) ---------------------------------------------------
    def allocate(self, nbytes):
        p = C.c_void_p()
        self.call("sysml_b200_malloc", int(nbytes), C.byref(p))
        return p.value or 0

    def cudaFreeHelper(self, ptr):
        if ptr:
            self.call("sysml_b200_free", C.c_void_p(ptr))

    def h2d(self, dptr, host):
        host = np.ascontiguousarray(host)
        self.call("sysml_b200_h2d", C.c_void_p(dptr), host.ctypes.data_as(C.c_void_p), host.nbytes)

    def d2h(self, dptr, shape, dtype=np.float64):
        out = np.empty(shape, dtype=dtype)
        if out.nbytes:
            self.call("sysml_b200_d2h", out.ctypes.data_as(C.c_void_p), C.c_void_p(dptr), out.nbytes)
        return out

    # -- timers ------------------------------------------------------------------------------------
    def timer_start(self):
        self.call("sysml_b200_timer_start")

    def timer_stop_ms(self):
        ms = C.c_double()
        self.call("sysml_b200_timer_stop_ms", C.byref(ms))
        return ms.value

    # -- convenience constructors ------------------------------------------------------------------
    def dense(self, host):
        return GPUObject.from_host_dense(self, host)

    def csr(self, rowptr, colind, val, shape):
        return GPUObject.from_host_csr(self, rowptr, colind, val, shape)

    def empty(self, rows, cols):
        return GPUObject.allocate_dense(self, rows, cols)

    def rand(self, rows, cols, seed, lo=-1.0, hi=1.0, offset=0):
        o = GPUObject.allocate_dense(self, rows, cols)
        self.call("sysml_b200_rand_uniform", C.c_void_p(o.ptr), rows * cols, seed, offset, lo, hi)
        return o


def _rand_csr(gctx, rows, cols, mean, seed, row_offset=0):
    """GPUObject holding rows [row_offset, row_offset+rows) of the synthetic CSR matrix (sysml_b200_rand_csr)."""
    rowptr = gctx.allocate((rows + 1) * 4)
    nnz = C.c_int64()
    gctx.call("sysml_b200_rand_csr", rows, cols, float(mean), seed, row_offset, C.c_void_p(rowptr), None, None, C.byref(nnz))
    o = GPUObject(gctx, rows, cols)
    o.csr = CSRPointer.adopt(gctx, rows, rowptr, gctx.allocate(max(nnz.value, 1) * 4), gctx.allocate(max(nnz.value, 1) * 8), nnz.value)
    gctx.call("sysml_b200_rand_csr", rows, cols, float(mean), seed, row_offset, C.c_void_p(rowptr), C.c_void_p(o.csr.colInd),
              C.c_void_p(o.csr.val), None)
    return o


GPUContext.rand_csr = _rand_csr


class CSRPointer:
    """val / rowPtr / colInd device arrays + nnz (CSRPointer.java:76-100); int32 indices, 0-based."""

    def __init__(self, gctx, nnz, rows):
        self.gctx = gctx
        self.nnz = int(nnz)
        self.rows = int(rows)
        self.val = gctx.allocate(max(self.nnz, 1) * 8)
        self.rowPtr = gctx.allocate((self.rows + 1) * 4)
        self.colInd = gctx.allocate(max(self.nnz, 1) * 4)

    @staticmethod
    def adopt(gctx, rows, rowPtr, colInd, val, nnz):
        """Takes ownership of arrays the library allocated (results of csr_add / matmult_csr_csr_sparse)."""
        c = CSRPointer.__new__(CSRPointer)
        c.gctx, c.nnz, c.rows = gctx, int(nnz), int(rows)
        c.rowPtr, c.colInd, c.val = rowPtr, colInd or 0, val or 0
        return c

    def deallocate(self):
        for p in (self.val, self.rowPtr, self.colInd):
            self.gctx.cudaFreeHelper(p)
        self.val = self.rowPtr = self.colInd = 0


class PSELLPointer:
    """Column-ordered second copy of a CSR matrix as panelled sliced ELL (csrc/psell.cu) -- what CSRPointer.transpose /
    csr2csc (CSRPointer.java:463-481) would give, cut into row panels so that the gathered vector fits in shared memory,
    and laid out so that a lane owns a column run: t(X) %*% y is a gather with no cross-lane work and no atomics."""
    PANEL_ROWS = 8192

    def __init__(self, gctx, csr, rows, cols, panel_rows=None):
        self.gctx = gctx
        self.rows, self.cols = int(rows), int(cols)
        self.panel_rows = int(panel_rows or self.PANEL_ROWS)
        n_seg = gctx._lib.sysml_b200_psell_segments(rows, self.panel_rows)
        self.sgrp_ptr = gctx.allocate(n_seg * 4)
        pv, pr, pc, pd = C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_void_p()
        ns, ng = C.c_int64(), C.c_int64()
        try:
            gctx.call("sysml_b200_csr_to_psell", C.c_void_p(csr.rowPtr), C.c_void_p(csr.colInd), C.c_void_p(csr.val), rows, cols, csr.nnz,
                      self.panel_rows, C.c_void_p(self.sgrp_ptr), C.byref(pv), C.byref(pr), C.byref(pc), C.byref(pd), C.byref(ns), C.byref(ng))
        except Exception:
            gctx.cudaFreeHelper(self.sgrp_ptr)
            raise
        self.val, self.lrow, self.gcol, self.gdesc = pv.value or 0, pr.value or 0, pc.value or 0, pd.value or 0
        self.slots, self.groups, self.nnz = ns.value, ng.value, csr.nnz

    @staticmethod
    def fits(gctx, rows, cols, panel_rows=None):
        return cols <= 65535 and (cols + 2 * ((panel_rows or PSELLPointer.PANEL_ROWS) + 1)) * 8 <= 226 * 1024

    def args(self):
        return (C.c_void_p(self.val), C.c_void_p(self.lrow), C.c_void_p(self.gcol), C.c_void_p(self.gdesc), C.c_void_p(self.sgrp_ptr))

    def vm(self, y_ptr, out_ptr, w_ptr=None, w_op=0):
        """out = t(X) %*% (y [* w | - w])"""
        self.gctx.call("sysml_b200_matmult_psell_vm", *self.args(), self.rows, self.cols, self.panel_rows, C.c_void_p(y_ptr),
                       C.c_void_p(w_ptr) if w_ptr else None, w_op, C.c_void_p(out_ptr))

    def mmchain(self, csr, v_ptr, w_ptr, tmp_ptr, out_ptr, chain_type):
        self.gctx.call("sysml_b200_mmchain_csr_psell", C.c_void_p(csr.rowPtr), C.c_void_p(csr.colInd), C.c_void_p(csr.val), *self.args(),
                       self.panel_rows, C.c_void_p(v_ptr), C.c_void_p(w_ptr) if w_ptr else None, C.c_void_p(tmp_ptr), C.c_void_p(out_ptr),
                       self.rows, self.cols, chain_type)

    def deallocate(self):
        for p in (self.val, self.lrow, self.gcol, self.gdesc, self.sgrp_ptr):
            self.gctx.cudaFreeHelper(p)
        self.val = self.lrow = self.gcol = self.gdesc = self.sgrp_ptr = 0


class GPUObject:
    """Device state of one matrix: dense row-major double* or a CSRPointer (GPUObject.java:52-99)."""

    def __init__(self, gctx, rows, cols):
        self.gctx = gctx
        self.rows = int(rows)
        self.cols = int(cols)
        self.ptr = 0          # jcudaDenseMatrixPtr
        self.csr = None       # jcudaSparseMatrixPtr
        self.colcopy = None   # column-ordered copy of a CSR matrix (PSELLPointer), built on its second transposed use
        self.transposed_uses = 0
        self.dirty = False
        self.owns = True

    # -- constructors ----------------------------------------------------------------------------
    @staticmethod
    def allocate_dense(gctx, rows, cols):
        o = GPUObject(gctx, rows, cols)
        o.ptr = gctx.allocate(max(rows * cols, 1) * 8)
        return o

    @staticmethod
    def from_host_dense(gctx, host):
        """copyFromHostToDevice, dense branch (GPUObject.java:755-870)."""
        host = np.ascontiguousarray(host, dtype=np.float64)
        if host.ndim == 1:
            host = host.reshape(-1, 1)
        o = GPUObject.allocate_dense(gctx, host.shape[0], host.shape[1])
        if host.size:
            gctx.h2d(o.ptr, host)
        return o

    @staticmethod
    def from_host_csr(gctx, rowptr, colind, val, shape):
        """copyFromHostToDevice, CSR branch (GPUObject.java:785-838; CSRPointer.copyToDevice :153-213)."""
        rows, cols = shape
        nnz = int(rowptr[-1])
        if nnz >= 2**31:
            raise native.NativeError(native.E_LIMIT, "CSR with nnz >= 2^31 (int32 row pointers, CSRPointer.java:76-100)")
        o = GPUObject(gctx, rows, cols)
        o.csr = CSRPointer(gctx, nnz, rows)
        gctx.h2d(o.csr.rowPtr, np.ascontiguousarray(rowptr, dtype=np.int32))
        if nnz:
            gctx.h2d(o.csr.colInd, np.ascontiguousarray(colind, dtype=np.int32))
            gctx.h2d(o.csr.val, np.ascontiguousarray(val, dtype=np.float64))
        return o

    @staticmethod
    def wrap(gctx, ptr, rows, cols):
        """Borrow a device pointer owned by someone else (e.g. a torch tensor)."""
        o = GPUObject(gctx, rows, cols)
        o.ptr = int(ptr)
        o.owns = False
        return o

    # -- state -----------------------------------------------------------------------------------
    def isSparse(self):
        return self.csr is not None

    @property
    def shape(self):
        return (self.rows, self.cols)

    def getNnz(self):
        """GPUObject.getNnz (GPUObject.java:672-691): exact count on the device."""
        if self.csr is not None:
            return self.csr.nnz
        n = C.c_int64()
        self.gctx.call("sysml_b200_count_nnz", C.c_void_p(self.ptr), self.rows * self.cols, C.byref(n))
        return n.value

    def to_host(self):
        """copyFromDeviceToHost (GPUObject.java:879-946).  Dense -> ndarray; CSR -> (rowptr, colind, val)."""
        if self.csr is None:
            return self.gctx.d2h(self.ptr, (self.rows, self.cols))
        rp = self.gctx.d2h(self.csr.rowPtr, (self.rows + 1,), np.int32)
        ci = self.gctx.d2h(self.csr.colInd, (self.csr.nnz,), np.int32)
        va = self.gctx.d2h(self.csr.val, (self.csr.nnz,), np.float64)
        return rp, ci, va

    def columnOrderedCopy(self):
        """The column-ordered copy of a CSR matrix, or None while it is not worth having: the first transposed use of an
        operand goes through the scatter kernels (a one-shot operand never pays for the sort), the second one builds
        the copy, and every later t(X) %*% y / mmchain gathers from it."""
        if self.csr is None or self.csr.nnz == 0 or not PSELLPointer.fits(self.gctx, self.rows, self.cols):
            return None
        if os.environ.get("SYSML_B200_NO_COLUMN_COPY"):   # A/B switch for measurements
            return None
        self.transposed_uses += 1
        if self.colcopy is None and self.transposed_uses >= 2:
            self.colcopy = PSELLPointer(self.gctx, self.csr, self.rows, self.cols)
        return self.colcopy

    def clearData(self):
        """GPUObject.clearData (GPUObject.java:948-970)."""
        if self.colcopy is not None:
            self.colcopy.deallocate()
            self.colcopy = None
        if self.owns and self.ptr:
            self.gctx.cudaFreeHelper(self.ptr)
        self.ptr = 0
        if self.csr is not None:
            self.csr.deallocate()
            self.csr = None

    free = clearData


class Communicator:
    """Sum exchange between the GPUs of one box through NVLink peer memory (csrc/comm.cu).

    The 64-byte IPC handle of every rank's mailbox has to reach all ranks once; `exchange` is any callable
    that maps this rank's bytes to the list of all ranks' bytes in rank order (e.g. built on
    torch.distributed.all_gather)."""

    def __init__(self, gctx, rank, world, max_values, exchange):
        self.gctx = gctx
        self.rank, self.world = rank, world
        lib = gctx._lib
        h = C.c_void_p()
        handle = (C.c_ubyte * 64)()
        native.check(gctx.handle, lib.sysml_b200_comm_create(gctx.handle, rank, world, int(max_values), C.byref(h), handle))
        self.handle = h
        if exchange is None:   # connected later by Communicator.connect_local
            return
        blobs = exchange(bytes(handle))
        assert len(blobs) == world and all(len(b) == 64 for b in blobs)
        joined = b"".join(blobs)
        buf = (C.c_ubyte * len(joined)).from_buffer_copy(joined)
        native.check(gctx.handle, lib.sysml_b200_comm_connect(self.handle, buf))

    @staticmethod
    def connect_local(gctxs, max_values):
        """Ranks that are contexts of ONE process (GPUContext.initializeThread, GPUContext.java:181-184): rank r
        lives on gctxs[r]; no IPC handles, peer access between the devices."""
        world = len(gctxs)
        comms = [Communicator(g, r, world, max_values, None) for r, g in enumerate(gctxs)]
        arr = (C.c_void_p * world)(*[c.handle for c in comms])
        native.check(gctxs[0].handle, gctxs[0]._lib.sysml_b200_comm_connect_local(arr, world))
        return comms

    def set_timeout_ms(self, ms):
        native.check(self.gctx.handle, self.gctx._lib.sysml_b200_comm_set_timeout_ms(self.handle, int(ms)))

    def check(self):
        """Synchronises the context's stream and raises if an exchange failed (peer missing / different call)."""
        native.check(self.gctx.handle, self.gctx._lib.sysml_b200_comm_check(self.handle))

    def allreduce_sum(self, ptr, length):
        self.gctx.call("sysml_b200_allreduce_sum", self.handle, C.c_void_p(ptr), int(length))

    def allreduce(self, ptr, length, op="sum"):
        """op: 'sum' | 'max' | 'min' -- in place, identical bits on every rank."""
        self.gctx.call("sysml_b200_allreduce", self.handle, C.c_void_p(ptr), int(length), {"sum": 0, "max": 1, "min": 2}[op])

    def destroy(self):
        if self.handle:
            self.gctx._lib.sysml_b200_comm_destroy(self.handle)
            self.handle = None


def torch_exchange(group=None):
    """exchange callable for Communicator built on torch.distributed (any backend)."""
    import torch
    import torch.distributed as dist

    def ex(mine):
        world = dist.get_world_size(group)
        backend = dist.get_backend(group)
        dev = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")
        t = torch.tensor(list(mine), dtype=torch.uint8, device=dev)
        outs = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(outs, t, group=group)
        return [bytes(o.cpu().tolist()) for o in outs]

    return ex
