"""systemml_b200 -- B200-native (sm_100a) implementation of SystemML's data-parallel linear-algebra
hot path (mmchain, tsmm, dense/CSR matmult, aggregate-unary, cellwise), behind the C ABI of
include/sysml_b200.h.  Python here is only the test/bench harness and a mirror of the reference's
host-side interface; all compute is in libsystemml_b200.so.  There is no CPU fallback.
"""
from . import native  # noqa: F401
from .gpu_context import GPUContext, GPUObject, CSRPointer, Communicator  # noqa: F401
from . import lib_matrix as LibMatrixB200  # noqa: F401

__all__ = ["native", "GPUContext", "GPUObject", "CSRPointer", "LibMatrixB200"]
