"""ctypes binding of libsystemml_b200.so (include/sysml_b200.h).

This is the only place Python touches the native library.  There is NO fallback: if the shared
library is missing or a symbol is absent, importing fails loudly.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(_HERE, "libsystemml_b200.so")

OK, E_INVALID, E_DIM, E_OOM, E_CUDA, E_UNSUPPORTED, E_LIMIT, E_WORKSPACE = 0, -1, -2, -3, -4, -5, -6, -7

vp, i64, u64, ci, dbl = C.c_void_p, C.c_int64, C.c_uint64, C.c_int, C.c_double
pvp, pi64, pint, pdbl = C.POINTER(C.c_void_p), C.POINTER(C.c_int64), C.POINTER(C.c_int), C.POINTER(C.c_double)

# name -> (restype, argtypes): mirrors include/sysml_b200.h declaration by declaration
SIGNATURES = {
    "sysml_b200_device_count": (ci, [pint]),
    "sysml_b200_version": (C.c_char_p, []),
    "sysml_b200_ctx_create": (ci, [ci, pvp]),
    "sysml_b200_ctx_create_on_stream": (ci, [ci, vp, pvp]),
    "sysml_b200_ctx_destroy": (ci, [vp]),
    "sysml_b200_ctx_set_workspace": (ci, [vp, vp, i64]),
    "sysml_b200_ctx_workspace_needed": (i64, [vp]),
    "sysml_b200_ctx_workspace_bytes": (i64, [vp]),
    "sysml_b200_sync": (ci, [vp]),
    "sysml_b200_last_error": (C.c_char_p, [vp]),
    "sysml_b200_ctx_stream": (vp, [vp]),
    "sysml_b200_ctx_device": (ci, [vp]),
    "sysml_b200_ctx_sm_count": (ci, [vp]),
    "sysml_b200_mem_info": (ci, [vp, pi64, pi64]),
    "sysml_b200_launch_count": (i64, [vp]),
    "sysml_b200_malloc": (ci, [vp, i64, pvp]),
    "sysml_b200_free": (ci, [vp, vp]),
    "sysml_b200_memset0": (ci, [vp, vp, i64]),
    "sysml_b200_h2d": (ci, [vp, vp, vp, i64]),
    "sysml_b200_d2h": (ci, [vp, vp, vp, i64]),
    "sysml_b200_d2d": (ci, [vp, vp, vp, i64]),
    "sysml_b200_host_alloc": (ci, [vp, i64, pvp]),
    "sysml_b200_host_free": (ci, [vp, vp]),
    "sysml_b200_timer_start": (ci, [vp]),
    "sysml_b200_timer_stop_ms": (ci, [vp, pdbl]),
    "sysml_b200_mmchain_dense": (ci, [vp, vp, vp, vp, vp, i64, i64, ci]),
    "sysml_b200_mmchain_dense_noalloc": (ci, [vp, vp, vp, vp, vp, i64, i64, ci]),
    "sysml_b200_mmchain_csr": (ci, [vp, vp, vp, vp, vp, vp, vp, i64, i64, ci]),
    "sysml_b200_mmchain_dense_hostvec": (ci, [vp, vp, vp, vp, vp, vp, i64, i64, ci]),
    "sysml_b200_tsmm_dense": (ci, [vp, vp, vp, i64, i64, ci, pi64]),
    "sysml_b200_tsmm_csr_left": (ci, [vp, vp, vp, vp, vp, i64, i64, pi64]),
    "sysml_b200_matmult_dense": (ci, [vp, vp, vp, vp, i64, i64, i64, i64, ci, ci]),
    "sysml_b200_matmult_csr_dense": (ci, [vp, vp, vp, vp, vp, vp, i64, i64, i64, i64, ci]),
    "sysml_b200_matmult_dense_csr": (ci, [vp, vp, vp, vp, vp, vp, i64, i64, i64, i64, ci]),
    "sysml_b200_matmult_csr_csr": (ci, [vp, vp, vp, vp, vp, vp, vp, vp, i64, i64, i64, i64]),
    "sysml_b200_tsmm_csr_right": (ci, [vp, vp, vp, vp, vp, i64, i64, pi64]),
    "sysml_b200_agg_unary_dense": (ci, [vp, vp, i64, i64, ci, ci, vp]),
    "sysml_b200_agg_unary_csr": (ci, [vp, vp, vp, vp, i64, i64, ci, ci, vp]),
    "sysml_b200_binary_mm": (ci, [vp, vp, vp, vp, i64, i64, ci, ci]),
    "sysml_b200_binary_ms": (ci, [vp, vp, dbl, vp, i64, ci, ci]),
    "sysml_b200_axpy": (ci, [vp, vp, vp, dbl, vp, i64, i64, ci]),
    "sysml_b200_unary": (ci, [vp, vp, vp, i64, ci]),
    "sysml_b200_cg_vector_update": (ci, [vp, vp, vp, vp, vp, vp, i64, dbl]),
    "sysml_b200_fill": (ci, [vp, vp, i64, dbl]),
    "sysml_b200_csr_transpose": (ci, [vp, vp, vp, vp, i64, i64, vp, vp, vp]),
    "sysml_b200_csr_add": (ci, [vp, vp, vp, vp, dbl, vp, vp, vp, dbl, i64, i64, vp, vp, vp, pi64]),
    "sysml_b200_matmult_csr_csr_sparse": (ci, [vp, vp, vp, vp, vp, vp, vp, i64, i64, i64, i64, vp, vp, vp, pi64]),
    "sysml_b200_matmult_ultra_sparse_left": (ci, [vp, vp, vp, vp, vp, i64, i64, i64, vp, vp, vp, pi64]),
    "sysml_b200_csr_scalar_op": (ci, [vp, vp, vp, vp, i64, i64, dbl, ci, ci, vp, vp, vp, pi64]),
    "sysml_b200_csr_unary_op": (ci, [vp, vp, vp, vp, i64, i64, ci, vp, vp, vp, pi64]),
    "sysml_b200_csr_dense_op": (ci, [vp, vp, vp, vp, i64, i64, vp, ci, ci, ci, vp, vp, vp, pi64]),
    "sysml_b200_csr_csr_op": (ci, [vp, vp, vp, vp, vp, vp, vp, i64, i64, ci, vp, vp, vp, pi64]),
    "sysml_b200_mmchain_rowfn_dense": (ci, [vp, vp, vp, vp, vp, vp, ci, vp, ci, ci, vp, vp, i64, i64]),
    "sysml_b200_fused_cellwise": (ci, [vp, vp, vp, ci, vp, ci, vp, ci, i64, i64, ci, vp]),
    "sysml_b200_fused_multi_agg": (ci, [vp, vp, vp, ci, vp, ci, vp, ci, i64, i64, ci, vp]),
    "sysml_b200_transpose": (ci, [vp, vp, vp, i64, i64]),
    "sysml_b200_count_nnz": (ci, [vp, vp, i64, pi64]),
    "sysml_b200_dense_to_csr": (ci, [vp, vp, i64, i64, vp, vp, vp, pi64]),
    "sysml_b200_csr_to_dense": (ci, [vp, vp, vp, vp, i64, i64, vp]),
    "sysml_b200_binaryblock_write": (ci, [C.c_char_p, vp, i64, i64, ci, ci, ci, i64]),
    "sysml_b200_binaryblock_read_dense": (ci, [vp, C.c_char_p, i64, i64, ci, ci, i64, i64, vp]),
    "sysml_b200_binaryblock_read_mmchain": (ci, [vp, C.c_char_p, i64, i64, ci, ci, i64, i64, vp, vp, vp, vp, ci]),
    "sysml_b200_rand_uniform": (ci, [vp, vp, i64, u64, i64, dbl, dbl]),
    "sysml_b200_rand_csr": (ci, [vp, i64, i64, dbl, u64, i64, vp, vp, vp, pi64]),
    "sysml_b200_kmeans_step": (ci, [vp, vp, vp, vp, i64, i64, i64, vp, vp, vp]),
    "sysml_b200_psell_segments": (i64, [i64, ci]),
    "sysml_b200_csr_to_psell": (ci, [vp, vp, vp, vp, i64, i64, i64, ci, vp, vp, vp, vp, vp, pi64, pi64]),
    "sysml_b200_matmult_psell_vm": (ci, [vp, vp, vp, vp, vp, vp, i64, i64, ci, vp, vp, ci, vp]),
    "sysml_b200_mmchain_csr_psell": (ci, [vp, vp, vp, vp, vp, vp, vp, vp, vp, ci, vp, vp, vp, vp, i64, i64, ci]),
    "sysml_b200_comm_create": (ci, [vp, ci, ci, i64, pvp, vp]),
    "sysml_b200_comm_connect": (ci, [vp, vp]),
    "sysml_b200_comm_connect_local": (ci, [vp, ci]),
    "sysml_b200_comm_set_timeout_ms": (ci, [vp, i64]),
    "sysml_b200_comm_check": (ci, [vp]),
    "sysml_b200_comm_destroy": (ci, [vp]),
    "sysml_b200_cg_step_dense": (ci, [vp, vp, vp, vp, vp, vp, vp, vp, vp, i64, i64, ci, dbl]),
    "sysml_b200_allreduce_sum": (ci, [vp, vp, vp, i64]),
    "sysml_b200_allreduce": (ci, [vp, vp, vp, i64, ci]),
    "sysml_b200_mmchain_dense_allreduce": (ci, [vp, vp, vp, vp, vp, vp, i64, i64, ci]),
    # ---- include/sysml_b200_host.h ----
    "sysml_b200_ec_create": (ci, [ci, dbl, ci, i64, pvp]),
    "sysml_b200_ec_destroy": (ci, [vp]),
    "sysml_b200_ec_last_error": (C.c_char_p, [vp]),
    "sysml_b200_ec_set_matrix_dense": (ci, [vp, C.c_char_p, vp, i64, i64]),
    "sysml_b200_ec_set_matrix_csr": (ci, [vp, C.c_char_p, vp, vp, vp, i64, i64]),
    "sysml_b200_ec_set_matrix_device": (ci, [vp, C.c_char_p, vp, i64, i64]),
    "sysml_b200_ec_set_scalar": (ci, [vp, C.c_char_p, dbl]),
    "sysml_b200_ec_remove": (ci, [vp, C.c_char_p]),
    "sysml_b200_ec_execute": (ci, [vp, C.c_char_p]),
    "sysml_b200_parse_instruction": (ci, [C.c_char_p, C.c_char_p, ci]),
    "sysml_b200_ec_get_matrix_dims": (ci, [vp, C.c_char_p, pi64, pi64, pi64]),
    "sysml_b200_ec_get_matrix_dense": (ci, [vp, C.c_char_p, vp]),
    "sysml_b200_ec_get_scalar": (ci, [vp, C.c_char_p, pdbl]),
    "sysml_b200_ec_stats": (ci, [vp, pi64]),
    "sysml_b200_ec_set_eviction_policy": (ci, [vp, C.c_char_p]),
    "sysml_b200_eval_sparse_format_in_memory": (ci, [i64, i64, i64]),
    "sysml_b200_cellwise_sparse_safe": (ci, [C.c_char_p, ci, dbl, ci]),
    "sysml_b200_ec_exam_sparsity": (ci, [vp, C.c_char_p, pint]),
    "sysml_b200_ec_is_sparse_on_device": (ci, [vp, C.c_char_p, pint]),
}


class NativeError(RuntimeError):
    """The Python face of DMLRuntimeException for a failed native call."""

    def __init__(self, code, msg):
        super().__init__("sysml_b200 error %d: %s" % (code, msg))
        self.code = code


_lib = None


def load():
    """Loads the shared library and binds EVERY symbol the header declares (raises if one is missing)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(SO_PATH):
        raise ImportError(
            "%s not found: build it with `python -m systemml_b200.build` (there is no CPU fallback)" % SO_PATH)
    lib = C.CDLL(SO_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is missing
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(ctx_handle, rc):
    if rc != 0:
        msg = load().sysml_b200_last_error(ctx_handle)
        raise NativeError(rc, msg.decode("utf-8", "replace") if msg else "?")
