"""The allocation-free mmchain entry point against the regular one (collected last on purpose: if it
fails under `pytest -x`, every other test has already run).

sysml_b200_mmchain_dense_noalloc is for callers that cannot let this library allocate device memory for
them -- a host runtime that budgets the whole card itself, like SystemML's GPUMemoryManager, or code
running where the CUDA runtime must not be asked for memory.  It reduces through a fixed array that is
part of the loaded library instead of cudaMalloc'ed scratch.  The array is shared by all calls and keeps
the previous call's contents, so the test makes two different calls in a row and checks both.
"""
import ctypes as C

import numpy as np
import pytest

from tests.util import assert_equal_matrices

pytestmark = pytest.mark.gpu


def _call(gctx, name, X, v, w, out, n, m, ct):
    fn = getattr(gctx._lib, name)
    return fn(gctx.handle, C.c_void_p(X.ptr), C.c_void_p(v.ptr), C.c_void_p(w.ptr) if w is not None else None,
              C.c_void_p(out.ptr), n, m, ct)


def test_noalloc_matches_regular_entry_point_twice_in_a_row(gctx):
    n, m = 60_000, 1000
    X = gctx.rand(n, m, seed=71)
    w = gctx.rand(n, 1, seed=72, lo=0.0, hi=1.0)
    out_ref, out_new = gctx.empty(m, 1), gctx.empty(m, 1)
    # two consecutive calls with different vectors AND different chain types; the second call finds the
    # first call's partial sums still sitting in the shared array
    for seed, ct in ((73, 1), (74, 2)):
        v = gctx.rand(m, 1, seed=seed)
        assert _call(gctx, "sysml_b200_mmchain_dense", X, v, w, out_ref, n, m, ct) == 0
        assert _call(gctx, "sysml_b200_mmchain_dense_noalloc", X, v, w, out_new, n, m, ct) == 0
        ref, new = out_ref.to_host(), out_new.to_host()
        assert np.isfinite(new).all()
        assert_equal_matrices(ref, new, what="noalloc vs regular, chain type %d" % ct)
        v.free()
    # a narrow matrix takes a different tile configuration (several row groups per CTA)
    n2, m2 = 20_001, 100
    X2 = gctx.rand(n2, m2, seed=75)
    v2 = gctx.rand(m2, 1, seed=76)
    o1, o2 = gctx.empty(m2, 1), gctx.empty(m2, 1)
    assert _call(gctx, "sysml_b200_mmchain_dense", X2, v2, None, o1, n2, m2, 0) == 0
    assert _call(gctx, "sysml_b200_mmchain_dense_noalloc", X2, v2, None, o2, n2, m2, 0) == 0
    assert_equal_matrices(o1.to_host(), o2.to_host(), what="noalloc vs regular, 20001 x 100")
    for o in (X, w, out_ref, out_new, X2, v2, o1, o2):
        o.free()


def test_noalloc_agrees_with_the_cpu_oracle_and_leaves_free_memory_alone():
    """On a context of its own that has never made a regular call: after the first (warm-up) call has made the
    library's device code resident, a call changes neither the free device memory nor anything else the library holds,
    and its result is the reference's (LibMatrixMult.matrixMultChain, LibMatrixMult.java:243-329) to 1e-9 per cell."""
    from oracle import oracle as O
    from systemml_b200 import GPUContext
    with GPUContext(0) as g:
        n, m = 30_000, 567
        Xh = O.rand_uniform(n * m, seed=81).reshape(n, m)
        vh, wh = O.rand_uniform(m, seed=82), O.rand_uniform(n, seed=83, lo=0.0, hi=1.0)
        X, v, w, out = g.dense(Xh), g.dense(vh.reshape(m, 1)), g.dense(wh.reshape(n, 1)), g.empty(m, 1)
        assert _call(g, "sysml_b200_mmchain_dense_noalloc", X, v, w, out, n, m, 1) == 0    # warm-up: loads the device code
        g.sync()
        free0, _ = g.getAvailableMemory()
        for ct in (1, 0, 2):
            assert _call(g, "sysml_b200_mmchain_dense_noalloc", X, v, w if ct else None, out, n, m, ct) == 0
            g.sync()
            free1, _ = g.getAvailableMemory()
            assert free1 == free0, "a call must not change the free device memory (%d -> %d)" % (free0, free1)
            got = out.to_host().reshape(-1)
            assert_equal_matrices(O.mmchain_dense(Xh, vh, wh if ct else None, ct, 4), got, what="noalloc vs oracle, chain type %d" % ct)
        for o in (X, v, w, out):
            o.free()


def test_noalloc_rejects_sizes_beyond_its_fixed_scratch(gctx):
    from systemml_b200 import native
    m = 1025
    X, v, out = gctx.rand(8, m, seed=1), gctx.rand(m, 1, seed=2), gctx.empty(m, 1)
    rc = _call(gctx, "sysml_b200_mmchain_dense_noalloc", X, v, None, out, 8, m, 0)
    assert rc == native.E_LIMIT
    msg = gctx._lib.sysml_b200_last_error(gctx.handle).decode()
    assert "1024" in msg and "1025" in msg
    # the regular entry point takes the same input
    assert _call(gctx, "sysml_b200_mmchain_dense", X, v, None, out, 8, m, 0) == 0
    for o in (X, v, out):
        o.free()
