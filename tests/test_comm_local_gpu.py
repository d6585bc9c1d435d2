"""The cross-rank exchange with ranks that are contexts of ONE process (sysml_b200_comm_connect_local): the way the
reference drives several GPUs (one GPUContext per parfor worker thread of one JVM, GPUContext.java:181-184).  Needs
one GPU only: the ranks share device 0 (different streams), so the protocol -- mailbox stores, 64-bit {signature, epoch}
flags, rank-ordered sum, both call parities, failure detection -- is exercised on any box; with several GPUs the same
body runs one rank per device.  Checks against the single-context result and the CPU oracle
(LibMatrixMult.matrixMultChain, LibMatrixMult.java:243-329) with the GPUTests.assertEqualMatrices rule."""
import ctypes as C
import threading

import numpy as np
import pytest

from tests.util import assert_equal_matrices, max_rel_err

pytestmark = pytest.mark.gpu


def _ngpus():
    from systemml_b200 import native
    n = C.c_int(0)
    native.load().sysml_b200_device_count(C.byref(n))
    return n.value


def _contexts(world, spread):
    from systemml_b200 import GPUContext
    ng = _ngpus()
    return [GPUContext(r % ng if spread else 0) for r in range(world)]


def _p(o):
    return C.c_void_p(o.ptr)


@pytest.mark.parametrize("world,spread", [(2, False), (3, False), (4, False), (2, True), (4, True), (8, True)])
def test_local_ranks_mmchain_allreduce_and_cg_step(world, spread):
    """spread=False: all ranks on device 0.  spread=True: one rank per GPU (skipped when the box has fewer)."""
    if spread and _ngpus() < world:
        pytest.skip("needs %d GPUs" % world)
    from oracle import oracle as O
    from systemml_b200 import Communicator, LibMatrixB200 as L
    from systemml_b200.sharding import row_shard
    n, m = 40_000, 200   # 7 exchange CTAs per rank: every rank's finishing kernel stays resident on a shared device
    gs = _contexts(world, spread)
    comms = Communicator.connect_local(gs, 4096)
    for c in comms:
        c.set_timeout_ms(20_000)
    Xh = O.rand_uniform(n * m, seed=1).reshape(n, m)
    wh = O.rand_uniform(n, seed=4, lo=0.0, hi=1.0)
    shards = [row_shard(n, world, r) for r in range(world)]
    X = [g.rand(b - a, m, seed=1, offset=a * m) for g, (a, b) in zip(gs, shards)]
    w = [g.rand(b - a, 1, seed=4, lo=0.0, hi=1.0, offset=a) for g, (a, b) in zip(gs, shards)]
    k = O.num_threads()
    for it, ct in enumerate((1, 0, 2, 1, 1)):   # both mailbox parities, growing epochs
        vh = O.rand_uniform(m, seed=10 + it)
        v = [g.dense(vh.reshape(m, 1)) for g in gs]
        out = [g.empty(m, 1) for g in gs]
        for r, g in enumerate(gs):              # every rank's call is queued before anybody synchronises
            a, b = shards[r]
            g.call("sysml_b200_mmchain_dense_allreduce", comms[r].handle, _p(X[r]), _p(v[r]), _p(w[r]) if ct else None, _p(out[r]), b - a, m, ct)
        res = [o.to_host().reshape(-1) for o in out]
        for r in range(1, world):
            assert np.array_equal(res[0], res[r]), "rank %d differs from rank 0" % r
        exp = O.mmchain_dense(Xh, vh, wh if ct else None, ct, k)
        print("world %d ct %d: max element-wise relative error vs oracle %.3e" % (world, ct, max_rel_err(exp, res[0])))
        assert_equal_matrices(exp, res[0], what="local-rank mmchain ct %d" % ct)
        for o in v + out:
            o.free()
    # whole CG iterations: cg_step_dense on the ranks against mmchain + cg_vector_update on one context holding all of X
    g0 = gs[0]
    Xf, wf = g0.rand(n, m, seed=1), g0.rand(n, 1, seed=4, lo=0.0, hi=1.0)
    ph = O.rand_uniform(m, seed=20)
    state = lambda g: [g.dense(ph.reshape(m, 1)), g.empty(m, 1), g.dense(-ph.reshape(m, 1)), g.dense(np.zeros((m, 1))),
                       g.dense(np.array([[float(ph @ ph)]]))]   # p, q, r, beta, nr2
    ref = state(g0)
    per_rank = [state(g) for g in gs]
    for step in range(4):
        g0.call("sysml_b200_mmchain_dense", _p(Xf), _p(ref[0]), _p(wf), _p(ref[1]), n, m, L.XtwXv)
        g0.call("sysml_b200_cg_vector_update", _p(ref[0]), _p(ref[1]), _p(ref[2]), _p(ref[3]), _p(ref[4]), m, 0.01)
        for r, g in enumerate(gs):
            a, b = shards[r]
            p_, q_, r_, b_, n2 = per_rank[r]
            g.call("sysml_b200_cg_step_dense", comms[r].handle, _p(X[r]), _p(w[r]), _p(p_), _p(q_), _p(r_), _p(b_), _p(n2), b - a, m, L.XtwXv, 0.01)
        for c in comms:
            c.check()
    want = [o.to_host().reshape(-1) for o in ref]
    for r in range(world):
        got = [o.to_host().reshape(-1) for o in per_rank[r]]
        for name, e, a_ in zip(("p", "q", "r", "beta", "nr2"), want, got):
            assert_equal_matrices(e, a_, rel=1e-9, what="cg_step rank %d %s" % (r, name))
        if r:
            first = [o.to_host().reshape(-1) for o in per_rank[0]]
            assert all(np.array_equal(x, y) for x, y in zip(first, got)), "CG state differs between ranks"
    # single context, no communicator: same entry point
    solo = state(g0)
    ref2 = state(g0)
    for step in range(3):
        g0.call("sysml_b200_mmchain_dense", _p(Xf), _p(ref2[0]), _p(wf), _p(ref2[1]), n, m, L.XtwXv)
        g0.call("sysml_b200_cg_vector_update", _p(ref2[0]), _p(ref2[1]), _p(ref2[2]), _p(ref2[3]), _p(ref2[4]), m, 0.0)
        g0.call("sysml_b200_cg_step_dense", None, _p(Xf), _p(wf), _p(solo[0]), _p(solo[1]), _p(solo[2]), _p(solo[3]), _p(solo[4]), n, m, L.XtwXv, 0.0)
    for name, e, a_ in zip(("p", "q", "r", "beta", "nr2"), ref2, solo):
        assert_equal_matrices(e.to_host(), a_.to_host(), rel=1e-9, what="cg_step (no comm) %s" % name)
    for c in comms:
        c.destroy()
    for g in gs:
        g.destroy()


def test_local_ranks_driven_by_threads():
    """One host thread per rank, as parfor workers do: each thread makes its own calls and synchronises on its own."""
    from oracle import oracle as O
    from systemml_b200 import Communicator
    world, length = 2, 1000
    gs = _contexts(world, _ngpus() >= world)
    comms = Communicator.connect_local(gs, 1024)
    out, errs = [None] * world, []
    # buffers are allocated up front and freed at the end: cudaFree waits for the whole device, and on a shared device
    # that includes the other rank's kernel, which may already be waiting for this rank's next call
    bufs = [[gs[r].rand(length, 1, seed=100 + 10 * it + r) for it in range(6)] for r in range(world)]

    def work(r):
        try:
            for it in range(6):
                comms[r].allreduce(bufs[r][it].ptr, length, "sum")
                out[r] = bufs[r][it].to_host().reshape(-1)   # synchronises this rank's stream only
        except Exception as e:  # noqa: BLE001
            errs.append(e)

    ts = [threading.Thread(target=work, args=(r,)) for r in range(world)]
    [t.start() for t in ts]
    [t.join(120) for t in ts]
    assert not errs, errs
    exp = O.rand_uniform(length, seed=150) + O.rand_uniform(length, seed=151)
    assert np.array_equal(out[0], exp) and np.array_equal(out[1], exp)
    for c in comms:
        c.destroy()
    for g in gs:
        g.destroy()


def test_missing_peer_is_reported_not_hung():
    """A rank that never makes the call: the other rank's kernel gives up after the budget and the error names the peer."""
    from systemml_b200 import Communicator, native
    gs = _contexts(2, False)
    comms = Communicator.connect_local(gs, 1024)
    comms[0].set_timeout_ms(300)
    buf = gs[0].rand(100, 1, seed=1)
    comms[0].allreduce(buf.ptr, 100, "sum")          # rank 1 stays silent
    with pytest.raises(native.NativeError) as ei:
        comms[0].check()
    assert "peer 1 did not arrive" in str(ei.value)
    with pytest.raises(native.NativeError):          # the communicator stays failed
        comms[0].allreduce(buf.ptr, 100, "sum")
    for c in comms:
        c.destroy()
    for g in gs:
        g.destroy()


def test_mismatched_call_sequence_is_reported():
    """Ranks that issue different calls (other length / combiner) must not be summed: the signature in the flag differs."""
    from systemml_b200 import Communicator, native
    gs = _contexts(2, False)
    comms = Communicator.connect_local(gs, 1024)
    for c in comms:
        c.set_timeout_ms(2000)
    a, b = gs[0].rand(64, 1, seed=1), gs[1].rand(64, 1, seed=2)
    comms[0].allreduce(a.ptr, 64, "sum")
    comms[1].allreduce(b.ptr, 64, "max")
    failures = 0
    for c in comms:
        try:
            c.check()
        except native.NativeError as e:
            failures += 1
            assert "different calls" in str(e)
    assert failures == 2
    for c in comms:
        c.destroy()
    for g in gs:
        g.destroy()
