"""Two real GPUs: row-sharded mmchain with the sum exchange fused into the finishing kernel (NVLink peer
memory, csrc/comm.cu) against the single-device result and the CPU oracle.  Skipped on a one-GPU box."""
import os
import socket

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _ngpus():
    import ctypes as C
    from systemml_b200 import native
    n = C.c_int(0)
    native.load().sysml_b200_device_count(C.byref(n))
    return n.value


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n, m, out_dir):
    import ctypes as C
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from systemml_b200 import Communicator, GPUContext, LibMatrixB200 as L
    from systemml_b200.gpu_context import torch_exchange
    from systemml_b200.sharding import row_shard
    g = GPUContext(rank)
    comm = Communicator(g, rank, world, 4096, torch_exchange())
    r0, r1 = row_shard(n, world, rank)
    X = g.rand(r1 - r0, m, seed=1, offset=r0 * m)
    w = g.rand(r1 - r0, 1, seed=4, lo=0.0, hi=1.0, offset=r0)
    res = {}
    for it, ct in enumerate((1, 0, 2, 1, 1)):  # several calls in a row: both mailbox parities, growing epochs
        v = g.rand(m, 1, seed=10 + it)
        out = g.empty(m, 1)
        g.call("sysml_b200_mmchain_dense_allreduce", comm.handle, C.c_void_p(X.ptr), C.c_void_p(v.ptr),
               C.c_void_p(w.ptr) if ct else None, C.c_void_p(out.ptr), r1 - r0, m, ct)
        res["q%d" % it] = out.to_host().reshape(-1)
    # generic small all-reduce (colSums plan): in place
    cs = L.unaryAggregate(g, "uack+", X, "uack+")
    comm.allreduce_sum(cs.ptr, m)
    res["cs"] = cs.to_host().reshape(-1)
    np.savez(os.path.join(out_dir, "rank%d.npz" % rank), **res)
    g.sync()
    dist.barrier()
    comm.destroy()
    g.destroy()
    dist.destroy_process_group()


@pytest.mark.skipif(_ngpus() < 2, reason="needs two GPUs")
def test_fused_allreduce_mmchain_two_gpus(tmp_path):
    import torch.multiprocessing as mp
    from oracle import oracle as O
    n, m, world = 200_000, 1000, 2
    mp.spawn(_worker, args=(world, _free_port(), n, m, str(tmp_path)), nprocs=world, join=True)
    r = [np.load(tmp_path / ("rank%d.npz" % k)) for k in range(world)]
    X = O.rand_uniform(n * m, seed=1).reshape(n, m)
    w = O.rand_uniform(n, seed=4, lo=0.0, hi=1.0)
    for it, ct in enumerate((1, 0, 2, 1, 1)):
        v = O.rand_uniform(m, seed=10 + it)
        exp = O.mmchain_dense(X, v, w if ct else None, ct, O.num_threads())
        a, b = r[0]["q%d" % it], r[1]["q%d" % it]
        assert np.array_equal(a, b), "all ranks must hold bit-identical results"
        assert np.max(np.abs((a - exp) / exp)) < 1e-9
    exp_cs = O.agg_unary_dense(X, O.SUM, O.COL, 4)
    assert np.array_equal(r[0]["cs"], r[1]["cs"])
    assert np.max(np.abs(r[0]["cs"] - exp_cs)) < 1e-9 * np.max(np.abs(exp_cs)) + 1e-9
