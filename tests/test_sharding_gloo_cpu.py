"""Host-side logic of the multi-GPU plan on CPU: row-block sharding + all-reduce of the per-shard
partial results (world_size 2, gloo).  The per-shard compute stands in through the oracle, which is
allowed in tests; what is checked is the partitioning, the exchange and that shard partials add up to the
single-device result (mmchain / t(X)y / colSums / tsmm need ONE sum exchange, MV /
rowSums need none)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from systemml_b200.sharding import all_shards, row_shard


def test_row_shard_covers_all_rows_in_whole_blocks():
    for n in (0, 1, 999, 1000, 1001, 10_000_000, 12_345_678):
        for world in (1, 2, 3, 4, 8):
            sh = all_shards(n, world)
            assert sh[0][0] == 0 and sh[-1][1] == n
            for (a, b), (c, d) in zip(sh, sh[1:]):
                assert b == c and a <= b
            for a, b in sh:
                if b > a:  # ranks beyond the last block get an empty range at the end
                    assert a % 1000 == 0 and (b % 1000 == 0 or b == n)
            sizes = [b - a for a, b in sh]
            assert max(sizes) - min(sizes) <= 2000  # one block of imbalance + the ragged last block
    assert row_shard(10_000_000, 8, 3) == (3_750_000, 5_000_000)
    with pytest.raises(ValueError):
        row_shard(10, 2, 2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n, m, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import oracle as O
    r0, r1 = row_shard(n, world, rank)
    # every rank regenerates exactly its rows of the global matrix (counter-based generator)
    X = O.rand_uniform((r1 - r0) * m, seed=1, offset=r0 * m).reshape(r1 - r0, m)
    w = O.rand_uniform(r1 - r0, seed=4, offset=r0, lo=0.0, hi=1.0)
    v = O.rand_uniform(m, seed=2)
    # mmchain: local partial + ONE all-reduce of m doubles
    q = torch.from_numpy(O.mmchain_dense(X, v, w, O.XtwXv, 1))
    dist.all_reduce(q)
    # colSums: same exchange; rowSums / X %*% v: no exchange, output stays row-sharded
    cs = torch.from_numpy(O.agg_unary_dense(X, O.SUM, O.COL, 1))
    dist.all_reduce(cs)
    mv = O.matmult_dense(X, v.reshape(m, 1), 1)[0].reshape(-1)
    # tsmm: all-reduce of the m x m partial
    T = torch.from_numpy(O.tsmm_dense(X, True, 1)[0])
    dist.all_reduce(T)
    shards = all_shards(n, world)
    pad = max(b - a for a, b in shards)  # gloo all_gather wants equal sizes: pad the ragged shard
    mine = torch.zeros(pad, dtype=torch.float64)
    mine[: r1 - r0] = torch.from_numpy(mv)
    gathered = [torch.zeros(pad, dtype=torch.float64) for _ in shards]
    dist.all_gather(gathered, mine)
    if rank == 0:
        full = torch.cat([g[: b - a] for g, (a, b) in zip(gathered, shards)])
        np.savez(os.path.join(out_dir, "res.npz"), q=q.numpy(), cs=cs.numpy(), T=T.numpy(), mv=full.numpy())
    dist.destroy_process_group()


def test_two_rank_shards_add_up_to_the_single_device_result(tmp_path):
    from oracle import oracle as O
    n, m, world = 5300, 37, 2
    port = _free_port()
    mp.spawn(_worker, args=(world, port, n, m, str(tmp_path)), nprocs=world, join=True)
    res = np.load(tmp_path / "res.npz")
    X = O.rand_uniform(n * m, seed=1).reshape(n, m)
    w = O.rand_uniform(n, seed=4, lo=0.0, hi=1.0)
    v = O.rand_uniform(m, seed=2)
    q = O.mmchain_dense(X, v, w, O.XtwXv, 1)
    assert np.max(np.abs((res["q"] - q) / q)) < 1e-12
    assert np.max(np.abs(res["cs"] - O.agg_unary_dense(X, O.SUM, O.COL, 1))) < 1e-10
    T, _ = O.tsmm_dense(X, True, 1)
    assert np.max(np.abs((res["T"] - T) / T)) < 1e-12
    assert np.array_equal(res["mv"], O.matmult_dense(X, v.reshape(m, 1), 1)[0].reshape(-1))
