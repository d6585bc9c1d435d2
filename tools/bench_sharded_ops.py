#!/usr/bin/env python
"""Strong scaling of the other row-sharded plans of SURVEY 8e on N GPUs of one node (one process per GPU):
tsmm t(X)%*%X (10M x 1000 total), t(P)%*%X (10M x 50 / 10M x 100), colSums (10M x 1000): local op on the row shard,
then ONE exchange through the peer-memory all-reduce (sysml_b200_allreduce).  CUDA events, max over ranks.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/bench_sharded_ops.py
"""
import ctypes as C
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from systemml_b200 import Communicator, GPUContext  # noqa: E402
from systemml_b200.gpu_context import torch_exchange  # noqa: E402
from systemml_b200.sharding import row_shard  # noqa: E402

rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", 0))
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
os.environ.setdefault("MASTER_PORT", "29533")
dist.init_process_group("gloo", rank=rank, world_size=world)
g = GPUContext(local)
comm = Communicator(g, rank, world, 1 << 20, torch_exchange())
p_ = lambda o: C.c_void_p(o.ptr)
n, m = int(os.environ.get("ROWS", 10_000_000)), 1000
r0, r1 = row_shard(n, world, rank)
nl = r1 - r0
X = g.rand(nl, m, seed=1, offset=r0 * m)
out = g.empty(m, m)
cs = g.empty(1, m)


def timed(fn, reps=5):
    fn()
    g.sync()
    dist.barrier()
    best = 1e30
    for _ in range(reps):
        g.timer_start()
        fn()
        ms = g.timer_stop_ms()
        t = torch.tensor([ms], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        best = min(best, float(t[0]))
    return best


def tsmm():
    g.call("sysml_b200_tsmm_dense", p_(X), p_(out), nl, m, 1, None)
    comm.allreduce(out.ptr, m * m, "sum")


def colsums():
    g.call("sysml_b200_agg_unary_dense", p_(X), nl, m, 0, 2, p_(cs))
    comm.allreduce(cs.ptr, m, "sum")


res = {"n_gpus": world, "rows_total": n}
t = timed(tsmm, 3)
res["tsmm_10Mx1000_ms"] = round(t, 3)
res["tsmm_tflops_symmetric_aggregate"] = round(n * m * (m + 1) / t / 1e9, 2)
t = timed(colsums)
res["colSums_10Mx1000_ms"] = round(t, 4)
res["colSums_GBs_aggregate"] = round(8.0 * n * m / t / 1e6, 1)
X.free()
out.free()
k, f = 50, 100
r0, r1 = row_shard(n, world, rank)
P, Xk, Ck = g.rand(nl, k, seed=2, offset=r0 * k), g.rand(nl, f, seed=3, offset=r0 * f), g.empty(k, f)


def tpx():
    g.call("sysml_b200_matmult_dense", p_(P), p_(Xk), p_(Ck), nl, k, nl, f, 1, 0)
    comm.allreduce(Ck.ptr, k * f, "sum")


t = timed(tpx)
res["tPX_50x10Mx100_ms"] = round(t, 4)
if rank == 0:
    print(json.dumps(res), flush=True)
g.sync()
dist.barrier()
comm.destroy()
g.destroy()
dist.destroy_process_group()
