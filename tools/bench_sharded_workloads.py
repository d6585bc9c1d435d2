#!/usr/bin/env python
"""Strong scaling of the row-sharded workloads of BASELINE configs C3, C4, C5 on the GPUs of one box:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/bench_sharded_workloads.py
    python tools/bench_sharded_workloads.py                        (N = 1)

Every rank generates its row range of the SAME synthetic inputs (counter-based generators with a row offset), runs the
algorithm over a Communicator (peer-memory exchange, csrc/comm.cu) and rank 0 prints one JSON line: wall seconds per
workload (host clock around the call, barrier + synchronise on both sides, max over ranks), the iteration counts and a
checksum of the result, which must not depend on N beyond the summation order of the exchange.
Not part of bench.py's headline; the numbers it prints are kept under profiles/."""
import ctypes as C
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from systemml_b200 import Communicator, GPUContext, LibMatrixB200 as L  # noqa: E402
from systemml_b200.algorithms import kmeans_lloyd, l2svm, multilogreg  # noqa: E402
from systemml_b200.gpu_context import torch_exchange  # noqa: E402
from systemml_b200.sharding import row_shard  # noqa: E402

_p = lambda o: C.c_void_p(o.ptr)


def main():
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    scale = float(sys.argv[1]) if len(sys.argv) > 1 else 1.0
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    g = GPUContext(local, stream=torch.cuda.current_stream().cuda_stream)
    comm = Communicator(g, rank, world, 1 << 17, torch_exchange()) if world > 1 else None

    def timed(fn):
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        r = fn()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        dt = time.perf_counter() - t0
        if world > 1:
            t = torch.tensor([dt], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t[0])
        return r, dt

    out = {"n_gpus": world, "scale": scale}
    # ---- C3: L2SVM on CSR 50M x 10k, ~10 entries per row ----
    n, m = int(50_000_000 * scale), 10_000
    r0, r1 = row_shard(n, world, rank)
    X = g.rand_csr(r1 - r0, m, 10, 2, row_offset=r0)
    v, sc, Y = g.rand(m, 1, seed=2), g.empty(r1 - r0, 1), g.empty(r1 - r0, 1)
    g.call("sysml_b200_matmult_csr_dense", C.c_void_p(X.csr.rowPtr), C.c_void_p(X.csr.colInd), C.c_void_p(X.csr.val), _p(v), _p(sc), r1 - r0, m, m, 1, 0)
    g.call("sysml_b200_binary_ms", _p(sc), 0.0, _p(Y), r1 - r0, L.BINARY_OPS[">"], 0)
    g.call("sysml_b200_binary_ms", _p(Y), 2.0, _p(Y), r1 - r0, L.BINARY_OPS["*"], 0)
    g.call("sysml_b200_binary_ms", _p(Y), 1.0, _p(Y), r1 - r0, L.BINARY_OPS["-"], 0)
    sc.free()
    l2svm(g, X, Y, reg=1.0, tol=0.001, maxiter=1, comm=comm)
    (w, its, obj), dt = timed(lambda: l2svm(g, X, Y, reg=1.0, tol=0.001, maxiter=20, comm=comm))
    out["C3_L2SVM"] = {"rows": n, "seconds": round(dt, 4), "outer_iterations": its, "objective": obj, "checksum": float(np.sum(w))}
    for o in (X, Y, v):
        o.free()
    # ---- C5: Kmeans 10M x 100, k = 50, exactly 10 Lloyd iterations ----
    n, m, k = int(10_000_000 * scale), 100, 50
    r0, r1 = row_shard(n, world, rank)
    X = g.rand(r1 - r0, m, seed=4, offset=r0 * m, lo=-10.0, hi=10.0)
    C0 = g.rand(k, m, seed=5, lo=-10.0, hi=10.0)
    kmeans_lloyd(g, X, C0, max_iter=1, eps=0.0, comm=comm, fused="step")
    (Cn, itk, wcss, term), dt = timed(lambda: kmeans_lloyd(g, X, C0, max_iter=10, eps=0.0, comm=comm, fused="step"))
    out["C5_Kmeans"] = {"rows": n, "seconds": round(dt, 4), "iterations": itk, "ms_per_iteration": round(1e3 * dt / max(itk, 1), 3), "wcss": wcss,
                        "checksum": float(np.sum(Cn))}
    for o in (X, C0):
        o.free()
    # ---- C4: MultiLogReg 1M x 5000, K = 20, 5 outer iterations ----
    E = L.Expr
    n, m, K = int(1_000_000 * scale), 5000, 20
    r0, r1 = row_shard(n, world, rank)
    X = g.rand(r1 - r0, m, seed=3, offset=r0 * m)
    W, sc = g.rand(m, K + 1, seed=7), g.empty(r1 - r0, K + 1)
    g.call("sysml_b200_matmult_dense", _p(X), _p(W), _p(sc), r1 - r0, m, m, K + 1, 0, 0)
    lab = L.unaryAggregate(g, "uarimax", sc, "uarimax")
    ids = g.dense(np.arange(1, K + 1, dtype=np.float64).reshape(1, K))
    Yk = L.fusedCellwise(g, "spoof", [lab, ids], E.input(0).binary(E.input(1), "=="), rows=r1 - r0, cols=K)
    yl = L.matrixScalarArithmetic(g, "==", lab, float(K + 1), "==")
    for o in (W, sc):
        o.free()
    multilogreg(g, X, Yk, yl, reg=1.0, tol=1e-6, maxiter=1, comm=comm, n_total=n)
    (B, itm, objm), dt = timed(lambda: multilogreg(g, X, Yk, yl, reg=1.0, tol=1e-6, maxiter=5, comm=comm, n_total=n))
    out["C4_MultiLogReg"] = {"rows": n, "seconds": round(dt, 4), "outer_iterations": itm, "objective": objm, "checksum": float(np.sum(B))}
    if rank == 0:
        print(json.dumps(out))
    if comm is not None:
        g.sync()
        dist.barrier()
        comm.destroy()
    g.destroy()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
