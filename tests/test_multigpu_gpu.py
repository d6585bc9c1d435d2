"""Two, four and eight real GPUs (whichever the box has): row-sharded mmchain with the sum exchange fused into the finishing kernel (NVLink peer
memory, csrc/comm.cu) against the single-device result and the CPU oracle, plus the other sharded plans of
SURVEY 8e (tsmm, t(P)%*%X, colSums, colMaxs/colMins, sum) as local op + one exchange.  World sizes the box cannot hold are skipped."""
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


def _multilogreg_problem():
    rng = np.random.default_rng(3)
    N, D, K = 24_000, 40, 4
    X = rng.uniform(-1, 1, (N, D))
    labels = np.argmax(X @ rng.normal(size=(D, K + 1)) + 0.5 * rng.gumbel(size=(N, K + 1)), axis=1)
    Yfull = np.eye(K + 1)[labels]
    return X, np.ascontiguousarray(Yfull[:, :K]), np.ascontiguousarray(Yfull[:, K:])


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
    # the other sharded plans of SURVEY 8e -- each: local op on the row shard, then ONE exchange
    comm2 = Communicator(g, rank, world, 1 << 16, torch_exchange())
    Xs = g.rand(r1 - r0, 160, seed=21, offset=r0 * 160)                 # narrower operand for tsmm / t(P)%*%X
    t, _ = L.matmultTSMM(g, "tsmm", Xs, True, with_nnz=True)            # tsmm LEFT: sum of the shards' m x m
    comm2.allreduce(t.ptr, 160 * 160, "sum")
    res["tsmm"] = t.to_host()
    Pm = g.rand(r1 - r0, 24, seed=22, offset=r0 * 24)                   # t(P) %*% X (Kmeans centroid update)
    tpx = L.matmult(g, "ba+*", Pm, Xs, isLeftTransposed=True)
    comm2.allreduce(tpx.ptr, 24 * 160, "sum")
    res["tpx"] = tpx.to_host()
    for opc, rop in (("uacmax", "max"), ("uacmin", "min")):           # colMaxs / colMins: exact
        o = L.unaryAggregate(g, opc, Xs, opc)
        comm2.allreduce(o.ptr, 160, rop)
        res[opc] = o.to_host().reshape(-1)
    sm = g.empty(1, 1)                                                  # sum(X): all-reduce of one value
    g.call("sysml_b200_agg_unary_dense", C.c_void_p(Xs.ptr), r1 - r0, 160, 0, 0, C.c_void_p(sm.ptr))
    comm2.allreduce(sm.ptr, 1, "sum")
    res["sum"] = sm.to_host().reshape(-1)
    # the workloads themselves, row-sharded: LinearRegCG and GLM Poisson/log (north_star: end to end on 1-8 GPUs)
    from systemml_b200.algorithms import glm_poisson_log, linreg_cg
    na, ma = 60_000, 200
    a0, a1 = row_shard(na, world, rank)
    Xa = g.rand(a1 - a0, ma, seed=40, offset=a0 * ma)
    wt = g.rand(ma, 1, seed=41)
    ya = g.empty(a1 - a0, 1)
    g.call("sysml_b200_matmult_dense", C.c_void_p(Xa.ptr), C.c_void_p(wt.ptr), C.c_void_p(ya.ptr), a1 - a0, ma, ma, 1, 0, 0)
    nz = g.rand(a1 - a0, 1, seed=42, offset=a0)
    g.call("sysml_b200_axpy", C.c_void_p(ya.ptr), C.c_void_p(nz.ptr), 0.1, C.c_void_p(ya.ptr), a1 - a0, 1, 0)
    beta, its = linreg_cg(g, Xa, ya, reg=0.01, tol=1e-4, comm=comm)
    res["linreg_beta"], res["linreg_its"] = beta, np.array([its])
    bt = g.rand(ma, 1, seed=43, lo=-0.05, hi=0.05)
    g.call("sysml_b200_matmult_dense", C.c_void_p(Xa.ptr), C.c_void_p(bt.ptr), C.c_void_p(ya.ptr), a1 - a0, ma, ma, 1, 0, 0)
    g.call("sysml_b200_unary", C.c_void_p(ya.ptr), C.c_void_p(ya.ptr), a1 - a0, 0)      # y = exp(X b): noiseless Poisson means
    gb, inner = glm_poisson_log(g, Xa, ya, reg=0.0, outer_iters=4, inner_iters=20, inner_tol=1e-8, comm=comm)
    res["glm_beta"], res["glm_inner"] = gb, np.array(inner)
    # Kmeans, row-sharded: the one-pass Lloyd step per rank + one exchange of the k x m sums, the counts and sum(minD)
    from systemml_b200.algorithms import kmeans_lloyd
    nk, mk, kk = 40_000, 20, 6
    k0, k1 = row_shard(nk, world, rank)
    Xk = g.rand(k1 - k0, mk, seed=50, offset=k0 * mk, lo=-10.0, hi=10.0)
    C0 = g.rand(kk, mk, seed=51, lo=-10.0, hi=10.0)
    Ck, itk, wk, tk = kmeans_lloyd(g, Xk, C0, max_iter=8, eps=1e-9, comm=comm2)
    res["kmeans_C"], res["kmeans_meta"] = Ck, np.array([itk, wk, tk])
    # L2SVM on a row-sharded CSR matrix: X %*% s local, t(X) %*% (out*Y) and the n-vector sums exchanged
    from systemml_b200.algorithms import l2svm
    ns, ms = 30_000, 300
    s0, s1 = row_shard(ns, world, rank)
    Xs_csr = g.rand_csr(s1 - s0, ms, 12, 60, row_offset=s0)
    vs = g.rand(ms, 1, seed=61)
    sc = g.empty(s1 - s0, 1)
    g.call("sysml_b200_matmult_csr_dense", C.c_void_p(Xs_csr.csr.rowPtr), C.c_void_p(Xs_csr.csr.colInd), C.c_void_p(Xs_csr.csr.val),
           C.c_void_p(vs.ptr), C.c_void_p(sc.ptr), s1 - s0, ms, ms, 1, 0)
    Ys = g.dense(np.where(sc.to_host() > 0, 1.0, -1.0))
    wsv, its_s, obj_s = l2svm(g, Xs_csr, Ys, reg=1.0, tol=0.001, maxiter=8, comm=comm2)
    res["l2svm_w"], res["l2svm_meta"] = wsv, np.array([its_s, obj_s])
    # MultiLogReg, row-sharded: X %*% V and the row-wise chains local, t(X) %*% (N x K) and the N-row sums exchanged
    from systemml_b200.algorithms import multilogreg
    Xm, Ym, ylm = _multilogreg_problem()
    m0, m1 = row_shard(Xm.shape[0], world, rank)
    Bm, itm, objm = multilogreg(g, g.dense(Xm[m0:m1]), g.dense(Ym[m0:m1]), g.dense(ylm[m0:m1]), reg=1.0, tol=1e-6, maxiter=6, comm=comm2,
                                n_total=Xm.shape[0])
    res["mlr_B"], res["mlr_meta"] = Bm, np.array([itm, objm])
    big = g.rand(200_000, 1, seed=30 + rank)                            # longer than the mailbox: goes in pieces
    comm2.allreduce(big.ptr, 200_000, "sum")
    res["big"] = big.to_host().reshape(-1)
    g.sync()
    dist.barrier()
    comm2.destroy()
    np.savez(os.path.join(out_dir, "rank%d.npz" % rank), **res)
    g.sync()
    dist.barrier()
    comm.destroy()
    g.destroy()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4, 8])
def test_fused_allreduce_mmchain_multi_gpu(tmp_path, world):
    if _ngpus() < world:
        pytest.skip("needs %d GPUs" % world)
    import torch.multiprocessing as mp
    from oracle import oracle as O
    n, m = 200_000, 1000
    mp.spawn(_worker, args=(world, _free_port(), n, m, str(tmp_path)), nprocs=world, join=True)
    r = [np.load(tmp_path / ("rank%d.npz" % k)) for k in range(world)]
    X = O.rand_uniform(n * m, seed=1).reshape(n, m)
    w = O.rand_uniform(n, seed=4, lo=0.0, hi=1.0)
    for it, ct in enumerate((1, 0, 2, 1, 1)):
        v = O.rand_uniform(m, seed=10 + it)
        exp = O.mmchain_dense(X, v, w if ct else None, ct, O.num_threads())
        a = r[0]["q%d" % it]
        for k in range(1, world):
            assert np.array_equal(a, r[k]["q%d" % it]), "all ranks must hold bit-identical results (rank %d)" % k
        rel = float(np.max(np.abs((a - exp) / exp)))
        print("world %d mmchain ct %d: max element-wise relative error %.3e" % (world, ct, rel))
        assert rel < 1e-9
    Xs = O.rand_uniform(n * 160, seed=21).reshape(n, 160)
    Pm = O.rand_uniform(n * 24, seed=22).reshape(n, 24)
    for key in ("tsmm", "tpx", "uacmax", "uacmin", "sum", "big", "cs", "linreg_beta", "glm_beta", "glm_inner", "kmeans_C", "kmeans_meta", "l2svm_w", "l2svm_meta", "mlr_B", "mlr_meta"):
        for k in range(1, world):
            assert np.array_equal(r[0][key], r[k][key]), "%s differs on rank %d" % (key, k)
    exp_t, _ = O.tsmm_dense(Xs, True, O.num_threads())
    assert np.max(np.abs(r[0]["tsmm"] - exp_t)) < 1e-9 * np.max(np.abs(exp_t))
    exp_tpx = (Pm.astype(np.longdouble).T @ Xs.astype(np.longdouble)).astype(np.float64)
    assert np.max(np.abs(r[0]["tpx"] - exp_tpx)) < 1e-9 * np.max(np.abs(exp_tpx))
    assert np.array_equal(r[0]["uacmax"], Xs.max(axis=0)) and np.array_equal(r[0]["uacmin"], Xs.min(axis=0))
    assert abs(r[0]["sum"][0] - float(np.sum(Xs.astype(np.longdouble)))) < 1e-9 * n
    big = O.rand_uniform(200_000, seed=30)
    for k in range(1, world):   # the exchange adds in rank order
        big = big + O.rand_uniform(200_000, seed=30 + k)
    assert np.array_equal(r[0]["big"], big)
    # sharded workloads: identical on both ranks, equal to the single-process CPU twins (oracle kernels)
    from oracle.algorithms_cpu import cpu_glm_poisson_log, cpu_linreg_cg
    na, ma = 60_000, 200
    Xa = O.rand_uniform(na * ma, seed=40).reshape(na, ma)
    ya = Xa @ O.rand_uniform(ma, seed=41) + 0.1 * O.rand_uniform(na, seed=42)
    eb, eits = cpu_linreg_cg(Xa, ya, 0.01, 1e-4, 0, O.num_threads())
    assert int(r[0]["linreg_its"][0]) == eits
    assert_close = lambda e, a, rel: np.max(np.abs(e - a)) <= rel * np.max(np.abs(e))
    assert assert_close(eb, r[0]["linreg_beta"], 1e-9)
    yg = np.exp(Xa @ O.rand_uniform(ma, seed=43, lo=-0.05, hi=0.05))
    egb = cpu_glm_poisson_log(Xa, yg, 0.0, 4, 20, 1e-8, O.num_threads())
    assert assert_close(egb, r[0]["glm_beta"], 1e-9)
    from oracle.algorithms_cpu import cpu_multilogreg
    Xm, Ym, ylm = _multilogreg_problem()
    eB, eitm, eobjm = cpu_multilogreg(Xm, Ym, ylm, 1.0, 1e-6, 6, 0)
    assert int(r[0]["mlr_meta"][0]) == eitm and abs(r[0]["mlr_meta"][1] - eobjm) <= 1e-9 * abs(eobjm)
    assert assert_close(eB, r[0]["mlr_B"], 1e-9)
    from oracle.algorithms_cpu import cpu_l2svm
    rp, ci, va = O.rand_csr(30_000, 300, 12, 60)
    Xd = np.zeros((30_000, 300))
    for i in range(30_000):
        Xd[i, ci[rp[i]:rp[i + 1]]] = va[rp[i]:rp[i + 1]]
    Yd = np.where(Xd @ O.rand_uniform(300, seed=61) > 0, 1.0, -1.0)
    ew, eits, eobj = cpu_l2svm(Xd, Yd, 1.0, 0.001, 8)
    assert int(r[0]["l2svm_meta"][0]) == eits and abs(r[0]["l2svm_meta"][1] - eobj) <= 1e-9 * abs(eobj)
    assert assert_close(ew, r[0]["l2svm_w"], 1e-9)
    from oracle.algorithms_cpu import cpu_kmeans_lloyd
    Xk = O.rand_uniform(40_000 * 20, seed=50, lo=-10.0, hi=10.0).reshape(40_000, 20)
    Ck0 = O.rand_uniform(6 * 20, seed=51, lo=-10.0, hi=10.0).reshape(6, 20)
    eC, eit, ew, et = cpu_kmeans_lloyd(Xk, Ck0, 8, 1e-9)
    assert int(r[0]["kmeans_meta"][0]) == eit and int(r[0]["kmeans_meta"][2]) == et
    assert assert_close(eC, r[0]["kmeans_C"], 1e-9) and abs(r[0]["kmeans_meta"][1] - ew) <= 1e-9 * abs(ew)
    exp_cs = O.agg_unary_dense(X, O.SUM, O.COL, 4)
    assert np.max(np.abs(r[0]["cs"] - exp_cs)) < 1e-9 * np.max(np.abs(exp_cs)) + 1e-9
