#!/usr/bin/env python
"""bench.py -- headline benchmark of the hot path (driver contract: one JSON line on rank 0).

Workload (BASELINE.json configs[1]): the CG inner iteration of GLM / LinearRegCG on a dense FP64
X of 10,000,000 x 1000 (80 GB), row-sharded over the N GPUs of one box (strong scaling: the total is
fixed; per-GPU rows = 10M / N in whole 1000-row blocks).  One "step" = one CG iteration:

    q = t(X) %*% (w * (X %*% p))            <- mmchain XtwXv, ONE pass over the shard (the hot kernel)
    [N > 1]  all-reduce(q) over NVLink (m = 1000 doubles)
    pq = t(p) %*% q ; a = nr2 / pq ; beta += a*p ; r += a*q ; nr2' = t(r) %*% r ; p = -r + (nr2'/nr2)*p

all on the device through the C ABI of include/sysml_b200.h (no host synchronisation inside a step), as ONE call,
sysml_b200_cg_step_dense = two launches: the streaming kernel and a finishing kernel that folds the per-CTA
partials, exchanges them with the peers over NVLink peer memory and applies the m-vector tail in its last CTA.
--three-launch runs the round-1 step (mmchain+exchange, then sysml_b200_cg_vector_update), --unfused-tail the ~11
separate m-sized instructions the reference's compiler would emit, --collective nccl a separate NCCL all-reduce.
For N > 1 a `parity` block (outside the timed regions) compares the fused exchange with local mmchain + NCCL
all-reduce, checks that all ranks hold identical bits and compares a 100k-row sub-result with the CPU oracle; the
process exits non-zero when it fails.

  value   : CG iterations per second, whole job, inputs resident in HBM (CUDA events, max over ranks)
  e2e     : the same iteration driven from HOST buffers through the reference-facing call shape: every
            step copies p from pinned host memory to the device, runs mmchain (+ all-reduce), reads q back
            to the host and does the m-vector CG algebra on the host, as SystemML's CP instructions do.
            X stays resident on the device across steps, exactly as GPUObject caches it between
            instructions (GPUObject.acquireDeviceRead copies host->device only when not resident,
            GPUObject.java:528-545); e2e.cold additionally times the first-touch upload of X from pinned
            host memory on a bounded row sample.
  roofline: mmchain kernel, algorithmic bytes 8nm+16m+8n per launch / mean launch time (CUDA events around
            every launch inside the timed region) against MEASURED_PEAKS.json hbm_gbs.
  cpu_baseline / --impl reference: the CPU restatement of LibMatrixMult.matrixMultChain (oracle/) with all
            host threads on a bounded row sample (the JVM reference cannot run in this image).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

N_ROWS, N_COLS = 10_000_000, 1000
METRIC = "cg_iterations_per_sec (GLM/LinearRegCG inner iteration: mmchain XtwXv on dense FP64 10M x 1000)"
UNIT = "iters/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--rows", type=int, default=N_ROWS)
    ap.add_argument("--cols", type=int, default=N_COLS)
    ap.add_argument("--no-tsmm", action="store_true", help="skip the auxiliary tsmm measurement")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the C3 / C4 / C5 block (N = 1 only)")
    ap.add_argument("--three-launch", action="store_true", help="round-1 step: mmchain(+exchange), then the fused m-vector kernel as a third launch")
    ap.add_argument("--unfused-tail", action="store_true", help="run the m-vector CG algebra as ~11 separate instructions instead of the fused kernel")
    ap.add_argument("--collective", default="fused", choices=["fused", "nccl"],
                    help="N>1: 'fused' = sum exchange inside the mmchain finishing kernel over NVLink peer memory; 'nccl' = separate all-reduce")
    return ap.parse_args()


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def alg_bytes_mmchain(n, m, weighted=True):
    return 8.0 * n * m + 16.0 * m + (8.0 * n if weighted else 0.0)


def host_memory_bytes():
    """(MemTotal, MemAvailable) of the host, clamped by the cgroup limit when there is one."""
    tot = avail = None
    try:
        for l in open("/proc/meminfo"):
            if l.startswith("MemTotal:"):
                tot = int(l.split()[1]) * 1024
            elif l.startswith("MemAvailable:"):
                avail = int(l.split()[1]) * 1024
    except OSError:
        pass
    for f in ("/sys/fs/cgroup/memory.max", "/sys/fs/cgroup/memory/memory.limit_in_bytes"):
        try:
            v = open(f).read().strip()
            if v.isdigit() and int(v) < (1 << 60):
                tot = min(tot, int(v)) if tot else int(v)
                avail = min(avail, int(v)) if avail else int(v)
        except OSError:
            pass
    return tot or (8 << 30), avail or (4 << 30)


def reference_arm_rows(n, m):
    """Rows of X the CPU arm (--impl reference) holds in host memory: the whole matrix when it fits in 60 % of the
    host's RAM, else the largest whole number of million-row blocks that does (never more than half of what is
    available right now).  Both arms compute it the same way and print it in `config`."""
    tot, avail = host_memory_bytes()
    budget = min(0.6 * tot, 0.5 * avail if 8.0 * n * m > 0.6 * tot else 0.9 * avail)
    rows = int(budget // (8 * m))
    if rows >= n:
        return n
    if rows >= 1_000_000:
        rows -= rows % 1_000_000
    return max(min(rows, n), 1)


def make_config(args, world):
    """The `config` object: identical in the b200 and reference arms for the same command-line arguments."""
    n, m = args.rows, args.cols
    # rows of rank 0's shard (systemml_b200.sharding.row_shard; restated so the reference arm imports nothing of the product)
    nblocks = (n + 999) // 1000
    r0, r1 = 0, min((nblocks // world + (1 if nblocks % world else 0)) * 1000, n)
    ref_rows = reference_arm_rows(n, m)
    coll = "none" if world == 1 else ("fused into the mmchain finishing kernel over NVLink peer memory" if args.collective == "fused" else "NCCL")
    return {
        "workload": "GLM/LinearRegCG CG iteration, mmchain XtwXv, dense FP64 %d x %d (BASELINE configs[1]), row-sharded %d x %d rows/GPU"
                    % (n, m, world, r1 - r0),
        "rows": n, "cols": m,
        "parallelism": "row-shard x%d + all-reduce(m doubles) [%s]" % (world, coll),
        "l2": "inputs (%.1f GB per GPU) larger than the 126 MB L2; no flush needed" % (8.0 * (r1 - r0) * m / 1e9),
        "x_distribution": "U(-1,1) counter-based generator, seed 1",
        "reference_arm": "CPU restatement of LibMatrixMult.matrixMultChain (oracle/) on every host CPU of this process's affinity mask; the JVM reference cannot run in this image",
        "reference_arm_sample_rows": ref_rows,
        "reference_arm_extrapolated": ref_rows < n,
    }


# --------------------------------------------------------------------------------------------------
# clocks / throttle sampling during the timed region (pynvml, falls back to nvidia-smi)
# --------------------------------------------------------------------------------------------------
class ClockSampler:
    def __init__(self, index):
        self.index = index
        self.samples, self.reasons = [], set()
        self.max_mhz = None
        self._stop = threading.Event()
        self._t = None

    def _loop(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM)
            names = {
                getattr(pynvml, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
                getattr(pynvml, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
                getattr(pynvml, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
                getattr(pynvml, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
            }
            while not self._stop.is_set():
                self.samples.append(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM))
                try:
                    r = pynvml.nvmlDeviceGetCurrentClocksEventReasons(h)
                except Exception:
                    r = pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for bit, nm in names.items():
                    if r & bit:
                        self.reasons.add(nm)
                time.sleep(0.004)
        except Exception:
            try:
                out = subprocess.check_output(["nvidia-smi", "-i", str(self.index), "--query-gpu=clocks.sm,clocks.max.sm",
                                               "--format=csv,noheader,nounits"], text=True)
                a, b = out.strip().split(",")
                self.samples.append(int(a))
                self.max_mhz = int(b)
            except Exception:
                pass

    def start(self):
        self._t = threading.Thread(target=self._loop, daemon=True)
        self._t.start()

    def stop(self):
        self._stop.set()
        if self._t:
            self._t.join(timeout=2)
        s = sorted(self.samples)
        med = s[len(s) // 2] if s else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(s)}


# --------------------------------------------------------------------------------------------------
# CPU arm: the oracle's mmchain (restatement of LibMatrixMult.matrixMultChain) on a bounded sample
# --------------------------------------------------------------------------------------------------
def cpu_cg_iterations(n_total, m, rows, steps=None, warmup=1, target_seconds=12.0):
    """Times CG iterations (mmchain XtwXv + m-vector algebra) with the oracle on every host CPU this process may use, on
    the first `rows` rows of the same synthetic X.  Returns a dict: iters/s scaled to n_total rows, threads, physical
    cores, sample description, median seconds per sample iteration."""
    import numpy as np
    from oracle import oracle as O
    k = O.use_all_host_threads()   # not OMP_NUM_THREADS: torchrun exports 1 to every rank
    logical, physical = O.host_cpus()
    X = O.rand_uniform(rows * m, seed=1).reshape(rows, m)
    w = O.rand_uniform(rows, seed=4, lo=0.0, hi=1.0)
    p = O.rand_uniform(m, seed=2)
    r = -p.copy()
    beta = np.zeros(m)
    nr2 = float(r @ r)
    t0 = time.perf_counter()
    O.mmchain_dense(X, p, w, O.XtwXv, k)  # first touch (page faults, thread pool)
    warm = time.perf_counter() - t0
    for _ in range(max(warmup - 1, 0)):
        O.mmchain_dense(X, p, w, O.XtwXv, k)
    reps = steps if steps else max(3, min(50, int(target_seconds / max(warm, 1e-3))))
    times = []
    for _ in range(reps):
        t0 = time.perf_counter()
        q = O.mmchain_dense(X, p, w, O.XtwXv, k)
        pq = float(p @ q)
        a = nr2 / pq if pq != 0 else 0.0
        beta += a * p
        r += a * q
        nr2n = float(r @ r)
        p = -r + (nr2n / nr2) * p if nr2 != 0 else -r
        nr2 = nr2n
        times.append(time.perf_counter() - t0)
        if not np.isfinite(nr2) or nr2 == 0:
            p = O.rand_uniform(m, seed=2)
            r = -p.copy()
            nr2 = float(r @ r)
    mean = sum(times) / len(times)
    times.sort()
    med = times[len(times) // 2]
    if rows == n_total:
        sample = "all %d rows x %d cols, %d timed iterations (mean)" % (rows, m, reps)
        per_iter = mean
    else:
        sample = "first %d of %d rows x %d cols, %d timed iterations (median), scaled to the full row count" % (rows, n_total, m, reps)
        per_iter = med * (n_total / float(rows))
    return {"value": 1.0 / per_iter, "threads": k, "physical_cores": physical, "sample": sample, "sample_rows": rows,
            "seconds_per_sample_iteration": med, "sample_gbs": alg_bytes_mmchain(rows, m) / med / 1e9}


def cpu_baseline_block(c):
    return {"value": c["value"], "unit": UNIT, "cores": c["threads"], "physical_cores": c["physical_cores"],
            "threads_are": "logical CPUs in the affinity mask (SMT siblings included when physical_cores < cores)",
            "kind": "port", "sample": c["sample"], "sample_rows": c["sample_rows"], "sample_gbs": c["sample_gbs"]}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    t_start = time.perf_counter()
    cfg = make_config(args, max(args.gpus, 1))
    rows = cfg["reference_arm_sample_rows"]
    c = cpu_cg_iterations(args.rows, args.cols, rows, steps=max(args.steps, 1), warmup=max(args.warmup, 1))
    val = c["value"]
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1000.0 / val if val > 0 else None, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": cfg,
        "extrapolated": rows < args.rows, "ms_per_step_on_sample": 1000.0 * c["seconds_per_sample_iteration"],
        "cpu_baseline": cpu_baseline_block(c),
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "wall_s": time.perf_counter() - t_start,
    }
    print(json.dumps(line), flush=True)
    return 0


# --------------------------------------------------------------------------------------------------
def main():
    args = parse()
    if args.impl == "reference":
        return run_reference_arm(args)

    import ctypes as C
    import numpy as np
    import torch
    import torch.distributed as dist
    from systemml_b200 import Communicator, GPUContext, GPUObject, LibMatrixB200 as L
    from systemml_b200.gpu_context import torch_exchange
    from systemml_b200.sharding import row_shard

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a GPU: there is no CPU fallback for the product path")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # keep stdout clean for the ONE JSON line: NCCL's version/debug banner goes to stderr
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=dev)

    n, m = args.rows, args.cols
    r0, r1 = row_shard(n, world, rank)
    nloc = r1 - r0
    # all library work is queued on torch's current stream so that torch events and NCCL see it in order
    stream = torch.cuda.current_stream()
    g = GPUContext(local_rank, stream=stream.cuda_stream)
    comm = None
    if world > 1 and args.collective == "fused":
        comm = Communicator(g, rank, world, max(m, 32), torch_exchange())

    def tensor(*shape):
        return torch.empty(*shape, dtype=torch.float64, device=dev)

    def obj(t, rows, cols):
        return GPUObject.wrap(g, t.data_ptr(), rows, cols)

    tX = tensor(max(nloc, 1) * m)
    X = obj(tX, nloc, m)
    g.call("sysml_b200_rand_uniform", C.c_void_p(X.ptr), nloc * m, 1, r0 * m, -1.0, 1.0)
    tw = tensor(max(nloc, 1))
    w = obj(tw, nloc, 1)
    g.call("sysml_b200_rand_uniform", C.c_void_p(w.ptr), nloc, 4, r0, 0.0, 1.0)
    vecs = {k: tensor(m) for k in ("p", "q", "r", "beta", "tmp")}
    scal = {k: tensor(1) for k in ("pq", "a", "nr2", "nr2n", "b")}
    P, Q, R, BETA, TMP = (obj(vecs[k], m, 1) for k in ("p", "q", "r", "beta", "tmp"))
    PQ, A_, NR2, NR2N, B_ = (obj(scal[k], 1, 1) for k in ("pq", "a", "nr2", "nr2n", "b"))

    def p_(o):
        return C.c_void_p(o.ptr)

    def reset_state():
        g.call("sysml_b200_rand_uniform", p_(P), m, 2, 0, -1.0, 1.0)
        g.call("sysml_b200_binary_ms", p_(P), -1.0, p_(R), m, L.BINARY_OPS["*"], 0)  # r = -p
        g.call("sysml_b200_fill", p_(BETA), m, 0.0)
        g.call("sysml_b200_matmult_dense", p_(R), p_(R), p_(NR2), 1, m, m, 1, 0, 0)   # nr2 = t(r) %*% r

    ev_pairs = []

    two_launch = not (args.unfused_tail or args.three_launch or (world > 1 and comm is None))

    def cg_step(record):
        if record:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
        if two_launch:
            # the whole iteration in two launches: streaming kernel, then the kernel that folds the per-CTA partials,
            # exchanges them with the peers and applies the m-vector tail (programmatic dependent launches)
            g.call("sysml_b200_cg_step_dense", comm.handle if comm is not None else None, p_(X), p_(w), p_(P), p_(Q), p_(R), p_(BETA),
                   p_(NR2), nloc, m, L.XtwXv, 0.0)
            if record:
                e1.record(stream)
                ev_pairs.append((e0, e1))
            return
        if comm is not None:
            g.call("sysml_b200_mmchain_dense_allreduce", comm.handle, p_(X), p_(P), p_(w), p_(Q), nloc, m, L.XtwXv)
        else:
            g.call("sysml_b200_mmchain_dense", p_(X), p_(P), p_(w), p_(Q), nloc, m, L.XtwXv)
        if record:
            e1.record(stream)
            ev_pairs.append((e0, e1))
        if world > 1 and comm is None:
            dist.all_reduce(vecs["q"])
        if args.unfused_tail:
            g.call("sysml_b200_matmult_dense", p_(P), p_(Q), p_(PQ), 1, m, m, 1, 0, 0)               # pq = t(p) %*% q
            g.call("sysml_b200_binary_mm", p_(NR2), p_(PQ), p_(A_), 1, 1, 0, L.BINARY_OPS["/"])     # a = nr2 / pq
            g.call("sysml_b200_binary_mm", p_(P), p_(A_), p_(TMP), m, 1, L.BCAST_ROW, L.BINARY_OPS["*"])
            g.call("sysml_b200_binary_mm", p_(BETA), p_(TMP), p_(BETA), m, 1, 0, L.BINARY_OPS["+"])  # beta += a*p
            g.call("sysml_b200_binary_mm", p_(Q), p_(A_), p_(TMP), m, 1, L.BCAST_ROW, L.BINARY_OPS["*"])
            g.call("sysml_b200_binary_mm", p_(R), p_(TMP), p_(R), m, 1, 0, L.BINARY_OPS["+"])        # r += a*q
            g.call("sysml_b200_matmult_dense", p_(R), p_(R), p_(NR2N), 1, m, m, 1, 0, 0)             # nr2' = t(r) %*% r
            g.call("sysml_b200_binary_mm", p_(NR2N), p_(NR2), p_(B_), 1, 1, 0, L.BINARY_OPS["/"])   # b = nr2'/nr2
            g.call("sysml_b200_binary_mm", p_(P), p_(B_), p_(TMP), m, 1, L.BCAST_ROW, L.BINARY_OPS["*"])
            g.call("sysml_b200_binary_mm", p_(TMP), p_(R), p_(P), m, 1, 0, L.BINARY_OPS["-"])        # p = b*p - r
            g.call("sysml_b200_d2d", p_(NR2), p_(NR2N), 8)
        else:
            # q += 0*p ; a = nr2/(p.q) ; beta += a p ; r += a q ; nr2' = r.r ; p = -r + (nr2'/nr2) p : one launch
            g.call("sysml_b200_cg_vector_update", p_(P), p_(Q), p_(R), p_(BETA), p_(NR2), m, 0.0)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---------------- resident run: `value` ----------------
    reset_state()
    for _ in range(max(args.warmup, 3)):
        cg_step(False)
    reset_state()
    sampler = ClockSampler(local_rank)
    barrier()
    sampler.start()
    launches0 = g.launch_count()
    t_evt0, t_evt1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_evt0.record(stream)
    for _ in range(args.steps):
        cg_step(False)     # no events between the launches of the timed region: they would serialise the dependent launches
    t_evt1.record(stream)
    barrier()
    clocks = sampler.stop()
    launches = g.launch_count() - launches0
    total_ms = t_evt0.elapsed_time(t_evt1)
    # second timed region, same K steps, CUDA events around every hot-path call: the kernel time behind `roofline`
    reset_state()
    barrier()
    k_evt0, k_evt1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    k_evt0.record(stream)
    for _ in range(args.steps):
        cg_step(True)
    k_evt1.record(stream)
    barrier()
    instr_total_ms = k_evt0.elapsed_time(k_evt1)
    kern_ms = sum(a.elapsed_time(b) for a, b in ev_pairs) / max(len(ev_pairs), 1)
    if world > 1:
        t = torch.tensor([total_ms, kern_ms, instr_total_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms, kern_ms, instr_total_ms = float(t[0]), float(t[1]), float(t[2])
    ms_per_step = total_ms / args.steps
    instr_ms_per_step = instr_total_ms / args.steps
    value = 1000.0 / ms_per_step
    finite = bool(torch.isfinite(vecs["beta"]).all().item())

    # ---------------- parity, outside every timed region ----------------
    # (a) the fused exchange against the plain plan "local mmchain, then NCCL all-reduce" on the same p: the sum the
    #     reference trusts is RDDAggregateUtils.sumStable (J/runtime/instructions/spark/utils/RDDAggregateUtils.java:58-70);
    # (b) every rank must hold the same bits (the sharded algorithms branch on these values);
    # (c) rank 0: the sub-result of the first 100k rows of its shard against the CPU oracle, element-wise relative
    #     (GPUTests.assertEqualMatrices, src/test/java/org/apache/sysml/test/gpu/GPUTests.java:254-302).
    parity = {"tolerance_rel": 1e-9}
    g.call("sysml_b200_rand_uniform", p_(P), m, 77, 0, -1.0, 1.0)
    if world > 1:
        g.call("sysml_b200_mmchain_dense", p_(X), p_(P), p_(w), p_(Q), nloc, m, L.XtwXv)
        q_nccl = vecs["q"].clone()
        dist.all_reduce(q_nccl)
        if comm is not None:
            g.call("sysml_b200_mmchain_dense_allreduce", comm.handle, p_(X), p_(P), p_(w), p_(Q), nloc, m, L.XtwXv)
            q_fused = vecs["q"].clone()
            parity["fused_vs_nccl_max_rel"] = float(((q_fused - q_nccl).abs() / q_nccl.abs()).max().item())
        else:
            q_fused = q_nccl
            parity["fused_vs_nccl_max_rel"] = None
        gathered = [torch.empty_like(q_fused) for _ in range(world)]
        dist.all_gather(gathered, q_fused)
        parity["ranks_bit_identical"] = all(torch.equal(t.view(torch.int64), gathered[0].view(torch.int64)) for t in gathered)
    if rank == 0 and not args.no_cpu:
        from oracle import oracle as O   # the checker, never the thing measured
        rows_o = min(nloc, 100_000)
        g.call("sysml_b200_mmchain_dense", p_(X), p_(P), p_(w), p_(Q), rows_o, m, L.XtwXv)
        got = vecs["q"].cpu().numpy()
        Xh = O.rand_uniform(rows_o * m, seed=1).reshape(rows_o, m)
        exp = O.mmchain_dense(Xh, O.rand_uniform(m, seed=77), O.rand_uniform(rows_o, seed=4, lo=0.0, hi=1.0), O.XtwXv, O.use_all_host_threads())
        parity["oracle_max_rel"] = float(np.max(np.abs((got - exp) / exp)))
        parity["oracle_rows"] = rows_o
        del Xh
    ok = [parity.get("oracle_max_rel", 0.0) < 1e-9]
    if world > 1:
        ok.append(bool(parity["ranks_bit_identical"]))
        if parity["fused_vs_nccl_max_rel"] is not None:
            ok.append(parity["fused_vs_nccl_max_rel"] < 1e-12)
    parity["ok"] = all(ok)

    # ---------------- host-driven run: `e2e` ----------------
    e2e = None
    if not args.no_e2e:
        hp = torch.empty(m, dtype=torch.float64).pin_memory()
        hq = torch.empty(m, dtype=torch.float64).pin_memory()
        p_np, q_np = hp.numpy(), hq.numpy()
        rng = np.random.default_rng(2)

        def host_state():
            p_np[:] = rng.uniform(-1, 1, m)
            r = -p_np.copy()
            return r, np.zeros(m), float(r @ r)

        def e2e_step(state):
            r, beta, nr2 = state
            g.call("sysml_b200_h2d", p_(P), C.c_void_p(hp.data_ptr()), 8 * m)          # host -> device: p
            if comm is not None:
                g.call("sysml_b200_mmchain_dense_allreduce", comm.handle, p_(X), p_(P), p_(w), p_(Q), nloc, m, L.XtwXv)
            else:
                g.call("sysml_b200_mmchain_dense", p_(X), p_(P), p_(w), p_(Q), nloc, m, L.XtwXv)
                if world > 1:
                    dist.all_reduce(vecs["q"])
            g.call("sysml_b200_d2h", C.c_void_p(hq.data_ptr()), p_(Q), 8 * m)          # device -> host: q (syncs)
            pq = float(p_np @ q_np)
            a = nr2 / pq if pq != 0 else 0.0
            beta += a * p_np
            r += a * q_np
            nr2n = float(r @ r)
            p_np[:] = -r + (nr2n / nr2) * p_np
            return r, beta, nr2n

        st = host_state()
        for _ in range(max(args.warmup, 3)):
            st = e2e_step(st)
        st = host_state()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            st = e2e_step(st)
        barrier()
        e2e_ms = (time.perf_counter() - t0) * 1000.0 / args.steps
        if world > 1:
            t = torch.tensor([e2e_ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            e2e_ms = float(t[0])
        e2e = {"value": 1000.0 / e2e_ms, "unit": UNIT, "h2d_bytes_per_step": 8 * m, "d2h_bytes_per_step": 8 * m,
               "ms_per_step": e2e_ms,
               "note": "X and w resident on the device across steps (GPUObject cache semantics); p in / q out through pinned host memory every step"}
        # cold first touch on a bounded row sample: X is not on the device yet.
        #  (a) the reference's sequence: one synchronous upload from pinned host memory, then the first iteration;
        #  (b) the same rows as a SystemML binaryblock matrix (8 part files in /dev/shm, 1000 x 1000 blocks) read straight
        #      into the device matrix with the first mmchain pass overlapped with the upload (sysml_b200_binaryblock_read_mmchain)
        if rank == 0:
            rows_c = min(nloc, 500_000)
            hx = torch.empty(rows_c * m, dtype=torch.float64).pin_memory()
            hx.uniform_(-1, 1)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            g.call("sysml_b200_h2d", p_(X), C.c_void_p(hx.data_ptr()), 8 * rows_c * m)
            g.call("sysml_b200_mmchain_dense", p_(X), p_(P), p_(w), p_(Q), rows_c, m, L.XtwXv)
            g.call("sysml_b200_d2h", C.c_void_p(hq.data_ptr()), p_(Q), 8 * m)
            dt = time.perf_counter() - t0
            e2e["cold"] = {"rows": rows_c, "seconds": dt, "h2d_gbs": 8.0 * rows_c * m / dt / 1e9,
                           "iters_per_sec_if_X_uploaded_every_step": (rows_c / float(n)) / dt * world,
                           "note": "bounded sample: first-touch upload of X (pinned host -> device) + one mmchain + q read-back"}
            q_serial = hq.clone()
            try:
                import shutil
                import tempfile
                from concurrent.futures import ThreadPoolExecutor
                bb_dir = tempfile.mkdtemp(prefix="sysml_b200_bb_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
                nparts = 8
                hxn = hx.numpy().reshape(rows_c, m)
                brs = (rows_c + 999) // 1000
                per = (brs + nparts - 1) // nparts

                def write_part(pi):
                    lo, hi = pi * per * 1000, min((pi + 1) * per * 1000, rows_c)
                    if hi > lo:
                        part = hxn[lo:hi]
                        rc = g._lib.sysml_b200_binaryblock_write(os.path.join(bb_dir, "0-m-%05d" % pi).encode(), part.ctypes.data_as(C.c_void_p),
                                                                 hi - lo, m, 1000, 1000, 1, pi * per)
                        assert rc == 0
                tw = time.perf_counter()
                with ThreadPoolExecutor(nparts) as ex:
                    list(ex.map(write_part, range(nparts)))
                tw = time.perf_counter() - tw
                g.call("sysml_b200_binaryblock_read_mmchain", bb_dir.encode(), rows_c, m, 1000, 1000, 0, rows_c, p_(X), p_(P), p_(w), p_(Q), L.XtwXv)  # warm-up (file cache, rings)
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                g.call("sysml_b200_binaryblock_read_mmchain", bb_dir.encode(), rows_c, m, 1000, 1000, 0, rows_c, p_(X), p_(P), p_(w), p_(Q), L.XtwXv)
                g.call("sysml_b200_d2h", C.c_void_p(hq.data_ptr()), p_(Q), 8 * m)
                db = time.perf_counter() - t0
                rel = float(((hq - q_serial).abs() / q_serial.abs()).max())
                e2e["cold_binaryblock"] = {"rows": rows_c, "seconds": db, "gbs": 8.0 * rows_c * m / db / 1e9, "part_files": nparts,
                                           "max_rel_vs_serial_path": rel, "write_seconds": tw,
                                           "note": "same rows as a binaryblock matrix (8 part files in /dev/shm): read + byte swap + placement on the device with the first mmchain pass overlapped with the upload + q read-back"}
                shutil.rmtree(bb_dir, ignore_errors=True)
            except Exception as ex:  # noqa: BLE001 -- an auxiliary measurement must not take the headline line down
                e2e["cold_binaryblock"] = {"error": "%s: %s" % (type(ex).__name__, ex)}
            del hx
            g.call("sysml_b200_rand_uniform", C.c_void_p(X.ptr), nloc * m, 1, r0 * m, -1.0, 1.0)  # restore X

    # ---------------- e2e through instruction strings (N = 1): the call shape SystemML itself uses ----------------
    if e2e is not None and world == 1:
        from systemml_b200.execution_context import ExecutionContext, inst, mat
        torch.cuda.synchronize()
        ecx = ExecutionContext(local_rank)
        ecx.setMatrixDevice("X", X.ptr, nloc, m)
        ecx.setMatrixDevice("w", w.ptr, nloc, 1)
        ins = inst("mmchain", mat("X"), mat("p"), mat("w"), mat("q"), "XtwXv", 16)

        def ec_step(state):
            r, beta, nr2 = state
            ecx.setMatrix("p", p_np)                 # host block; uploaded when the instruction acquires it
            ecx.execute(ins)                          # GPU°mmchain°X°p°w°q°XtwXv°16 (+ sync.postProcess)
            q_h = ecx.getMatrix("q").reshape(-1)      # acquireHostRead: device -> host
            pq = float(p_np @ q_h)
            a = nr2 / pq if pq != 0 else 0.0
            beta += a * p_np
            r += a * q_h
            nr2n = float(r @ r)
            p_np[:] = -r + (nr2n / nr2) * p_np
            return r, beta, nr2n

        st = host_state()
        for _ in range(3):
            st = ec_step(st)
        st = host_state()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            st = ec_step(st)
        ms_i = (time.perf_counter() - t0) * 1000.0 / args.steps
        e2e["instruction_level"] = {"value": 1000.0 / ms_i, "unit": UNIT, "ms_per_step": ms_i,
                                    "note": "same iteration through sysml_b200_ec_execute('GPU°mmchain°X°p°w°q°XtwXv°16'): parser, GPUObject locks, lazy H2D of p, D2H of q, sync after the instruction"}
        ecx.destroy()

    # ---------------- auxiliary: tsmm FP64 TFLOP/s on this rank's shard (N = 1 only) ----------------
    tsmm = None
    if world == 1 and not args.no_tsmm:
        tout = tensor(m * m)
        g.call("sysml_b200_tsmm_dense", p_(X), C.c_void_p(tout.data_ptr()), nloc, m, 1, None)  # warm-up (+ workspace growth)
        torch.cuda.synchronize()
        reps = 5
        ts_sampler = ClockSampler(local_rank)
        ts_sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(reps):
            g.call("sysml_b200_tsmm_dense", p_(X), C.c_void_p(tout.data_ptr()), nloc, m, 1, None)
        e1.record(stream)
        torch.cuda.synchronize()
        ts_clocks = ts_sampler.stop()
        ms = e0.elapsed_time(e1) / reps
        flops = float(nloc) * m * (m + 1)
        tsmm = {"shape": [nloc, m], "ms": ms, "tflops_symmetric": flops / ms / 1e9, "tflops_gemm_equivalent": 2.0 * nloc * m * m / ms / 1e9,
                "flops_counted": "n*m*(m+1) (upper triangle incl. diagonal, as the reference computes)", "reps": reps, "clocks": ts_clocks}
        peaks = os.path.join(ROOT, "profiles", "fp64_peak.json")
        if os.path.exists(peaks):
            try:
                pk = json.load(open(peaks))
                tsmm["fp64_dmma_peak_tflops"] = pk["dmma_peak_tflops"]
                tsmm["fp64_dmma_peak_source"] = "profiles/fp64_peak.json: this repo's own DMMA issue-rate probe (tools/probe_peaks.cu), round 1; NOT measured in this run and not a MEASURED_PEAKS.json figure"
                tsmm["frac_of_measured_dmma_peak"] = tsmm["tflops_symmetric"] / pk["dmma_peak_tflops"]
            except Exception:
                pass
        del tout

    # ---------------- roofline + cpu baseline + output ----------------
    peak, peak_src = measured_peak()
    bytes_launch = alg_bytes_mmchain(nloc, m)
    achieved = bytes_launch / (kern_ms * 1e-3) / 1e9
    traffic, traffic_src = None, None
    tp = os.path.join(ROOT, "profiles", "mmchain_traffic.json")
    if os.path.exists(tp):
        try:
            tj = json.load(open(tp))
            if tj.get("rows") == nloc and tj.get("cols") == m:
                traffic = tj.get("dram_bytes_per_launch")
                traffic_src = "profiles/mmchain_traffic.json: dram__bytes_read.sum + dram__bytes_write.sum of one `ncu --set full` capture of this kernel at this shape (%s); NOT measured in this run" % tj.get("source", "committed capture")
        except Exception:
            pass
    roofline = {"bound": "hbm", "kernel": "rowstream_kernel<CHAIN> (mmchain XtwXv)" + (" + finish_exchange_kernel (fold of per-CTA partials, exchange, CG tail; a few microseconds)" if two_launch else ""),
                "timed_how": "CUDA events around every hot-path call in a second timed region of the same K steps (events inside the first one would serialise its programmatic dependent launches); kernel_share_of_step is taken within that second region",
                "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src, "algorithmic_bytes_per_launch": bytes_launch,
                "kernel_ms": kern_ms, "kernel_share_of_step": kern_ms / instr_ms_per_step, "instrumented_ms_per_step": instr_ms_per_step}
    # ---------------- BASELINE configs C3, C4, C5 at their actual sizes (N = 1): per-op rooflines + end-to-end runs ----------------
    configs = None
    if world == 1 and not args.no_configs:
        del tX, tw
        torch.cuda.empty_cache()     # the library allocates with cudaMalloc: torch must give the 80 GB back first
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import bench_configs
        dmma_peak = 37.0
        try:
            dmma_peak = float(json.load(open(os.path.join(ROOT, "profiles", "fp64_peak.json")))["dmma_peak_tflops"])
        except Exception:
            pass
        configs = bench_configs.run_configs(g, peak, dmma_peak, cpu=not args.no_cpu)
    if rank == 0:
        cpu = None
        if world == 1 and not args.no_cpu:
            # bounded sample (about 12 s of host time); the --impl reference arm holds reference_arm_sample_rows rows
            cpu = cpu_baseline_block(cpu_cg_iterations(n, m, min(n, 2_000_000)))
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": make_config(args, world), "parity": parity,
            "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu,
            "mmchain_gbs_aggregate": achieved * world, "result_finite": finite,
            "launches_per_step": 2 if two_launch else None,
        }
        if tsmm:
            line["tsmm"] = tsmm
        if configs:
            line["configs"] = configs
        print(json.dumps(line), flush=True)
    if world > 1:
        torch.cuda.synchronize()
        flag = torch.tensor([0 if parity["ok"] else 1], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MAX)
        parity["ok"] = int(flag.item()) == 0
        dist.barrier()
        if comm is not None:
            comm.destroy()
        dist.destroy_process_group()
    if not parity["ok"]:
        sys.stderr.write("bench.py: PARITY FAILED: %s\n" % json.dumps(parity))
        return 3
    return 0


if __name__ == "__main__":
    sys.exit(main())
