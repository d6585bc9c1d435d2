"""bench.py's reference arm runs on host cores only: check the one-JSON-line contract and its keys on a small shape."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_one_json_line():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "2", "--warmup", "1",
                          "--rows", "40000", "--cols", "200"], capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, "exactly one line on stdout"
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["n_gpus"] == 1 and d["steps"] == 2 and d["warmup"] == 1
    for k in ("metric", "value", "unit", "ms_per_step", "higher_is_better", "scaling", "vs_baseline", "dtype", "data", "config",
              "cpu_baseline", "e2e"):
        assert k in d, k
    assert d["vs_baseline"] is None and d["dtype"] == "f64" and d["higher_is_better"] is True
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["value"] == d["value"]
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0 and d["e2e"]["value"] == d["value"]
    assert "workload" in d["config"] and d["value"] > 0


def test_reference_arm_ignores_torchrun_omp_num_threads():
    """torchrun exports OMP_NUM_THREADS=1 to every rank; the CPU arm must still use every CPU of the affinity mask
    (LibMatrixMult.matrixMultChain's k = number of virtual cores, LibMatrixMult.java:284-329), and only rank 0 prints."""
    env = dict(os.environ, RANK="0", LOCAL_RANK="0", WORLD_SIZE="2", OMP_NUM_THREADS="1", MASTER_ADDR="127.0.0.1", MASTER_PORT="29999")
    cmd = [sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "2", "--warmup", "1",
           "--rows", "40000", "--cols", "200"]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT, env=env)
    assert out.returncode == 0, out.stderr[-2000:]
    d = json.loads(out.stdout.strip())
    assert d["cpu_baseline"]["cores"] == len(os.sched_getaffinity(0))
    assert d["cpu_baseline"]["physical_cores"] >= 1 and d["n_gpus"] == 2
    assert d["config"]["reference_arm_sample_rows"] == 40000 and d["config"]["reference_arm_extrapolated"] is False
    env["RANK"] = "1"
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT, env=env)
    assert out.returncode == 0 and out.stdout.strip() == ""


def test_both_arms_print_the_same_config():
    """`config` must be identical in the b200 and reference arms for the same arguments (the driver compares them)."""
    sys.path.insert(0, ROOT)
    import bench
    import argparse
    for gpus in (1, 2, 8):
        a = argparse.Namespace(rows=10_000_000, cols=1000, gpus=gpus, collective="fused")
        c1, c2 = bench.make_config(a, gpus), bench.make_config(a, gpus)
        assert c1 == c2 and c1["rows"] == 10_000_000 and "row-sharded %d x %d rows/GPU" % (gpus, 10_000_000 // gpus) in c1["workload"]
        assert 1 <= c1["reference_arm_sample_rows"] <= 10_000_000
        assert c1["reference_arm_extrapolated"] == (c1["reference_arm_sample_rows"] < 10_000_000)
