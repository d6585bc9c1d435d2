"""Builds libsystemml_b200.so in-tree with nvcc for sm_100a (no torch, no cmake).

    python -m systemml_b200.build [--force]

The .so is git-ignored but travels to the GPU box with the gpurun snapshot.
"""
import concurrent.futures as cf
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
SO = os.path.join(HERE, "libsystemml_b200.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden", "-ccbin", "/usr/bin/g++",
]


def sources():
    return sorted(f for f in os.listdir(CSRC) if f.endswith((".cu", ".cpp")))


def _newer(a, b):
    return (not os.path.exists(b)) or os.path.getmtime(a) > os.path.getmtime(b)


def embed_ops_header():
    """cellwise_ops.cuh as a C++ raw string literal, for the NVRTC-compiled fused cellwise kernels (fused.cu)."""
    src = os.path.join(CSRC, "cellwise_ops.cuh")
    dst = os.path.join(CSRC, "cellwise_ops_embed.inc")
    text = 'R"SYSMLOPS(' + open(src).read() + ')SYSMLOPS"\n'
    if not os.path.exists(dst) or open(dst).read() != text:
        with open(dst, "w") as f:
            f.write(text)


def build(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    embed_ops_header()
    hdrs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h", ".inc"))]
    hdrs.append(os.path.join(HERE, "..", "include", "sysml_b200.h"))
    hdrs.append(os.path.join(HERE, "..", "include", "sysml_b200_host.h"))
    hdr_time = max(os.path.getmtime(h) for h in hdrs)
    jobs = []
    for s in sources():
        src = os.path.join(CSRC, s)
        obj = os.path.join(OBJ, os.path.splitext(s)[0] + ".o")
        if force or _newer(src, obj) or hdr_time > os.path.getmtime(obj):
            jobs.append((src, obj))

    def compile_one(job):
        src, obj = job
        cmd = [NVCC] + FLAGS + (["-Xptxas", "-v"] if src.endswith(".cu") else []) + ["-c", src, "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        out = r.stdout + r.stderr
        rc = r.returncode
        # ptxas occasionally falls off a cliff on the warp-specialised kernels (32 registers, kilobytes of
        # spills); a kernel that spills is a build failure, not a slow kernel
        import re as _re
        spills = []
        for l in out.splitlines():
            mm = _re.search(r"(\d+) bytes spill stores", l)
            if mm and int(mm.group(1)) > 256:  # tens of bytes outside the inner loops are harmless; kilobytes mean ptxas gave up
                spills.append(l)
        if rc == 0 and spills:
            out += "\nERROR: register spills in %s:\n%s\n" % (src, "\n".join(spills[:5]))
            rc = 1
            if os.path.exists(obj):
                os.remove(obj)
        if not verbose and rc == 0:
            out = ""
        return src, rc, out

    if jobs:
        with cf.ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            for src, rc, out in ex.map(compile_one, jobs):
                if verbose or rc:
                    sys.stderr.write(out)
                if rc:
                    raise RuntimeError("nvcc failed on %s" % src)
    objs = [os.path.join(OBJ, os.path.splitext(s)[0] + ".o") for s in sources()]
    if jobs or not os.path.exists(SO):
        cmd = [NVCC, "-shared", "-o", SO] + objs + ["-ccbin", "/usr/bin/g++", "-Xlinker", "--no-undefined", "-ldl"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode:
            sys.stderr.write(r.stdout + r.stderr)
            raise RuntimeError("link failed")
    return SO


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
