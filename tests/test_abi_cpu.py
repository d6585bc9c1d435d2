"""CPU-side checks of the drop-in boundary: the shared library loads, exports every symbol that
include/sysml_b200.h declares, the ctypes table matches the header, and errors come back as codes."""
import ctypes as C
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADERS = [os.path.join(ROOT, "include", h) for h in ("sysml_b200.h", "sysml_b200_host.h")]


def _declared():
    names = set()
    for h in HEADERS:
        txt = open(h).read()
        txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
        names |= set(re.findall(r"\b(sysml_b200_\w+)\s*\(", txt))
    return sorted(names)


@pytest.fixture(scope="module")
def so_path():
    from systemml_b200 import build
    return build.build()


def test_header_declares_functions():
    names = _declared()
    assert len(names) >= 40
    for must in ("sysml_b200_mmchain_dense", "sysml_b200_tsmm_dense", "sysml_b200_matmult_dense",
                 "sysml_b200_agg_unary_dense", "sysml_b200_binary_mm", "sysml_b200_mmchain_csr"):
        assert must in names


def test_library_exports_every_declared_symbol(so_path):
    out = subprocess.check_output(["nm", "-D", "--defined-only", so_path], text=True)
    exported = set(l.split()[-1] for l in out.splitlines() if " T " in l)
    missing = [n for n in _declared() if n not in exported]
    assert not missing, "declared in the header but not exported: %s" % missing
    # nothing but the ABI is visible
    stray = [n for n in exported if not n.startswith("sysml_b200_")]
    assert not stray, "unexpected exported symbols: %s" % stray[:5]


def test_ctypes_table_matches_header(so_path):
    from systemml_b200 import native
    assert sorted(native.SIGNATURES) == _declared()
    lib = native.load()
    assert lib.sysml_b200_version().startswith(b"systemml_b200")


def test_header_is_plain_c(tmp_path):
    src = tmp_path / "t.c"
    src.write_text('#include "sysml_b200.h"\n#include "sysml_b200_host.h"\nint main(void){return SYSML_OK;}\n')
    subprocess.check_call(["/usr/bin/gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"),
                           "-c", str(src), "-o", str(tmp_path / "t.o")])


def test_errors_are_codes_not_aborts(so_path):
    """Without a GPU every entry point must fail with a code and a message (never abort / throw)."""
    from systemml_b200 import native
    lib = native.load()
    n = C.c_int(-1)
    rc = lib.sysml_b200_device_count(C.byref(n))
    if rc == 0 and n.value > 0:
        pytest.skip("a GPU is visible; the no-device error path is not reachable")
    h = C.c_void_p()
    rc = lib.sysml_b200_ctx_create(0, C.byref(h))
    assert rc < 0 and not h.value
    assert len(lib.sysml_b200_last_error(None)) > 0
    # null context -> SYSML_E_INVALID
    assert lib.sysml_b200_sync(None) == native.E_INVALID
    assert lib.sysml_b200_mmchain_dense(None, None, None, None, None, 1, 1, 0) == native.E_INVALID


def test_no_oracle_in_product():
    """The product must never import, link or call the oracle (or any CPU fallback)."""
    pkg = os.path.join(ROOT, "systemml_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                txt = open(os.path.join(dirpath, f), errors="replace").read()
                assert "oracle" not in txt.lower() or f == "rand.cu", "%s mentions the oracle" % f
    out = subprocess.check_output(["ldd", os.path.join(pkg, "libsystemml_b200.so")], text=True)
    assert "oracle" not in out
