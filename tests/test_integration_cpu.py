"""integration/: the Java natives class and the JNI forwarders cover every export of include/sysml_b200.h with matching
arity, are up to date with the header, type-check against it, and the four patches apply to the reference tree."""
import os
import re
import shutil
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
INTEG = os.path.join(ROOT, "integration")
JNI = os.path.join(INTEG, "cpp", "systemml_b200_jni.cpp")
JAVA = os.path.join(INTEG, "java", "org", "apache", "sysml", "runtime", "matrix", "data", "LibMatrixB200.java")
sys.path.insert(0, INTEG)


def test_generated_files_are_up_to_date():
    assert subprocess.run([sys.executable, os.path.join(INTEG, "gen_bindings.py"), "--check"]).returncode == 0, \
        "run `python integration/gen_bindings.py` after changing include/sysml_b200.h"


def test_every_export_has_a_native_and_a_forwarder_of_matching_arity():
    import gen_bindings as G
    decls = G.declarations()
    assert len(decls) >= 60
    jni, java = open(JNI).read(), open(JAVA).read()
    for ret, name, params in decls:
        jname = G.camel(name)
        m = re.search(r"JNICALL %s%s\(JNIEnv\* env, jclass(.*?)\) \{" % (G.JNI_PREFIX, jname), jni)
        assert m, "no JNI forwarder for %s" % name
        n_jni = len([p for p in m.group(1).split(",") if p.strip()])
        assert n_jni == len(params), "%s: %d C parameters, %d JNI parameters" % (name, len(params), n_jni)
        mj = re.search(r"public static native \S+ %s\((.*?)\);" % jname, java)
        assert mj, "no Java native for %s" % name
        n_java = len([p for p in mj.group(1).split(",") if p.strip()])
        assert n_java == len(params), "%s: %d C parameters, %d Java parameters" % (name, len(params), n_java)
        assert "%s(" % name in jni, "%s is never called" % name
    # the host-runtime header is the C++ twin of what Java already has (parser, ExecutionContext): no natives for it
    assert "sysml_b200_ec_" not in jni


def test_forwarders_type_check_against_the_c_header():
    r = subprocess.run(["/usr/bin/g++", "-std=c++17", "-fsyntax-only", "-Wall", "-Werror", "-DSYSML_JNI_SYNTAX_CHECK",
                        "-I", os.path.join(INTEG, "cpp"), "-I", os.path.join(ROOT, "include"), JNI], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]


def test_java_sources_are_well_formed():
    """No JDK here: at least balanced braces / parentheses, a package line, and the natives the hand-written classes call exist."""
    natives = set(re.findall(r"public static native \S+ (\w+)\(", open(JAVA).read()))
    for dirpath, _, files in os.walk(os.path.join(INTEG, "java")):
        for f in files:
            src = open(os.path.join(dirpath, f)).read()
            code = re.sub(r"/\*.*?\*/|//[^\n]*|\"(?:\\.|[^\"\\])*\"", "", src, flags=re.S)
            assert code.count("{") == code.count("}") and code.count("(") == code.count(")"), f
            assert re.search(r"^package org\.apache\.sysml\.", code, flags=re.M), f
            for call in re.findall(r"LibMatrixB200\.(\w+)\(", code):
                assert call in natives or call in ("check",), "%s calls LibMatrixB200.%s which does not exist" % (f, call)


@pytest.mark.skipif(not os.path.isdir("/root/reference/src/main/java"), reason="the reference tree is only present in the build container")
def test_patches_apply_to_the_reference_tree(tmp_path):
    files = set()
    pdir = os.path.join(INTEG, "patches")
    patches = sorted(p for p in os.listdir(pdir) if p.endswith(".diff"))
    assert len(patches) == 4
    for p in patches:
        files |= set(re.findall(r"^--- a/(\S+)", open(os.path.join(pdir, p)).read(), flags=re.M))
    for f in files:
        dst = tmp_path / f
        dst.parent.mkdir(parents=True, exist_ok=True)
        shutil.copy(os.path.join("/root/reference", f), dst)
    for p in patches:
        r = subprocess.run(["patch", "-p1", "-i", os.path.join(pdir, p)], cwd=tmp_path, capture_output=True, text=True)
        assert r.returncode == 0, "%s: %s" % (p, r.stdout + r.stderr)
    parser = (tmp_path / "src/main/java/org/apache/sysml/runtime/instructions/GPUInstructionParser.java").read_text()
    assert '"mmchain"' in parser and '"uarimax"' in parser and "MMChainGPUInstruction.parseInstruction" in parser
