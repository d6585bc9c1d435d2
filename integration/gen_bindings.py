#!/usr/bin/env python
"""Generates the Java natives class and the JNI forwarders from include/sysml_b200.h -- one native and one forwarder
per exported function, nothing hand-maintained:

    integration/java/org/apache/sysml/runtime/matrix/data/LibMatrixB200.java
    integration/cpp/systemml_b200_jni.cpp

Convention (the reference's own native plugin: J/utils/NativeHelper.java:165-196,265-327; src/main/cpp/systemml.h:27-64,
systemml.cpp:50-124): symbols Java_<package>_<class>_<method>, status codes returned as int, Java heap arrays pinned with
GetPrimitiveArrayCritical and released with JNI_ABORT when they are inputs.  Handles (sysml_b200_ctx*, sysml_b200_comm*)
and device pointers travel as jlong.

    python integration/gen_bindings.py            # rewrites both files
    python integration/gen_bindings.py --check    # exits 1 when the committed files are stale
"""
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "sysml_b200.h")
JAVA_OUT = os.path.join(ROOT, "integration", "java", "org", "apache", "sysml", "runtime", "matrix", "data", "LibMatrixB200.java")
JNI_OUT = os.path.join(ROOT, "integration", "cpp", "systemml_b200_jni.cpp")
JNI_PREFIX = "Java_org_apache_sysml_runtime_matrix_data_LibMatrixB200_"

# host arrays that do not carry the h_ prefix: (function, parameter) -> kind
HOST_ARRAYS = {
    ("sysml_b200_fused_cellwise", "inputs"): "long_in",       # array of device pointers
    ("sysml_b200_fused_cellwise", "input_kinds"): "int_in",
    ("sysml_b200_fused_cellwise", "consts"): "double_in",
    ("sysml_b200_fused_cellwise", "program"): "int_in",
    ("sysml_b200_comm_create", "handle_out64"): "byte_out",
    ("sysml_b200_comm_connect", "all_handles"): "byte_in",
    ("sysml_b200_comm_connect_local", "comms"): "long_in",
    ("sysml_b200_device_count", "count"): "int_out",
    ("sysml_b200_timer_stop_ms", "ms"): "double_out",
    ("sysml_b200_mem_info", "free_bytes"): "long_out",
    ("sysml_b200_mem_info", "total_bytes"): "long_out",
    ("sysml_b200_csr_add", "c_nnz"): "long_out",
    ("sysml_b200_matmult_csr_csr_sparse", "c_nnz"): "long_out",
    ("sysml_b200_matmult_ultra_sparse_left", "c_nnz"): "long_out",
    ("sysml_b200_csr_scalar_op", "c_nnz"): "long_out",
    ("sysml_b200_csr_unary_op", "c_nnz"): "long_out",
    ("sysml_b200_csr_dense_op", "c_nnz"): "long_out",
    ("sysml_b200_csr_csr_op", "c_nnz"): "long_out",
    ("sysml_b200_csr_to_psell", "n_slots"): "long_out",
    ("sysml_b200_csr_to_psell", "n_groups"): "long_out",
    ("sysml_b200_mmchain_rowfn_dense", "program"): "int_in",
    ("sysml_b200_mmchain_rowfn_dense", "consts"): "double_in",
    ("sysml_b200_fused_multi_agg", "inputs"): "long_in",
    ("sysml_b200_fused_multi_agg", "input_kinds"): "int_in",
    ("sysml_b200_fused_multi_agg", "consts"): "double_in",
    ("sysml_b200_fused_multi_agg", "program"): "int_in",
    # generic byte copies: the Java side holds double[] blocks (DenseBlockDRB) -- int[] variants are written by hand below
    ("sysml_b200_binaryblock_write", "h_X"): "double_in",
    ("sysml_b200_h2d", "h_src"): "double_in",
    ("sysml_b200_d2h", "h_dst"): "double_out",
}


def declarations():
    txt = open(HEADER).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    txt = re.sub(r"#.*", "", txt)
    out = []
    for m in re.finditer(r"([\w\s\*]+?)\b(sysml_b200_\w+)\s*\(([^;{]*?)\)\s*;", txt, flags=re.S):
        ret = " ".join(m.group(1).split())
        name = m.group(2)
        params = []
        ptxt = " ".join(m.group(3).split())
        if ptxt and ptxt != "void":
            for p in ptxt.split(","):
                p = p.strip()
                pm = re.match(r"(.*?)(\w+)$", p)
                params.append((pm.group(1).strip(), pm.group(2)))
        out.append((ret, name, params))
    return out


def camel(name):
    parts = name[len("sysml_b200_"):].split("_")
    return parts[0] + "".join(x.capitalize() if not x[0].isdigit() else x for x in parts[1:])


def classify(fn, ctype, pname):
    """-> (kind, java type, jni type).  kinds: handle, handle_out, dev, dev_out, scalar, <elem>_in / <elem>_out, string"""
    t = ctype.replace("const ", "").replace(" ", "")
    if (fn, pname) in HOST_ARRAYS:
        k = HOST_ARRAYS[(fn, pname)]
        elem = k.split("_")[0]
        return k, {"long": "long[]", "int": "int[]", "double": "double[]", "byte": "byte[]"}[elem], "j%sArray" % elem
    if t in ("sysml_b200_ctx*", "sysml_b200_comm*"):
        return "handle", "long", "jlong"
    if t in ("sysml_b200_ctx**", "sysml_b200_comm**"):
        return "handle_out", "long[]", "jlongArray"
    if t.endswith("**"):
        return "dev_out", "long[]", "jlongArray"
    if t == "void*" and pname == "h_ptr":
        return "dev", "long", "jlong"                  # pinned host block from host_alloc: an address, not a Java array       # the library hands back a device (or pinned host) address
    if t == "constchar*" or t == "char*":
        return "string", "String", "jstring"
    if t.endswith("*"):
        if pname.startswith("h_"):
            if t == "void*":
                raise ValueError("untyped host buffer %s of %s: add it to HOST_ARRAYS" % (pname, fn))
            elem = {"double*": "double", "int64_t*": "long", "int32_t*": "int", "int*": "int"}[t]
            inp = ctype.strip().startswith("const")
            return "%s_%s" % (elem, "in" if inp else "out"), elem + "[]", "j%sArray" % elem
        return "dev", "long", "jlong"
    if t in ("int",):
        return "scalar", "int", "jint"
    if t in ("int64_t", "uint64_t"):
        return "scalar", "long", "jlong"
    if t == "double":
        return "scalar", "double", "jdouble"
    raise ValueError("unclassified parameter %s %s of %s" % (ctype, pname, fn))


def java_ret(ret):
    r = ret.replace("const ", "").replace(" ", "")
    return {"int": ("int", "jint"), "int64_t": ("long", "jlong"), "char*": ("String", "jstring"), "void*": ("long", "jlong")}[r]


def generate():
    decls = declarations()
    J = []
    J.append("""/* GENERATED by integration/gen_bindings.py from include/sysml_b200.h -- do not edit.
 *
 * Natives of libsystemml_b200_jni.so, one per function of the C ABI.  Takes the place of the jcuda.* calls inside
 * LibMatrixCUDA / LibMatrixCuMatMult / GPUObject / GPUMemoryManager; loaded the way NativeHelper loads its BLAS shim
 * (utils/NativeHelper.java:165-196, 296-320).  Every native returns the C status code (0 = ok); check() turns a
 * failure into the DMLRuntimeException JCudaKernels.checkResult used to throw (JCudaKernels.java:115-119).
 * Handles and device pointers are longs; arrays are Java heap arrays (pinned for the duration of the call). */
package org.apache.sysml.runtime.matrix.data;

import org.apache.sysml.runtime.DMLRuntimeException;

public class LibMatrixB200 {
	static {
		System.loadLibrary("systemml_b200_jni");
	}

	public static final int OK = 0, E_INVALID = -1, E_DIM = -2, E_OOM = -3, E_CUDA = -4, E_UNSUPPORTED = -5, E_LIMIT = -6;
	public static final int CHAIN_XtXv = 0, CHAIN_XtwXv = 1, CHAIN_XtXvy = 2;
	public static final int AGG_SUM = 0, AGG_SUMSQ = 1, AGG_MAX = 2, AGG_MIN = 3, AGG_MAXINDEX = 4, AGG_MININDEX = 5, AGG_MEAN = 6, AGG_PROD = 7, AGG_VAR = 8;
	public static final int DIR_ALL = 0, DIR_ROW = 1, DIR_COL = 2;
	public static final int BCAST_NONE = 0, BCAST_COL = 1, BCAST_ROW = 2, BCAST_OUTER = 3;

	public static void check(long ctx, int rc) {
		if( rc != OK )
			throw new DMLRuntimeException("GPU : " + lastError(ctx) + " (code " + rc + ")");
	}
""")
    C = []
    C.append("""// GENERATED by integration/gen_bindings.py from include/sysml_b200.h -- do not edit.
//
// JNI forwarders of libsystemml_b200_jni.so: one per function of the C ABI, no arithmetic, no allocation.  Built with the
// JDK's <jni.h> and linked against libsystemml_b200.so; same conventions as src/main/cpp/systemml.cpp:50-124
// (GetPrimitiveArrayCritical pins, JNI_ABORT for inputs, status code returned to Java).
#ifdef SYSML_JNI_SYNTAX_CHECK
#include "jni_min.h"  // stand-in with the handful of JNI declarations used here, for images without a JDK
#else
#include <jni.h>
#endif
#include <stdint.h>

#include "sysml_b200.h"

namespace {
template <class T>
inline T* ptr(jlong h) { return reinterpret_cast<T*>(static_cast<intptr_t>(h)); }
// pins a Java primitive array for the duration of a call; mode JNI_ABORT for inputs (nothing to copy back), 0 for outputs
struct Pin {
  JNIEnv* env;
  jarray arr;
  void* p;
  jint mode;
  Pin(JNIEnv* e, jarray a, jint m) : env(e), arr(a), p(a ? e->GetPrimitiveArrayCritical(a, nullptr) : nullptr), mode(m) {}
  ~Pin() { if (p) env->ReleasePrimitiveArrayCritical(arr, p, mode); }
};
}  // namespace

extern "C" {
""")
    for ret, name, params in decls:
        jname = camel(name)
        jret, cret = java_ret(ret)
        jparams, cparams, decl, pre, args, post = [], [], [], [], [], []
        for ctype, pname in params:
            kind, jt, ct = classify(name, ctype, pname)
            jparams.append("%s %s" % (jt, pname))
            cparams.append("%s %s" % (ct, pname))
            base = ctype.replace("const ", "").strip()
            if kind == "handle":
                args.append("ptr<%s>(%s)" % (base.rstrip("*").strip(), pname))
            elif kind == "handle_out":
                inner = base[:-1].strip()  # T*
                decl.append("  %s %s_v = nullptr;" % (inner, pname))
                args.append("&%s_v" % pname)
                post.append("  if (%s) { jlong v = (jlong)(intptr_t)%s_v; env->SetLongArrayRegion(%s, 0, 1, &v); }" % (pname, pname, pname))
            elif kind == "dev_out":
                inner = base[:-1].strip()
                decl.append("  %s %s_v = nullptr;" % (inner, pname))
                args.append("&%s_v" % pname)
                post.append("  if (%s) { jlong v = (jlong)(intptr_t)%s_v; env->SetLongArrayRegion(%s, 0, 1, &v); }" % (pname, pname, pname))
            elif kind == "dev":
                args.append("(%s)(intptr_t)%s" % (ctype.strip(), pname))
            elif kind == "scalar":
                args.append("(%s)%s" % (base, pname))
            elif kind == "string":
                if "const" in ctype:
                    # a JNI call: must happen before any array is pinned (no JNI calls inside a critical region)
                    decl.append("  const char* %s_c = %s ? env->GetStringUTFChars(%s, nullptr) : nullptr;" % (pname, pname, pname))
                    args.append("%s_c" % pname)
                    post.append("  if (%s_c) env->ReleaseStringUTFChars(%s, %s_c);" % (pname, pname, pname))
                else:
                    raise ValueError("output string parameter %s of %s" % (pname, name))
            elif kind == "long_in" and (name, pname) in (("sysml_b200_fused_cellwise", "inputs"), ("sysml_b200_comm_connect_local", "comms")):
                # jlong[] of addresses -> array of pointers (both are 8 bytes on every platform this library targets)
                pre.append("  static_assert(sizeof(jlong) == sizeof(void*), \"64-bit only\");")
                pre.append("  Pin %s_p(env, %s, JNI_ABORT);" % (pname, pname))
                args.append("(%s)%s_p.p" % (ctype.strip(), pname))
            elif kind.endswith("_in"):
                pre.append("  Pin %s_p(env, %s, JNI_ABORT);" % (pname, pname))
                args.append("(%s)%s_p.p" % (ctype.strip(), pname))
            elif kind.endswith("_out"):
                if base in ("int64_t*",) and kind == "long_out":
                    # int64_t and jlong are both 64-bit signed
                    pre.append("  Pin %s_p(env, %s, 0);" % (pname, pname))
                    args.append("(int64_t*)%s_p.p" % pname)
                else:
                    pre.append("  Pin %s_p(env, %s, 0);" % (pname, pname))
                    args.append("(%s)%s_p.p" % (ctype.strip(), pname))
            else:
                raise ValueError(kind)
        J.append("\tpublic static native %s %s(%s);" % (jret, jname, ", ".join(jparams)))
        sig = ", ".join(["JNIEnv* env", "jclass"] + cparams)
        C.append("JNIEXPORT %s JNICALL %s%s(%s) {" % (cret, JNI_PREFIX, jname, sig))
        C.append("  (void)env;")
        call = "%s(%s)" % (name, ", ".join(args))
        if jret == "String":
            C.extend(decl)
            C.extend(pre)
            C.append("  const char* s = %s;" % call)
            C.extend(post)
            C.append("  return env->NewStringUTF(s ? s : \"\");")
        elif cret == "jlong" and ret.replace(" ", "").endswith("void*"):
            C.extend(pre)
            C.append("  jlong r = (jlong)(intptr_t)%s;" % call)
            C.extend(post)
            C.append("  return r;")
        else:
            if pre or post or decl:
                C.append("  %s r;" % cret)
                C.extend(decl)
                C.append("  {")
                C.extend("  " + l for l in pre)
                C.append("    r = (%s)%s;" % (cret, call))
                C.append("  }  // pins released before any JNI call below")
                C.extend(post)
                C.append("  return r;")
            else:
                C.append("  return (%s)%s;" % (cret, call))
        C.append("}")
        C.append("")
    # int[] variants of the generic copies (CSR row pointers / column indices live in int[] on the Java side)
    J.append("\tpublic static native int h2dInts(long ctx, long d_dst, int[] h_src, long bytes);")
    J.append("\tpublic static native int d2hInts(long ctx, int[] h_dst, long d_src, long bytes);")
    C.append("""JNIEXPORT jint JNICALL %sh2dInts(JNIEnv* env, jclass, jlong ctx, jlong d_dst, jintArray h_src, jlong bytes) {
  Pin p(env, h_src, JNI_ABORT);
  return (jint)sysml_b200_h2d(ptr<sysml_b200_ctx>(ctx), (void*)(intptr_t)d_dst, p.p, (int64_t)bytes);
}

JNIEXPORT jint JNICALL %sd2hInts(JNIEnv* env, jclass, jlong ctx, jintArray h_dst, jlong d_src, jlong bytes) {
  Pin p(env, h_dst, 0);
  return (jint)sysml_b200_d2h(ptr<sysml_b200_ctx>(ctx), p.p, (const void*)(intptr_t)d_src, (int64_t)bytes);
}
""" % (JNI_PREFIX, JNI_PREFIX))
    J.append("}")
    C.append("}  // extern \"C\"")
    return "\n".join(J) + "\n", "\n".join(C) + "\n"


if __name__ == "__main__":
    java, jni = generate()
    if "--check" in sys.argv:
        ok = os.path.exists(JAVA_OUT) and os.path.exists(JNI_OUT) and open(JAVA_OUT).read() == java and open(JNI_OUT).read() == jni
        sys.exit(0 if ok else 1)
    os.makedirs(os.path.dirname(JAVA_OUT), exist_ok=True)
    os.makedirs(os.path.dirname(JNI_OUT), exist_ok=True)
    open(JAVA_OUT, "w").write(java)
    open(JNI_OUT, "w").write(jni)
    print("wrote", JAVA_OUT, "and", JNI_OUT)
