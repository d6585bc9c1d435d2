"""binaryblock files -> device matrix (csrc/binaryblock.cu) against the host matrix they were written from: all four block
formats, ragged last blocks, row ranges (a rank's shard), a directory of part files, blocks missing from the file, and the
upload with the first mmchain pass overlapped against mmchain on the resident matrix and the CPU oracle
(ReaderBinaryBlock.java:124-207 + LibMatrixMult.matrixMultChain, LibMatrixMult.java:243-329)."""
import ctypes as C
import os

import numpy as np
import pytest

from oracle import oracle as O
from tests.util import assert_equal_matrices

pytestmark = pytest.mark.gpu


def _write(path, X, br, bc, dense_only=0, first_block_row=0):
    from systemml_b200 import native
    X = np.ascontiguousarray(X)
    rc = native.load().sysml_b200_binaryblock_write(str(path).encode(), X.ctypes.data_as(C.c_void_p), X.shape[0], X.shape[1], br, bc, dense_only,
                                                    first_block_row)
    assert rc == 0


def _matrix(rows, cols, seed):
    rng = np.random.default_rng(seed)
    X = rng.uniform(-1, 1, (rows, cols))
    X[: rows // 5, : cols // 3] = 0.0                                        # empty blocks
    m = slice(rows // 5, 2 * rows // 5)
    X[m, :] *= rng.uniform(size=X[m, :].shape) < 0.03                        # sparse blocks
    X[2 * rows // 5: 3 * rows // 5, cols // 2:] = 0.0
    X[2 * rows // 5 + 3, cols // 2 + 1] = -4.25                              # ultra-sparse
    return X


@pytest.mark.parametrize("rows,cols,br,bc", [(2300, 1700, 1000, 1000), (530, 90, 100, 64), (1000, 1000, 1000, 1000), (77, 5, 1000, 1000)])
def test_read_dense_all_formats_and_row_ranges(gctx, tmp_path, rows, cols, br, bc):
    X = _matrix(rows, cols, rows + cols)
    for dense_only in (0, 1):
        p = tmp_path / ("X%d.bb" % dense_only)
        _write(p, X, br, bc, dense_only)
        d = gctx.empty(rows, cols)
        gctx.call("sysml_b200_binaryblock_read_dense", str(p).encode(), rows, cols, br, bc, 0, rows, C.c_void_p(d.ptr))
        assert np.array_equal(d.to_host(), X), "whole matrix, dense_only=%d" % dense_only
        d.free()
    # a rank's shard: whole blocks from a block boundary, possibly ending inside a block
    for r0, r1 in ((0, rows), (br if rows > br else 0, rows), (0, min(rows, br + br // 2))):
        d = gctx.empty(max(r1 - r0, 1), cols)
        gctx.call("sysml_b200_binaryblock_read_dense", str(tmp_path / "X0.bb").encode(), rows, cols, br, bc, r0, r1, C.c_void_p(d.ptr))
        assert np.array_equal(d.to_host()[: r1 - r0], X[r0:r1]), (r0, r1)
        d.free()


def test_directory_of_part_files_and_missing_blocks(gctx, tmp_path):
    """Writers split a matrix over part files (IOUtilFunctions.getSequenceFilePaths) and may omit empty blocks."""
    rows, cols, br = 3000, 400, 1000
    X = _matrix(rows, cols, 5)
    X[1000:2000] = 0.0                      # this block row is written to neither part: it must read as zeros
    d = tmp_path / "X.dir"
    d.mkdir()
    _write(d / "0-m-00000", X[:1000], br, 1000, 0)
    _write(d / "0-m-00001", X[2000:], br, 1000, 1, first_block_row=2)     # second part: rows 2000..2999
    (d / "_SUCCESS").write_bytes(b"")
    dev = gctx.empty(rows, cols)
    gctx.call("sysml_b200_binaryblock_read_dense", str(d).encode(), rows, cols, br, 1000, 0, rows, C.c_void_p(dev.ptr))
    assert np.array_equal(dev.to_host(), X)
    dev.free()


def test_first_pass_overlapped_with_the_upload(gctx, tmp_path):
    rows, cols, br = 12_345, 300, 1000
    X = O.rand_uniform(rows * cols, seed=11).reshape(rows, cols)
    v = O.rand_uniform(cols, seed=12)
    w = O.rand_uniform(rows, seed=13, lo=0.0, hi=1.0)
    p = tmp_path / "X.bb"
    _write(p, X, br, 1000, 1)
    dX, dv, dw, out = gctx.empty(rows, cols), gctx.dense(v.reshape(-1, 1)), gctx.dense(w.reshape(-1, 1)), gctx.empty(cols, 1)
    for ct in (1, 0, 2):
        gctx.call("sysml_b200_binaryblock_read_mmchain", str(p).encode(), rows, cols, br, 1000, 0, rows, C.c_void_p(dX.ptr), C.c_void_p(dv.ptr),
                  C.c_void_p(dw.ptr) if ct else None, C.c_void_p(out.ptr), ct)
        assert np.array_equal(dX.to_host(), X)
        assert_equal_matrices(O.mmchain_dense(X, v, w if ct else None, ct, 4), out.to_host().reshape(-1), what="upload-overlapped mmchain ct=%d" % ct)
    for o in (dX, dv, dw, out):
        o.free()


def test_bad_files_are_refused(gctx, tmp_path):
    from systemml_b200 import native
    p = tmp_path / "junk"
    p.write_bytes(b"not a sequence file at all")
    d = gctx.empty(10, 10)
    with pytest.raises(native.NativeError) as ei:
        gctx.call("sysml_b200_binaryblock_read_dense", str(p).encode(), 10, 10, 1000, 1000, 0, 10, C.c_void_p(d.ptr))
    assert "SequenceFile" in str(ei.value)
    with pytest.raises(native.NativeError):
        gctx.call("sysml_b200_binaryblock_read_dense", str(tmp_path / "missing").encode(), 10, 10, 1000, 1000, 0, 10, C.c_void_p(d.ptr))
    X = np.ones((20, 20))
    _write(tmp_path / "ok", X, 1000, 1000, 1)
    with pytest.raises(native.NativeError) as ei:   # declared smaller than the blocks in the file
        gctx.call("sysml_b200_binaryblock_read_dense", str(tmp_path / "ok").encode(), 10, 10, 1000, 1000, 0, 10, C.c_void_p(d.ptr))
    assert ei.value.code == native.E_DIM
    d.free()


def test_parallel_part_files_give_the_same_bits_every_time(gctx, tmp_path):
    """Eight part files are read by eight threads (as ReaderBinaryBlockParallel does); which one is faster must not show in
    the overlapped first pass: group partials are added in group order."""
    rows, cols, br = 40_000, 200, 1000
    X = O.rand_uniform(rows * cols, seed=21).reshape(rows, cols)
    v = O.rand_uniform(cols, seed=22)
    d = tmp_path / "X.parts"
    d.mkdir()
    nparts = 8
    per = rows // br // nparts                      # block rows per part
    for p in range(nparts):
        lo = p * per * br
        _write(d / ("0-m-%05d" % p), X[lo:lo + per * br], br, 1000, 1, first_block_row=p * per)
    dX, dv, out = gctx.empty(rows, cols), gctx.dense(v.reshape(-1, 1)), gctx.empty(cols, 1)
    res = []
    for _ in range(3):
        gctx.call("sysml_b200_binaryblock_read_mmchain", str(d).encode(), rows, cols, br, 1000, 0, rows, C.c_void_p(dX.ptr), C.c_void_p(dv.ptr), None,
                  C.c_void_p(out.ptr), 0)
        res.append(out.to_host().reshape(-1))
    assert np.array_equal(dX.to_host(), X)
    assert np.array_equal(res[0], res[1]) and np.array_equal(res[1], res[2])
    assert_equal_matrices(O.mmchain_dense(X, v, None, 0, 4), res[0], what="parallel parts, overlapped mmchain")
    for o in (dX, dv, out):
        o.free()
