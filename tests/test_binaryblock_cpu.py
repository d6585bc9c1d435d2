"""The binaryblock writer (host code of the library, no GPU needed) against the format the reference defines: a Hadoop
SequenceFile (version 6, uncompressed) of MatrixIndexes -> MatrixBlock records.  The file is parsed here in plain Python,
byte by byte, following MatrixIndexes.write (MatrixIndexes.java:117-130), MatrixBlock.write (MatrixBlock.java:1924-2070) and the
SequenceFile class comment of hadoop-common 2.6.0."""
import ctypes as C
import struct

import numpy as np


def _parse(path):
    b = open(path, "rb").read()
    assert b[:3] == b"SEQ" and b[3] == 6
    pos = 4

    def text():
        nonlocal pos
        n = b[pos]
        assert n < 128          # single-byte vint: both class names are shorter than 128 characters
        s = b[pos + 1:pos + 1 + n].decode()
        pos += 1 + n
        return s
    assert text() == "org.apache.sysml.runtime.matrix.data.MatrixIndexes"
    assert text() == "org.apache.sysml.runtime.matrix.data.MatrixBlock"
    assert b[pos] == 0 and b[pos + 1] == 0      # no compression, no block compression
    pos += 2
    assert struct.unpack(">i", b[pos:pos + 4])[0] == 0   # empty metadata
    pos += 4
    sync = b[pos:pos + 16]
    pos += 16
    blocks = {}
    while pos < len(b):
        reclen = struct.unpack(">i", b[pos:pos + 4])[0]
        pos += 4
        if reclen == -1:
            assert b[pos:pos + 16] == sync
            pos += 16
            continue
        keylen = struct.unpack(">i", b[pos:pos + 4])[0]
        assert keylen == 16
        bi, bj = struct.unpack(">qq", b[pos + 4:pos + 20])
        rlen, clen, typ = struct.unpack(">iib", b[pos + 20:pos + 29])
        payload = b[pos + 29:pos + 4 + reclen]
        pos += 4 + reclen
        blk = np.zeros((rlen, clen))
        if typ == 3:
            blk = np.frombuffer(payload, dtype=">f8").reshape(rlen, clen).astype(np.float64)
        elif typ == 2:
            q = 4
            for r in range(rlen):
                nr = struct.unpack(">i", payload[q:q + 4])[0]
                q += 4
                for _ in range(nr):
                    c, v = struct.unpack(">id", payload[q:q + 12])
                    blk[r, c] = v
                    q += 12
            assert q == len(payload)
        elif typ == 1:
            nnz = struct.unpack(">i", payload[:4])[0]
            q = 4
            for _ in range(nnz):
                r, c, v = struct.unpack(">iid", payload[q:q + 16])
                blk[r, c] = v
                q += 16
            assert q == len(payload)
        else:
            assert typ == 0 and len(payload) == 0
        blocks[(bi, bj)] = (typ, blk)
    return blocks


def test_writer_produces_the_reference_layout(tmp_path):
    from systemml_b200 import native
    lib = native.load()
    rng = np.random.default_rng(0)
    rows, cols, br, bc = 230, 170, 100, 64
    X = rng.uniform(-1, 1, (rows, cols))
    X[:100, 64:128] = 0.0                                   # an empty block
    X[100:200, :64] *= rng.uniform(size=(100, 64)) < 0.05   # a sparse block
    X[200:230, 128:170] = 0.0
    X[205, 130] = 7.5                                       # an ultra-sparse block (nnz < rows)
    p = str(tmp_path / "X.bb").encode()
    assert lib.sysml_b200_binaryblock_write(p, X.ctypes.data_as(C.c_void_p), rows, cols, br, bc, 0, 0) == 0
    blocks = _parse(p.decode())
    assert len(blocks) == 3 * 3
    types = {k: v[0] for k, v in blocks.items()}
    assert types[(1, 2)] == 0 and types[(2, 1)] == 2 and types[(3, 3)] == 1 and types[(1, 1)] == 3
    out = np.zeros_like(X)
    for (bi, bj), (_, blk) in blocks.items():               # 1-based block indexes
        out[(bi - 1) * br:(bi - 1) * br + blk.shape[0], (bj - 1) * bc:(bj - 1) * bc + blk.shape[1]] = blk
    assert np.array_equal(out, X)
    # dense_only: every block is a DENSE_BLOCK
    assert lib.sysml_b200_binaryblock_write(p, X.ctypes.data_as(C.c_void_p), rows, cols, br, bc, 1, 0) == 0
    assert all(v[0] == 3 for v in _parse(p.decode()).values())
