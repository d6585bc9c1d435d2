"""Synthetic CSR matrices for the benchmarks (C3 of BASELINE.json: each cell non-zero with a small probability, so
row lengths are ~Poisson(mean)); built group-by-row-length so that 10^8+ entries take seconds in numpy."""
import numpy as np


def poisson_rows_csr(rows, cols, mean, seed, uniform=False):
    """Returns (rowptr int32, colind int32, val float64): per-row counts ~Poisson(mean) (or exactly `mean` when
    uniform), columns sorted and distinct within a row (stratified: entry j of a row with c entries falls in
    stratum j of width cols // c)."""
    rng = np.random.default_rng(seed)
    if uniform:
        cnt = np.full(rows, int(mean), dtype=np.int64)
    else:
        cnt = np.minimum(rng.poisson(mean, rows), cols).astype(np.int64)
    rp = np.zeros(rows + 1, dtype=np.int64)
    np.cumsum(cnt, out=rp[1:])
    nnz = int(rp[-1])
    assert nnz < 2 ** 31
    ci = np.empty(nnz, dtype=np.int32)
    for c in np.unique(cnt):
        if c == 0:
            continue
        r = np.nonzero(cnt == c)[0]
        w = cols // int(c)
        cc = (np.arange(c, dtype=np.int64)[None, :] * w + rng.integers(0, w, size=(len(r), int(c)))).astype(np.int32)
        pos = rp[r][:, None] + np.arange(c, dtype=np.int64)[None, :]
        ci[pos.reshape(-1)] = cc.reshape(-1)
    va = rng.uniform(-1, 1, nnz)
    return rp.astype(np.int32), ci, va
