"""Shared helpers for the parity tests."""
import os

import numpy as np
from scipy.sparse import csr_matrix

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
GOLDEN_CASES = ["toy_hic", "toy_porec", "porec_small", "porec_highorder", "pairs_small"]


def load_golden(name):
    return dict(np.load(os.path.join(GOLDEN, f"ref_{name}.npz")))


def csr_from(g, prefix):
    shape = tuple(int(x) for x in g[prefix + "_shape"])
    return csr_matrix((g[prefix + "_data"], g[prefix + "_indices"], g[prefix + "_indptr"]), shape=shape)


def canon(M):
    """canonical CSR: sorted indices, duplicates kept (stable by column)."""
    M = M.tocsr().copy()
    M.sort_indices()
    return M


def canon_dup(M):
    """canonical form when duplicated (row,col) entries exist: sort by (col, value) inside rows."""
    M = M.tocsr()
    rows = np.repeat(np.arange(M.shape[0]), np.diff(M.indptr))
    o = np.lexsort((M.data, M.indices, rows))
    return rows[o], M.indices[o], M.data[o]


def assert_same_structure_close(X, Y, rtol, thr=None, atol=0.0):
    """X, Y csr.  Same sparsity pattern and values within rtol.  If `thr` is given, entries whose
    value is within rtol of the threshold may be present in only one of the two."""
    X = canon(X); Y = canon(Y)
    assert X.shape == Y.shape
    if thr is None:
        assert np.array_equal(X.indptr, Y.indptr), "row pointer mismatch"
        assert np.array_equal(X.indices, Y.indices), "column index mismatch"
        np.testing.assert_allclose(X.data, Y.data, rtol=rtol, atol=atol)
        return
    D = (X - Y).tocoo()
    Xd = X.todok() if X.nnz < 2_000_000 else None
    bad = 0
    for i, j, d in zip(D.row, D.col, D.data):
        x = X[i, j]; y = Y[i, j]
        if x != 0 and y != 0:
            if abs(x - y) > rtol * max(abs(x), abs(y)) + atol:
                bad += 1
        else:
            v = x if x != 0 else y
            if abs(v - thr) > 4 * rtol * abs(thr):
                bad += 1
    assert bad == 0, f"{bad} entries differ beyond tolerance"
