"""ctypes door to oracle/community.c (sequential Louvain / Leiden restatement) -- TEST INFRASTRUCTURE.
Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liboracle.so")
_lib = None


def lib():
    global _lib
    if _lib is None:
        src = os.path.join(_HERE, "community.c")
        if not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
            subprocess.check_call(["make", "-C", _HERE])
        L = C.CDLL(_SO)
        L.oracle_community.restype = C.c_int
        L.oracle_community.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_double, C.c_int, C.c_int, C.c_void_p]
        L.oracle_modularity.restype = C.c_double
        L.oracle_modularity.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_double]
        _lib = L
    return _lib


def _csr(A):
    """symmetric scipy matrix -> (n, indptr i64, indices i32, data f64) with the diagonal removed
    (IRMM: setdiag(0), algorithms/hypergraph.py:1175)."""
    A = A.tocsr().astype(np.float64).copy()
    A.setdiag(0)
    A.eliminate_zeros()
    A.sort_indices()
    return (A.shape[0], np.ascontiguousarray(A.indptr, dtype=np.int64), np.ascontiguousarray(A.indices, dtype=np.int32),
            np.ascontiguousarray(A.data, dtype=np.float64))


def community(A, resolution=1.0, method="louvain", n_iterations=2, initial=None):
    n, ip, ix, w = _csr(A)
    mem = np.full(n, -1, dtype=np.int32) if initial is None else np.ascontiguousarray(initial, dtype=np.int32).copy()
    g = lib().oracle_community(n, ip.ctypes.data, ix.ctypes.data, w.ctypes.data, float(resolution),
                               0 if method == "louvain" else 1, int(n_iterations), mem.ctypes.data)
    return mem, g


def modularity(A, membership, resolution=1.0):
    n, ip, ix, w = _csr(A)
    mem = np.ascontiguousarray(membership, dtype=np.int32)
    return lib().oracle_modularity(n, ip.ctypes.data, ix.ctypes.data, w.ctypes.data, mem.ctypes.data, float(resolution))


def ari(a, b):
    """adjusted Rand index (independent of label names)."""
    a = np.asarray(a); b = np.asarray(b)
    _, ai = np.unique(a, return_inverse=True)
    _, bi = np.unique(b, return_inverse=True)
    M = np.zeros((ai.max() + 1, bi.max() + 1), dtype=np.int64)
    np.add.at(M, (ai, bi), 1)
    c2 = lambda x: x * (x - 1) / 2.0
    s = c2(M).sum(); sa = c2(M.sum(1)).sum(); sb = c2(M.sum(0)).sum(); t = c2(len(a))
    e = sa * sb / t if t else 0.0
    mx = 0.5 * (sa + sb)
    return 1.0 if mx == e else (s - e) / (mx - e)
