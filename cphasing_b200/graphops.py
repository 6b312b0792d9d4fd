"""CSR(A) epilogues on the device: pair re-weighting, node-weight normalisation, degree filters.

Host objects in / out are the scipy matrices the reference's callers expect; the arithmetic is in
csrc/graphops.cu.  Reference: cphasing/algorithms/hypergraph.py:81-97, 464-481, 526-602, 632-709.
"""
from __future__ import annotations

import numpy as np
import torch

from . import device as D
from ._lib import check, lib, ptr, stream


def upload_adjacency(A) -> D.DeviceCSR:
    """scipy matrix (or DeviceBackedCSR with a live twin) -> DeviceCSR with sorted indices."""
    from .algorithms.hypergraph import device_handle
    h = device_handle(A)
    if isinstance(h, D.DeviceCSR):
        return h
    dev = D._dev()
    Ac = A.tocsr()
    if not Ac.has_sorted_indices:
        Ac = Ac.copy()
        Ac.sort_indices()
    n = Ac.shape[0]
    indptr = D.to_dev(Ac.indptr, torch.int64, dev)
    indices = D.to_dev(Ac.indices, torch.int32, dev)
    d64 = D.to_dev(Ac.data, torch.float64, dev)
    d32 = D.to_dev(Ac.data, torch.float32, dev)
    return D.DeviceCSR(indptr, indices, d32, d64, n, int(Ac.nnz))


def clone_adjacency(a: D.DeviceCSR) -> D.DeviceCSR:
    return D.DeviceCSR(a.indptr, a.indices, None if a.data32 is None else a.data32.clone(),
                       None if a.data64 is None else a.data64.clone(), a.n, a.nnz)


def _pairs(P, dev):
    if P is None:
        return None, None, 0
    if hasattr(P, "dev"):                                    # algorithms.hypergraph.DevicePairs: never leaves the device
        r = P.dev[0].to(device=dev, dtype=torch.int64).contiguous()
        c = P.dev[1].to(device=dev, dtype=torch.int64).contiguous()
        return (r, c, int(r.numel())) if r.numel() else (None, None, 0)
    if isinstance(P, (tuple, list)) and len(P) == 2 and isinstance(P[0], torch.Tensor):      # already on the device
        r = P[0].to(device=dev, dtype=torch.int64).contiguous()
        c = P[1].to(device=dev, dtype=torch.int64).contiguous()
        return (r, c, int(r.numel())) if r.numel() else (None, None, 0)
    if isinstance(P, (tuple, list)) and len(P) == 2 and np.asarray(P[0]).ndim == 1:
        r, c = np.asarray(P[0], dtype=np.int64).ravel(), np.asarray(P[1], dtype=np.int64).ravel()
    else:
        arr = np.asarray(P)
        if arr.size == 0:
            return None, None, 0
        if arr.ndim != 2 or arr.shape[1] != 2:
            raise ValueError(f"Invalid pair index format: shape={arr.shape}")
        r, c = arr[:, 0].astype(np.int64), arr[:, 1].astype(np.int64)
    if r.size == 0:
        return None, None, 0
    return D.to_dev(r, torch.int64, dev), D.to_dev(c, torch.int64, dev), int(r.size)


def find_pairs(a: D.DeviceCSR, rows, cols):
    """device find_csr_data_offsets: int64 offsets (-1 = not stored), one per pair."""
    dev = a.indptr.device
    r, c, k = _pairs((rows, cols), dev)
    off = torch.empty(k, dtype=torch.int64, device=dev)
    if k:
        check(lib().cph_csr_find_pairs(ptr(a.indptr), ptr(a.indices), a.n, ptr(r), ptr(c), k, ptr(off), stream()))
    return off


def reweight_device(a: D.DeviceCSR, P_allelic_idx, P_weak_idx, allelic_factor, weak_factor):
    """in place on a's data arrays."""
    dev = a.indptr.device
    ar, ac, na = _pairs(P_allelic_idx, dev)
    wr, wc, nw = _pairs(P_weak_idx, dev)
    if a.nnz == 0 or (na == 0 and nw == 0):
        return a
    mark = torch.empty(a.nnz, dtype=torch.uint8, device=dev)
    check(lib().cph_csr_reweight_pairs(ptr(a.indptr), ptr(a.indices), a.n, a.nnz, ptr(ar), ptr(ac), na,
                                       float(allelic_factor), ptr(wr), ptr(wc), nw, float(weak_factor),
                                       ptr(a.data32), ptr(a.data64), ptr(mark), stream()))
    return a


def reweight_pairs(A, P_allelic_idx, P_weak_idx, allelic_factor, weak_factor):
    from .algorithms.hypergraph import wrap_adjacency
    if not isinstance(allelic_factor, (int, float)) or allelic_factor > 0:
        raise NotImplementedError("only the numeric allelic_factor <= 0 branch of reweight_adjacency_matrix "
                                  "(the CLI default path) is implemented on the device")
    a = clone_adjacency(upload_adjacency(A))
    reweight_device(a, P_allelic_idx, P_weak_idx, allelic_factor, weak_factor)
    return wrap_adjacency(a, dtype=A.dtype if A.dtype in (np.float32, np.float64) else np.float32)


def scale_by_node_weights(A, w):
    from .algorithms.hypergraph import wrap_adjacency
    a = clone_adjacency(upload_adjacency(A))
    wd = D.to_dev(np.asarray(w, dtype=np.float64), torch.float64, a.indptr.device)
    if a.nnz:
        check(lib().cph_csr_scale_nodes(ptr(a.indptr), ptr(a.indices), a.n, ptr(wd), ptr(a.data32), ptr(a.data64),
                                        stream()))
    return wrap_adjacency(a, dtype=A.dtype if A.dtype in (np.float32, np.float64) else np.float32)


def abs_degree(a: D.DeviceCSR, skip_diag=False, use32=False):
    """(deg float64[n], diag float64[n]) device tensors."""
    dev = a.indptr.device
    deg = torch.empty(a.n, dtype=torch.float64, device=dev)
    diag = torch.empty(a.n, dtype=torch.float64, device=dev)
    d32 = a.data32 if (use32 or a.data64 is None) else None
    d64 = None if d32 is not None else a.data64
    check(lib().cph_csr_abs_degree(ptr(a.indptr), ptr(a.indices), a.n, ptr(d32), ptr(d64), int(skip_diag), ptr(deg),
                                   ptr(diag), stream()))
    return deg, diag


def zero_diagonal(a: D.DeviceCSR):
    """A.setdiag(0) on stored entries (algorithms/hypergraph.py:1172); in place."""
    if a.nnz == 0:
        return a
    dev = a.indptr.device
    row_of = torch.empty(a.nnz, dtype=torch.int32, device=dev)
    check(lib().cph_csr_row_ids(ptr(a.indptr), a.n, ptr(row_of), stream()))
    on_diag = row_of == a.indices
    if a.data32 is not None:
        a.data32.masked_fill_(on_diag, 0)
    if a.data64 is not None:
        a.data64.masked_fill_(on_diag, 0)
    return a
