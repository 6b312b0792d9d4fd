"""Louvain / Leiden modularity clustering on the device (csrc/community.cu) behind the protocol the
reference's callers use on igraph's ``VertexClustering`` (``.membership``, iteration over groups, ``len``).

Reference: cphasing/algorithms/hypergraph.py:1157-1256 (graph preparation inside IRMM and the
``community_multilevel`` / ``community_leiden`` calls).  There is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np
import torch

from . import device as D
from ._lib import nvtx, call_ws, ptr

LOUVAIN, LEIDEN = 0, 1
LOUVAIN_ITERATIONS = 1     # multilevel passes restarted from the partition found (igraph's community_multilevel does one)


class VertexClustering:
    """What IRMM hands back as ``cluster_assignments``: the subset of igraph.VertexClustering the reference touches
    (``.membership`` list, ``len()``, iteration over lists of vertex ids; ``.modularity`` / ``.q`` as a bonus)."""

    def __init__(self, membership, quality=None, stats=None):
        self._m = np.asarray(membership, dtype=np.int64)
        self.quality = quality
        self.stats = stats or {}
        self._groups = None

    @property
    def membership(self):
        return self._m.tolist()

    @property
    def modularity(self):
        return self.quality

    q = modularity

    def _g(self):
        if self._groups is None:
            order = np.argsort(self._m, kind="stable")
            k = int(self._m.max()) + 1 if len(self._m) else 0
            cuts = np.searchsorted(self._m[order], np.arange(1, k))
            self._groups = [g.tolist() for g in np.split(order, cuts)] if k else []
        return self._groups

    def __len__(self):
        return len(self._g())

    def __iter__(self):
        return iter(self._g())

    def __getitem__(self, i):
        return self._g()[i]


def _settings():
    """n_phases / max_sweeps are tuning knobs of the synchronous sweeps (results stay deterministic for fixed values)."""
    return int(os.environ.get("CPH_COMMUNITY_PHASES", "0")), int(os.environ.get("CPH_COMMUNITY_MAX_SWEEPS", "0"))


@nvtx("cph:community")
def community_device(a: D.DeviceCSR, resolution=1.0, method="louvain", n_iterations=2, initial_membership=None,
                     use32=False):
    """a: symmetric CSR(A) on the device (both triangles, sorted indices; the diagonal is ignored).
    -> (membership int32 device tensor, quality float, stats dict)."""
    dev = a.indptr.device
    n = a.n
    mem = torch.empty(max(n, 1), dtype=torch.int32, device=dev)
    init = None
    if initial_membership is not None:
        im = np.asarray(initial_membership, dtype=np.int64)
        if len(im) != n:
            raise ValueError("initial_membership must have one label per vertex")
        _, inv = np.unique(im, return_inverse=True)           # labels into [0, n)
        init = D.to_dev(inv.astype(np.int32), torch.int32, dev)
    q = C.c_double(0.0)
    st = (C.c_uint32 * 4)()
    d32 = a.data32 if (use32 or a.data64 is None) else None
    d64 = None if d32 is not None else a.data64
    phases, sweeps = _settings()
    m = LEIDEN if method == "leiden" else LOUVAIN
    iters = int(os.environ.get("CPH_COMMUNITY_ITERS", "0")) or (int(n_iterations) if m == LEIDEN else LOUVAIN_ITERATIONS)
    call_ws("cph_community", dev, ptr(a.indptr), ptr(a.indices), ptr(d32), ptr(d64), n, a.nnz, float(resolution), m,
            iters, phases, sweeps, ptr(init), ptr(mem), C.byref(q), st)
    stats = {"groups": int(st[0]), "levels": int(st[1]), "sweeps": int(st[2]), "refine_rounds": int(st[3])}
    return mem[:n], float(q.value), stats


def modularity_device(a: D.DeviceCSR, membership, resolution=1.0, use32=False):
    dev = a.indptr.device
    m = np.asarray(membership, dtype=np.int64)
    _, inv = np.unique(m, return_inverse=True)
    md = D.to_dev(inv.astype(np.int32), torch.int32, dev)
    q = C.c_double(0.0)
    d32 = a.data32 if (use32 or a.data64 is None) else None
    d64 = None if d32 is not None else a.data64
    call_ws("cph_modularity", dev, ptr(a.indptr), ptr(a.indices), ptr(d32), ptr(d64), a.n, a.nnz, float(resolution),
            ptr(md), C.byref(q))
    return float(q.value)


def cluster(A, resolution=1.0, method="louvain", n_iterations=2, initial_membership=None) -> VertexClustering:
    """scipy matrix or DeviceBackedCSR -> VertexClustering.  A must be symmetric (IRMM passes the symmetrised
    upper triangle, as igraph sees it)."""
    from . import graphops
    a = graphops.upload_adjacency(A)
    mem, q, stats = community_device(a, resolution, method, n_iterations, initial_membership)
    return VertexClustering(mem.cpu().numpy(), q, stats)
