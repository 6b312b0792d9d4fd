"""IRMM rounds 2.. on the device (reference: cphasing/algorithms/hypergraph.py:1258-1397).

State carried from round to round: two per-pem-entry masks (the zeroed monomers of ``H_reweighted`` and the subset
that also reached ``H_T`` through the reference's shared data array), a per-pem-entry weight (``W``, constant inside a hyperedge) and the dense fp64 ``A``.  Kernels: csrc/irmm.cu.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import device as D
from ._lib import call_ws, check, lib, ptr, stream


def ari(a, b):
    """adjusted Rand index of two label vectors (sklearn.metrics.adjusted_rand_score of :1381-1383)."""
    a = np.asarray(a); b = np.asarray(b)
    _, ai = np.unique(a, return_inverse=True)
    _, bi = np.unique(b, return_inverse=True)
    M = np.zeros((ai.max() + 1, bi.max() + 1), dtype=np.int64)
    np.add.at(M, (ai, bi), 1)
    c2 = lambda x: x * (x - 1) / 2.0
    s = c2(M).sum(); sa = c2(M.sum(1)).sum(); sb = c2(M.sum(0)).sum(); t = c2(len(a))
    e = sa * sb / t if t else 0.0
    mx = 0.5 * (sa + sb)
    return 1.0 if mx == e else float((s - e) / (mx - e))


def _dedup_pairs(P, n, dev):
    """(rows, cols) -> deduplicated int64 device vectors (``A.data[offsets] *= f`` acts once per stored entry)."""
    if P is None:
        return None
    r = np.asarray(P[0], dtype=np.int64).ravel(); c = np.asarray(P[1], dtype=np.int64).ravel()
    if r.size == 0:
        return None
    ok = (r >= 0) & (r < n) & (c >= 0) & (c < n)
    key = np.unique(r[ok] * n + c[ok])
    return (torch.from_numpy(key // n).to(dev), torch.from_numpy(key % n).to(dev))


class Rounds:
    """Device state of the iterative re-weighting for one (H, A) pair."""

    def __init__(self, dH: D.DeviceH, a: D.DeviceCSR, node_w=None, P_allelic=None, P_weak=None, allelic_factor=-1.0,
                 min_weight=1.0):
        self.H = dH
        self.n = dH.n_rows
        dev = dH.device
        self.dev = dev
        self.mask = torch.zeros(max(dH.nnz, 1), dtype=torch.uint8, device=dev)       # zeros of H_reweighted
        self.mask_t = torch.zeros(max(dH.nnz, 1), dtype=torch.uint8, device=dev)     # zeros that reached H_T (aliasing)
        self.alias = True           # H still is the object H_T was taken from (:1135, 1298-1310)
        self.w = torch.ones(max(dH.nnz, 1), dtype=torch.float32, device=dev)
        self.dense = torch.empty((self.n, self.n), dtype=torch.float64, device=dev)
        check(lib().cph_dense_from_csr(ptr(a.indptr), ptr(a.indices), ptr(a.data32) if a.data64 is None else None,
                                       ptr(a.data64), self.n, ptr(self.dense), stream()))
        self.accF = torch.empty((self.n, self.n), dtype=torch.int64, device=dev)
        self.node_w = D.to_dev(node_w, torch.float64, dev) if node_w is not None else None
        self.Pa = _dedup_pairs(P_allelic, self.n, dev)
        self.Pw = _dedup_pairs(P_weak, self.n, dev)
        self.allelic_factor = float(allelic_factor) if isinstance(allelic_factor, (int, float)) else -1.0
        self.min_weight = float(min_weight)
        self.suppressed = torch.zeros(1, dtype=torch.int64, device=dev)
        self._bound = dH.value_bound()

    def reweight(self, membership, n_clusters):
        """:1270-1317 -> updates mask and w; returns max |W - W_prev| (:1389)."""
        lab = D.to_dev(np.asarray(membership, dtype=np.int32), torch.int32, self.dev)
        w_prev = self.w.clone()
        self.suppressed.zero_()
        check(lib().cph_irmm_edge_weights(ptr(self.H.pem_row), ptr(self.H.pem_pk), self.H.nnz, ptr(lab), int(n_clusters),
                                          ptr(self.mask), ptr(self.mask_t), int(self.alias), ptr(self.w),
                                          ptr(self.suppressed), stream()))
        if self.alias and int(self.suppressed.item()) > 0:
            self.alias = False
        return float((self.w - w_prev).abs().max().item()) if self.H.nnz else 0.0

    def blend(self, damping=0.5):
        """:1321-1349: A_new = H' W D_e^-1 H^T, threshold, A <- (1-d) A + d A_new, normalise, zero diagonal, pairs."""
        wmax = float(self.w.max().item()) if self.H.nnz else 1.0
        fb = D.frac_bits_for(int(self._bound * max(1.0, np.ceil(wmax))))
        check(lib().cph_expand_full(ptr(self.H.pem_row), ptr(self.H.pem_pk), ptr(self.w), ptr(self.mask), ptr(self.mask_t),
                                    self.H.nnz, self.n, fb, ptr(self.accF), stream()))
        check(lib().cph_irmm_blend(ptr(self.accF), fb, self.min_weight, ptr(self.node_w), self.n, float(damping),
                                   ptr(self.dense), stream()))
        if self.Pa is not None:
            check(lib().cph_dense_scale_pairs(ptr(self.dense), self.n, ptr(self.Pa[0]), ptr(self.Pa[1]),
                                              self.Pa[0].numel(), self.allelic_factor, stream()))
        if self.Pw is not None:
            check(lib().cph_dense_scale_pairs(ptr(self.dense), self.n, ptr(self.Pw[0]), ptr(self.Pw[1]),
                                              self.Pw[0].numel(), 0.0, stream()))

    def to_csr(self, mode) -> D.DeviceCSR:
        """mode 0: A as the reference returns it; mode 1: the undirected graph of triu(A) (:1351-1366)."""
        n = self.n
        need = C.c_size_t(0)
        hn = C.c_uint64(0)
        fn = lib().cph_dense_to_csr_count
        check(fn(ptr(self.dense), n, int(mode), C.byref(hn), None, 0, C.byref(need), None))
        ws = torch.empty(need.value, dtype=torch.uint8, device=self.dev)
        check(fn(ptr(self.dense), n, int(mode), C.byref(hn), ptr(ws), ws.numel(), C.byref(need), stream()))
        nnz = int(hn.value)
        indptr = torch.empty(n + 1, dtype=torch.int64, device=self.dev)
        indices = torch.empty(max(nnz, 1), dtype=torch.int32, device=self.dev)
        d32 = torch.empty(max(nnz, 1), dtype=torch.float32, device=self.dev)
        d64 = torch.empty(max(nnz, 1), dtype=torch.float64, device=self.dev)
        check(lib().cph_dense_to_csr_emit(ptr(self.dense), n, int(mode), ptr(ws), ptr(indptr), ptr(indices), ptr(d32),
                                          ptr(d64), stream()))
        return D.DeviceCSR(indptr, indices[:nnz], d32[:nnz], d64[:nnz], n, nnz)

    def edge_weights(self):
        """W as one float per hyperedge (head entries), for inspection / tests."""
        heads = (self.H.pem_row.view(torch.int32) < 0)
        return self.w[:self.H.nnz][heads]

    def masked_entries(self):
        return int(self.mask[:self.H.nnz].sum().item())
