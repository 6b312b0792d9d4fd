"""Drop-in for ``cphasing.algorithms.hypergraph`` (reference: cphasing/algorithms/hypergraph.py).

Same names, arguments and return contracts as the reference; all arithmetic runs in
libcph's CUDA kernels (``cphasing_b200.device``).  Matrices handed back to callers are real
``scipy.sparse.csr_matrix`` objects (a lazy subclass that also carries the device handle, so the
next stage does not pay a host round trip).  There is no CPU fallback.
"""
from __future__ import annotations

import gc
import logging

import msgspec
import numpy as np
import torch
from scipy.sparse import csr_matrix

from .. import device as D
from .._lib import CphError

logger = logging.getLogger(__name__)

HYPERGRAPH_ORDER_DTYPE = "int8"     # cphasing/_config.py:17
HYPERGRAPH_COL_DTYPE = "int32"      # cphasing/_config.py:18


# ------------------------------------------------------------------------------------------------
# msgpack hooks + HyperEdges (algorithms/hypergraph.py:66-123)
# ------------------------------------------------------------------------------------------------
def numpy_enc_hook(obj):
    if isinstance(obj, np.ndarray):
        return obj.tolist()
    if isinstance(obj, (np.integer, np.floating)):
        return obj.item()
    raise TypeError(f"Encoding objects of type {type(obj)} is unsupported")


def numpy_dec_hook(type, obj):
    if type is np.ndarray or type is list:
        return np.array(obj)
    return obj


class HyperEdges(msgspec.Struct, dict=True):
    """The wire / disk type (algorithms/hypergraph.py:100-123).  ``dict=True`` lets the GPU
    extractors hang the device-resident copy (``_raw``) on the instance without changing the
    msgpack layout."""
    idx: dict
    row: np.ndarray
    col: np.ndarray
    count: np.ndarray
    contigsizes: dict
    mapq: np.ndarray

    @classmethod
    def from_lists(cls, idx, row, col, count, contigsizes, mapq):
        return cls(idx=idx, row=np.array(row, dtype=HYPERGRAPH_COL_DTYPE), col=np.array(col, dtype=HYPERGRAPH_COL_DTYPE),
                   count=np.array(count, dtype=np.uint32), contigsizes=contigsizes, mapq=np.array(mapq, dtype=np.int8))

    def to_numpy(self):
        self.row = np.asarray(self.row, dtype=HYPERGRAPH_COL_DTYPE)
        self.col = np.asarray(self.col, dtype=HYPERGRAPH_COL_DTYPE)
        self.count = np.asarray(self.count, dtype=np.uint32)
        self.mapq = np.asarray(self.mapq, dtype=np.int8)


class LazyTriplets:
    """numpy-like stand-in for a HyperEdges column that still lives on the device: materialises on
    first array use (``np.asarray``, indexing), answers ``len``/``shape``/``size`` without a copy."""

    def __init__(self, provider, n, np_dtype):
        self._p = provider          # callable -> device tensor (the wire-order sort runs on first touch only)
        self._n = int(n)
        self._a = None
        self.dtype = np.dtype(np_dtype)

    def _arr(self):
        if self._a is None:
            t = self._p()
            if t.dtype == torch.uint32:
                t = t.view(torch.int32)
            a = t.cpu().numpy()
            self._a = a.view(self.dtype) if a.dtype != self.dtype else a
            self._p = None
        return self._a

    def __len__(self):
        return self._n

    @property
    def shape(self):
        return (self._n,)

    @property
    def size(self):
        return self._n

    def __array__(self, dtype=None, copy=None):
        a = self._arr()
        return a.astype(dtype) if dtype is not None and np.dtype(dtype) != a.dtype else a

    def __getitem__(self, k):
        return self._arr()[k]

    def __getattr__(self, name):
        if name.startswith("_"):
            raise AttributeError(name)
        return getattr(self._arr(), name)


# ------------------------------------------------------------------------------------------------
# scipy CSR that is filled from the device on first touch and remembers its device twin
# ------------------------------------------------------------------------------------------------
class DeviceBackedCSR(csr_matrix):
    _loader = None
    _cph = None      # DeviceH (incidence) or DeviceCSR (adjacency) twin; dropped when the host copy is re-assigned

    def __init__(self, *args, loader=None, lazy_shape=None, handle=None, **kw):
        self._parts = [None, None, None]
        if loader is not None:
            self._loader = loader
            self._cph = handle
            self._shape = tuple(int(x) for x in lazy_shape)
            self.maxprint = 50
        else:
            csr_matrix.__init__(self, *args, **kw)

    def _get(self, i):
        if self._loader is not None:
            ld, self._loader = self._loader, None
            self._parts = list(ld())
        return self._parts[i]

    def _set(self, i, v):
        if self._loader is not None:
            self._get(0)
        if self._parts[i] is not None and v is not self._parts[i]:
            self._cph = None      # host copy diverged from the device twin
        self._parts[i] = v

    data = property(lambda s: s._get(0), lambda s, v: s._set(0, v))
    indices = property(lambda s: s._get(1), lambda s, v: s._set(1, v))
    indptr = property(lambda s: s._get(2), lambda s, v: s._set(2, v))

    def __reduce__(self):
        return (csr_matrix, ((self.data, self.indices, self.indptr), self.shape))


def device_handle(M):
    return getattr(M, "_cph", None) if isinstance(M, DeviceBackedCSR) else None


def _idx_dtype(nnz):
    return torch.int32 if nnz < 2 ** 31 - 1 else torch.int64


def wrap_incidence(H: D.DeviceH):
    def load():
        ip = H.indptr.to(_idx_dtype(H.nnz)).cpu().numpy()
        return H.counts_f32().cpu().numpy(), H.indices.cpu().numpy(), ip
    return DeviceBackedCSR(loader=load, lazy_shape=(H.n_rows, H.n_cols), handle=H)


def wrap_adjacency(A: D.DeviceCSR, dtype=np.float32):
    def load():
        ip = A.indptr.to(_idx_dtype(A.nnz)).cpu().numpy()
        d = A.data32 if (dtype == np.float32 and A.data32 is not None) else A.data64
        return d.cpu().numpy().astype(dtype, copy=False), A.indices.cpu().numpy(), ip
    return DeviceBackedCSR(loader=load, lazy_shape=(A.n, A.n), handle=A)


def upload_incidence(H) -> D.DeviceH:
    """Any scipy matrix H (n x m, integer-valued counts) -> DeviceH with every row and column kept."""
    h = device_handle(H)
    if isinstance(h, D.DeviceH):
        return h
    Hc = H.tocsr()
    n, m = Hc.shape
    rows = np.repeat(np.arange(n, dtype=np.int32), np.diff(Hc.indptr))
    cnt = np.asarray(Hc.data)
    if not np.all(cnt == np.floor(cnt)) or (cnt.size and cnt.min() < 0):
        raise CphError("clique_expansion_init: H must hold non-negative integer counts "
                       "(the reference builds it from alignment counts)")
    e = D.triplets_to_edges(rows, Hc.indices, cnt.astype(np.uint32), None, n, n_cols=max(m, 1))
    dev = e.row.device
    row_map = torch.arange(n, dtype=torch.int32, device=dev)
    h = D._incidence_build(e.row, e.col, e.count, None, e.nnz, n, 0, row_map, n, 0, False, None, None, m)
    h.pem_col = e.col          # nothing is filtered on this path, so pem order == raw edge-major order
    return h


# ------------------------------------------------------------------------------------------------
# HyperGraph (algorithms/hypergraph.py:145-753)
# ------------------------------------------------------------------------------------------------
class HyperGraph:
    """Same constructor / attributes / methods as the reference class; H and A are built on the GPU."""

    def __init__(self, edges, min_quality=1, mapq_filter=True):
        self.edges = edges
        self.contigsizes = self.edges.contigsizes
        self.min_quality = min_quality
        self.mapq_filter = mapq_filter
        self.get_data()

    # -- algorithms/hypergraph.py:162-205
    def get_data(self):
        if len(self.edges.row) == 0:
            raise ValueError("No hyperedges found.")
        names = np.array(list(self.edges.idx.keys()))
        indices = np.array(list(self.edges.idx.values()))
        self.nodes = names[np.argsort(indices)]
        n = len(self.nodes)
        raw = getattr(self.edges, "_raw", None)
        if raw is None:      # edges that arrive from the host (.hg files, user arrays)
            raw = D.triplets_to_edges(self.edges.row, self.edges.col, self.edges.count, self.edges.mapq, n)
            self.edges._raw = raw
        if raw.n_contigs != n:
            raw = D.RawEdges(raw.row, raw.col, raw.count, raw.mapq, raw.nnz, raw.n_cols, n)
        self._raw = raw
        self._row_drop = None
        self._use_mapq = bool(self.min_quality > 0 and len(self.edges.mapq) > 1 and self.mapq_filter)
        if self._use_mapq:
            rs, _ = D.row_sums(raw, self.min_quality, None, True)
            has = (rs > 0).cpu().numpy()
            self.remove_contigs = self.nodes[~has]
            logger.info(f"Total {int((~has).sum()):,} low-quality (mapq < {self.min_quality}) contigs were removed. ")
            kept_cols = raw.col.view(torch.int32)[raw.mapq >= self.min_quality]
            ncol = int(kept_cols.max().item()) + 1 if kept_cols.numel() else 0
        else:
            self.remove_contigs = np.array([])
            ncol = raw.n_cols
        self.shape = (n, ncol)
        self.removed_count = 0
        self.remove_contig_idx = np.array([])
        self._trip = None

    # the reference exposes the (filtered) triplets as numpy attributes; materialised only on demand
    def _triplets(self):
        if self._trip is None:
            row, col, cnt, mq = D.edges_to_triplets(self._raw)
            keep = torch.ones(row.numel(), dtype=torch.bool, device=row.device)
            if self._use_mapq:
                keep &= mq >= self.min_quality
            if self._row_drop is not None:
                keep &= self._row_drop[row.long()] == 0
            self._trip = (row[keep].cpu().numpy(), col[keep].cpu().numpy(),
                          cnt.view(torch.int32)[keep].cpu().numpy().view(np.uint32),
                          mq[keep].cpu().numpy() if (mq is not None and self._use_mapq) else np.array([]))
        return self._trip

    row = property(lambda s: s._triplets()[0])
    col = property(lambda s: s._triplets()[1])
    count = property(lambda s: s._triplets()[2])
    mapq = property(lambda s: s._triplets()[3])

    # -- algorithms/hypergraph.py:207-223
    def remove_rows(self, contigs):
        idx = [self.edges.idx.get(c) for c in contigs]
        idx = np.asarray([i for i in idx if i is not None], dtype=np.int64)
        self.removed_count += len(idx)
        if self._row_drop is None:
            self._row_drop = torch.zeros(len(self.nodes), dtype=torch.uint8, device=self._raw.row.device)
        if len(idx):
            self._row_drop[torch.from_numpy(idx).to(self._row_drop.device)] = 1
        self._trip = None

    # -- algorithms/hypergraph.py:225-247
    def extract_rows(self, contigs):
        idx = np.asarray([self.edges.idx[c] for c in contigs], dtype=np.int64)
        self.removed_count += len(idx)
        drop = torch.ones(len(self.nodes), dtype=torch.uint8, device=self._raw.row.device)
        if len(idx):
            drop[torch.from_numpy(idx).to(drop.device)] = 0
        self._row_drop = drop if self._row_drop is None else torch.maximum(self._row_drop, drop)
        self._trip = None

    # -- algorithms/hypergraph.py:249-314
    def incidence_matrix(self, min_contacts=1):
        H = D.incidence(self._raw, min_contacts=min_contacts, min_mapq=self.min_quality if self._use_mapq else 0,
                        row_drop=self._row_drop, mapq_filter=self._use_mapq)
        if min_contacts:
            keep = H.row_keep.cpu().numpy()
            self.remove_contig_idx = np.where(~keep)[0]
            n0 = len(self.nodes)
            self.nodes = self.nodes[keep]
            logger.info(f"Total {n0 - H.n_rows:,} low-signal (contacts < {min_contacts}) contigs were removed "
                        f"(--min-contacts). ")
            self.shape = (H.n_rows, H.n_cols)
        else:
            self.remove_contig_idx = np.array([])
            self.remove_contigs = np.array([])
            H.n_cols = self.shape[1]
        self.device_H = H
        return wrap_incidence(H)

    def clear(self):
        self._trip = None
        self._raw = None
        if hasattr(self.edges, "_raw"):
            self.edges._raw = None
        gc.collect()

    # -- algorithms/hypergraph.py:358-485
    @staticmethod
    def clique_expansion_init(H, cis_count=None, P_allelic_idx=None, P_weak_idx=None, NW=None, W=None,
                              allelic_factor=-1, min_weight=0.1, use_high_order=True):
        if cis_count is not None:
            raise NotImplementedError("cis_count is never passed by the reference's live callers "
                                      "(hyperpartition.py:350 sets it to None)")
        dH = upload_incidence(H)
        entry_w = None
        if W is not None:                                                 # :427-430  W.data[0] * d_e
            wv = np.asarray(W.data[0] if hasattr(W, "data") and getattr(W.data, "ndim", 1) == 2 else W, dtype=np.float32)
            wd = D.to_dev(wv, torch.float32, dH.device)
            entry_w = _edge_major_weights(dH, wd)
        node_w = None
        if NW is not None:
            node_w = _node_weights_from_NW(NW)
        A, _acc = D.clique_expansion(dH, use_high_order=use_high_order, min_weight=float(min_weight), node_w=node_w,
                                     entry_w=entry_w, want32=True, want64=True)
        out = wrap_adjacency(A)
        if P_allelic_idx or P_weak_idx:                                   # :464-481
            out = HyperGraph.reweight_pairs(out, P_allelic_idx, P_weak_idx, allelic_factor)
        return out

    @staticmethod
    def reweight_pairs(A, P_allelic_idx, P_weak_idx, allelic_factor):
        """algorithms/hypergraph.py:464-481: A[rows, cols] *= allelic_factor ; A[weak] = 0 (stored entries only)."""
        from .. import graphops
        return graphops.reweight_pairs(A, P_allelic_idx, P_weak_idx, allelic_factor, 0.0)

    # -- algorithms/hypergraph.py:590-602
    @staticmethod
    def normalize_adjacency_matrix(A, NW):
        from .. import graphops
        return graphops.scale_by_node_weights(A, _node_weights_from_NW(NW))


def _node_weights_from_NW(NW):
    """The reference passes the dense outer product NW = w w^T (hyperpartition.py:568); the kernels
    take the vector.  Accept either: a vector is used as is, a matrix gives w = sqrt(diag)."""
    NW = np.asarray(NW)
    if NW.ndim == 1:
        return NW.astype(np.float64)
    return np.sqrt(np.diag(NW).astype(np.float64))


def _edge_major_weights(dH: D.DeviceH, w_col):
    """per pem entry weight: W of the entry's hyperedge.  Hyperedge ids in edge-major order are the
    running count of head bits."""
    if dH.pem_col is not None:
        return w_col[dH.pem_col.view(torch.int32).long()].contiguous()
    head = (dH.pem_row.view(torch.int32) < 0).to(torch.int64)        # head bit = sign bit
    col = torch.cumsum(head, 0) - 1
    if dH.n_cols != int(head.sum().item()):
        # columns were not renumbered (min_contacts = 0 path): recover ids from the CSR view instead
        raise CphError("per-hyperedge weights need a renumbered incidence matrix (build H with min_contacts > 0)")
    return w_col[col].contiguous()


# -- algorithms/hypergraph.py:938-954
def extract_incidence_matrix2(mat, idx):
    h = device_handle(mat)
    if not isinstance(h, D.DeviceH):
        h = upload_incidence(mat)
    idx = np.asarray(idx, dtype=np.int64)
    sub = D.sub_incidence(h, idx)
    kept = _kept_columns(h, idx)
    removed = np.setdiff1d(np.arange(h.n_cols), kept, assume_unique=True) if len(kept) < h.n_cols else np.array([])
    return wrap_incidence(sub), removed, kept


def _kept_columns(h: D.DeviceH, idx):
    """ids (in the parent numbering) of the hyperedges extract_incidence_matrix2 keeps: summed count > 1."""
    dev = h.device
    sel = torch.zeros(h.n_rows, dtype=torch.bool, device=dev)
    sel[torch.from_numpy(idx).to(dev)] = True
    rows = torch.repeat_interleave(torch.arange(h.n_rows, device=dev), h.indptr[1:] - h.indptr[:-1])
    m = sel[rows]
    cnt = (h.pk.view(torch.int32) & 0x00FFFFFF).to(torch.int64)
    s = torch.zeros(h.n_cols, dtype=torch.int64, device=dev)
    s.index_add_(0, h.indices.long()[m], cnt[m])
    return torch.nonzero(s > 1).flatten().cpu().numpy()
