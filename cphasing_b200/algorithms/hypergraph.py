"""Drop-in for ``cphasing.algorithms.hypergraph`` (reference: cphasing/algorithms/hypergraph.py).

Same names, arguments and return contracts as the reference; all arithmetic runs in
libcph's CUDA kernels (``cphasing_b200.device``).  Matrices handed back to callers are real
``scipy.sparse.csr_matrix`` objects (a lazy subclass that also carries the device handle, so the
next stage does not pay a host round trip).  There is no CPU fallback.
"""
from __future__ import annotations

import gc
import logging
import os

import msgspec
import numpy as np
import pandas as pd
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
            a = D.to_host(self._p())
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
    """scipy CSR whose three arrays are fetched from the device on first touch and that remembers its device twin.

    While the twin is attached the host arrays are READ-ONLY views: an in-place edit through numpy
    (``A.data[mask] = 0``, ``A.data *= f`` -- idioms the reference uses on its own matrices, algorithms/hypergraph.py:1165,
    kprune.py:419) raises instead of silently leaving the device copy, which the next stage would read, out of date.
    scipy's own in-place methods (setdiag, eliminate_zeros, sum_duplicates, sort_indices, prune, item assignment, *=, /=)
    first DETACH the matrix -- private writable host arrays, device twin dropped -- and then behave as usual; so does
    re-assigning ``.data`` / ``.indices`` / ``.indptr``.  ``A.copy()`` is a plain, writable matrix."""
    _loader = None
    _cph = None      # DeviceH (incidence) or DeviceCSR (adjacency) twin; dropped when the host copy is detached

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
            if self._cph is not None:
                for a in self._parts:
                    a.setflags(write=False)
        return self._parts[i]

    def _detach(self):
        """private writable host arrays; the device twin no longer describes this matrix"""
        if self._loader is not None:
            self._get(0)
        if self._cph is not None:
            self._cph = None
            self._parts = [a if a.flags.writeable else a.copy() for a in self._parts]

    def setdiag(self, *a, **k):
        self._detach()
        return csr_matrix.setdiag(self, *a, **k)

    def eliminate_zeros(self):
        self._detach()
        return csr_matrix.eliminate_zeros(self)

    def sum_duplicates(self):
        if self._cph is not None and self.has_canonical_format:      # nothing to change: stay attached
            return
        self._detach()
        return csr_matrix.sum_duplicates(self)

    def sort_indices(self):
        if self._cph is not None and self.has_sorted_indices:
            return
        self._detach()
        return csr_matrix.sort_indices(self)

    def prune(self):
        self._detach()
        return csr_matrix.prune(self)

    def __setitem__(self, key, x):
        self._detach()
        return csr_matrix.__setitem__(self, key, x)

    def __imul__(self, other):
        self._detach()
        return csr_matrix.__imul__(self, other)

    def __itruediv__(self, other):
        self._detach()
        return csr_matrix.__itruediv__(self, other)

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
        ip = D.to_host(H.indptr.to(_idx_dtype(H.nnz)))
        return D.to_host(H.counts_f32()), D.to_host(H.indices), ip
    return DeviceBackedCSR(loader=load, lazy_shape=(H.n_rows, H.n_cols), handle=H)


def wrap_adjacency(A: D.DeviceCSR, dtype=np.float32):
    def load():
        ip = D.to_host(A.indptr.to(_idx_dtype(A.nnz)))
        d = A.data32 if (dtype == np.float32 and A.data32 is not None) else A.data64
        return D.to_host(d).astype(dtype, copy=False), D.to_host(A.indices), ip
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
            raw = D.RawEdges(raw.row, raw.col, raw.count, raw.mapq, raw.nnz, raw.n_cols, n, raw.packed, raw.n_edges,
                             raw.min_order)
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
    #: one process per GPU, each holding the hyperedges of ITS concatemer range: set to True on every rank
    #: and incidence_matrix all-reduces the per-contig contact sums (the min_contacts filter is global) while
    #: clique_expansion_init sums the partial accumulators over the ranks (cphasing_b200/parallel.py) -- H stays the
    #: rank's column shard, A comes out global and identical everywhere.  Not in the reference (single process).
    column_shard = False

    def incidence_matrix(self, min_contacts=1):
        hook = None
        if self.column_shard:
            from .. import parallel as par
            hook = par.allreduce_rowsums
        H = D.incidence(self._raw, min_contacts=min_contacts, min_mapq=self.min_quality if self._use_mapq else 0,
                        row_drop=self._row_drop, mapq_filter=self._use_mapq, rowsum_hook=hook)
        H.column_shard = bool(self.column_shard)
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
        if getattr(dH, "column_shard", False) and entry_w is None:       # H is this rank's shard: A = sum over the ranks
            from .. import parallel as par
            acc = par.expand_accumulate_sum(dH, use_high_order=use_high_order)
            A = D.expand_emit(acc, float(min_weight), node_w, True, True)
        else:
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
    def normalize_adjacency_matrix(A, NW=None):
        if NW is None:                  # the reference's callers pass self.NW = None when --normalize is off (:1046, :1779)
            return A
        from .. import graphops
        return graphops.scale_by_node_weights(A, _node_weights_from_NW(NW))


    # -- algorithms/hypergraph.py:486-524
    @staticmethod
    def _to_series_contigsizes(contigsizes, nodes):
        import pandas as pd
        if contigsizes is None:
            raise ValueError("contigsizes is required")
        if isinstance(contigsizes, dict):
            return pd.Series(contigsizes, name="size")
        if isinstance(contigsizes, pd.Series):
            return contigsizes.rename("size") if contigsizes.name is None else contigsizes
        if isinstance(contigsizes, pd.DataFrame):
            df = contigsizes.copy()
            cols = {str(c).lower(): c for c in df.columns}
            if "chrom" in cols and "length" in cols:
                df = df.rename(columns={cols["chrom"]: "contig", cols["length"]: "size"})
            elif not ("contig" in df.columns and "size" in df.columns):
                if df.shape[1] >= 2:
                    df = df.iloc[:, :2].copy()
                    df.columns = ["contig", "size"]
                else:                                   # read_chrom_sizes: index = contig, one column = length
                    df = df.reset_index()
                    df.columns = ["contig", "size"]
            return pd.Series(df["size"].values, index=df["contig"].astype(str).values, name="size")
        raise TypeError(f"Unsupported contigsizes type: {type(contigsizes)}")

    # -- algorithms/hypergraph.py:526-580: degree on the device (cph_csr_abs_degree), n-vector statistics on the host
    @staticmethod
    def filter_adjacency_matrix(A, nodes, contigsizes, invert=False, method="quantile", high_q=0.999, low_q=None,
                                mad_z=3.0):
        import pandas as pd
        from .. import graphops
        n = A.shape[0]
        if A is None or n == 0 or A.nnz == 0:
            return np.arange(n, dtype=np.int64)
        a = graphops.upload_adjacency(A)
        deg, _ = graphops.abs_degree(a, skip_diag=False, use32=(A.dtype == np.float32))
        deg = deg.cpu().numpy()
        cs = HyperGraph._to_series_contigsizes(contigsizes, nodes)
        sizes = np.array(pd.Index(np.asarray(nodes).astype(str)).map(cs), dtype=np.float64)   # writable copy
        sizes[~np.isfinite(sizes)] = np.nan
        sizes[sizes <= 0] = np.nan
        norm = deg / sizes * 10000.0
        x = norm[np.isfinite(norm)]
        if x.size == 0:
            return np.arange(n, dtype=np.int64)
        keep = np.ones(n, dtype=bool)
        if method == "quantile":
            if high_q is not None:
                hi = float(np.nanquantile(x, float(high_q)))
                keep &= ~(np.isfinite(norm) & (norm > hi))
            if low_q is not None:
                lo = float(np.nanquantile(x, float(low_q)))
                keep &= ~(np.isfinite(norm) & (norm < lo))
        elif method == "mad":
            med = float(np.nanmedian(x))
            mad = float(np.nanmedian(np.abs(x - med)))
            if mad <= 0:
                hi = float(np.nanquantile(x, 0.999))
                keep &= ~(np.isfinite(norm) & (norm > hi))
            else:
                robust_z = 0.6745 * (norm - med) / mad
                keep &= ~(np.isfinite(robust_z) & (robust_z > float(mad_z)))
        else:
            raise ValueError("method must be 'quantile' or 'mad'")
        return np.where(keep)[0].astype(np.int64)

    # -- algorithms/hypergraph.py:582-588
    @staticmethod
    def filter_by_weight(A, min_weight=0.1):
        if min_weight > 0:
            A = csr_matrix(A, copy=True)          # the caller's matrix (and its device twin) stays as it was
            A.data[A.data < min_weight] = 0
            A.eliminate_zeros()
        return A

    # -- algorithms/hypergraph.py:604-630
    @staticmethod
    def _normalize_pair_index(P):
        if P is None:
            return None, None
        if isinstance(P, (tuple, list)) and len(P) == 2 and not (
                isinstance(P[0], (tuple, list, np.ndarray)) and len(np.asarray(P[0]).shape) == 2):
            return np.asarray(P[0], dtype=np.int64).ravel(), np.asarray(P[1], dtype=np.int64).ravel()
        arr = np.asarray(P)
        if arr.ndim != 2 or arr.shape[1] != 2:
            raise ValueError(f"Invalid pair index format for P_allelic_idx: shape={arr.shape}")
        return arr[:, 0].astype(np.int64, copy=False), arr[:, 1].astype(np.int64, copy=False)

    # -- algorithms/hypergraph.py:632-709 (numeric allelic_factor <= 0: the CLI default path; the string modes and
    #    positive factors are experimental branches no live caller selects)
    @staticmethod
    def reweight_adjacency_matrix(A, P_allelic_idx=None, P_weak_idx=None, allelic_factor=-1, cross_allelic_factor=0):
        if P_allelic_idx and isinstance(allelic_factor, (int, float)) and not isinstance(allelic_factor, bool) \
                and allelic_factor > 0:
            # :674-681 -- the branch `--allelic-positive-factor` (CLI default 3.0) takes when the first round merges
            # homologous clusters (hyperpartition.py:1903-1907): listed pairs become (A_ij + half the 90th percentile of the
            # stored weights) x factor, INCLUDING pairs that had no contact (the assignment creates the entry).  A few
            # thousand pairs, once per partition: done on the host copy, the structure of A changes.
            out = HyperGraph._reweight_positive(A, P_allelic_idx, float(allelic_factor))
            if P_weak_idx:
                out = HyperGraph.reweight_adjacency_matrix(out, None, P_weak_idx, 0, cross_allelic_factor)
            return out
        if P_allelic_idx or P_weak_idx:
            from .. import graphops
            return graphops.reweight_pairs(A, P_allelic_idx if P_allelic_idx else None,
                                           P_weak_idx if P_weak_idx else None, allelic_factor, cross_allelic_factor)
        return A

    @staticmethod
    def _reweight_positive(A, P, factor):
        Ah = csr_matrix(A, copy=True)
        Ah.sum_duplicates()
        n = Ah.shape[0]
        r, c = HyperGraph._normalize_pair_index(P)
        key = np.unique(r * n + c)                               # a pair listed twice is assigned once
        r, c = key // n, key % n
        baseline = 0.5 * (float(np.percentile(Ah.data, 90)) if Ah.nnz else 1.0)
        vals = np.asarray(Ah[r, c]).ravel().astype(np.float64)
        delta = (vals + baseline) * factor - vals                # what has to be ADDED to reach the assigned value
        from scipy.sparse import coo_matrix
        out = (Ah + coo_matrix((delta.astype(Ah.dtype), (r, c)), shape=Ah.shape).tocsr()).tocsr()
        out.sort_indices()
        return out

    # -- algorithms/hypergraph.py:712-753 (TSV writer: host I/O)
    @staticmethod
    def to_contacts(A, nodes=None, NW=None, min_weight=1.0, contigsizes=None, output="hypergraph.expansion.contacts"):
        import pandas as pd
        Ac = A.tocoo()
        mask = (Ac.row <= Ac.col) & (np.abs(Ac.data) >= min_weight) & np.isfinite(Ac.data)
        r, c, d = Ac.row[mask], Ac.col[mask], Ac.data[mask]
        if nodes is not None:
            nodes = np.asarray(nodes)
            df = pd.DataFrame({"0": nodes[r], "1": nodes[c], "2": d})
            if contigsizes is not None:
                cs = HyperGraph._to_series_contigsizes(contigsizes, nodes)
                l0 = df["0"].map(cs).to_numpy(dtype=np.float64)
                l1 = df["1"].map(cs).to_numpy(dtype=np.float64)
                df["2"] = df["2"] / np.sqrt(l0 * l1)
        else:
            df = pd.DataFrame({"0": r, "1": c, "2": d})
        if output:
            df.to_csv(output, sep="\t", header=False, index=False)
            logger.info(f"Export hypergraph into `{output}`")
            return
        return df


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
    # the reference's callers on this path drop the two index arrays (`H, _, _ = ...`, hyperpartition.py:1009, 1761):
    # they are computed when somebody looks (they need the CSR view of the parent and nnz-sized passes)
    box = {}

    def kept():
        if "k" not in box:
            box["k"] = _kept_columns(h, idx)
        return box["k"]

    def removed():
        k = kept()
        return np.setdiff1d(np.arange(h.n_cols), k, assume_unique=True) if len(k) < h.n_cols else np.array([])
    return wrap_incidence(sub), _LazyArray(removed), _LazyArray(kept)


class _LazyArray:
    """a numpy array computed on first use (len, indexing, iteration, np.asarray, attribute access)"""

    def __init__(self, provider):
        self._p, self._a = provider, None

    def _arr(self):
        if self._a is None:
            self._a = np.asarray(self._p())
            self._p = None
        return self._a

    def __array__(self, dtype=None, copy=None):
        a = self._arr()
        return a if dtype is None else a.astype(dtype, copy=False)

    def __len__(self):
        return len(self._arr())

    def __getitem__(self, k):
        return self._arr()[k]

    def __iter__(self):
        return iter(self._arr())

    def __getattr__(self, name):
        if name.startswith("_"):
            raise AttributeError(name)
        return getattr(self._arr(), name)


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


# ------------------------------------------------------------------------------------------------
# IRMM (algorithms/hypergraph.py:1067-1399)
# ------------------------------------------------------------------------------------------------
def _upper_pairs_symmetric(P):
    """igraph only ever sees triu(A) (algorithms/hypergraph.py:1178): a listed pair acts on the graph through its
    upper-triangle entry only.  Keep the pairs listed with r < c and mirror them, so the symmetric device graph carries
    exactly the edge weights igraph would see.  (The reference's callers pass symmetrised lists,
    hyperpartition.py:438-441, for which this is the whole list.)"""
    r, c = HyperGraph._normalize_pair_index(P)
    if r is None or len(r) == 0:
        return None
    up = r < c
    ru, cu = r[up], c[up]
    if len(ru) == 0:
        return None
    return np.concatenate([ru, cu]), np.concatenate([cu, ru])


class DevicePairs(list):
    """A pair list `[rows, cols]` whose entries live on the DEVICE (int64 vectors); behaves like the two-element list of
    pandas Series the reference passes around (hyperpartition.py:438-461: P_allelic_idx / P_weak_idx) -- indexing,
    iteration, len() and truth value -- but the Series are only built when a host caller really looks, and IRMM /
    re-weighting take the device vectors as they are.  Millions of cross-allelic pairs then never cross PCIe."""

    def __init__(self, rows, cols):
        super().__init__([None, None])
        self.dev = (rows, cols)
        self._upper = None

    def _series(self, i):
        if list.__getitem__(self, i) is None:
            list.__setitem__(self, i, pd.Series(D.to_host(self.dev[i])))
        return list.__getitem__(self, i)

    def __getitem__(self, i):
        if isinstance(i, slice):
            return [self._series(k) for k in range(2)][i]
        return self._series(i if i >= 0 else 2 + i)

    def __iter__(self):
        return iter((self._series(0), self._series(1)))

    def __bool__(self):
        return True

    def __reduce__(self):
        return (list, ([self._series(0), self._series(1)],))

    def upper_symmetric(self):
        """the pairs listed with r < c, mirrored (what `_upper_pairs_symmetric` builds on the host)"""
        if self._upper is None:
            r, c = self.dev
            up = r < c
            ru, cu = r[up], c[up]
            self._upper = (torch.cat([ru, cu]), torch.cat([cu, ru])) if ru.numel() else False
        return self._upper or None


_PAIR_CACHE = []        # (list object 0, list object 1, device, result): the resolution search calls IRMM with the SAME pair lists


def _upper_pairs_device(P, dev):
    """`_upper_pairs_symmetric` as int64 device vectors, remembered for the last few pair lists by object identity (the
    resolution auto-search hands the same two Series to every IRMM call; preparing 5 M pairs on the host each time cost
    more than the clustering).  A caller that edits a list IN PLACE between calls must pass a new object."""
    if not P:
        return None
    if isinstance(P, DevicePairs):
        up = P.upper_symmetric()
        return None if up is None else (up[0].to(dev), up[1].to(dev))
    p0, p1 = (P[0], P[1]) if isinstance(P, (list, tuple)) and len(P) == 2 else (P, None)
    for a, b, d, n, res in _PAIR_CACHE:
        if a is p0 and b is p1 and d == dev and n == len(p0):
            return res
    up = _upper_pairs_symmetric(P)
    res = None if up is None else (D.to_dev(up[0], torch.int64, dev), D.to_dev(up[1], torch.int64, dev))
    _PAIR_CACHE.append((p0, p1, dev, len(p0), res))
    del _PAIR_CACHE[:-4]
    return res


def IRMM(H, A=None, NW=None, P_allelic_idx=None, P_weak_idx=None, allelic_factor=-1, cross_allelic_factor=0.3,
         resolution=1, min_weight=1, threshold=0.01, max_round=1, chunk=10000, threads=10, method="louvain",
         outprefix="all"):
    """Same signature and return contract as the reference: ``(raw_A, A, cluster_assignments, cluster_results)``.

    Graph preparation (:1120-1178) and the clustering (:1239-1252) run on the device; ``cluster_assignments`` offers
    ``.membership`` and iterates over groups like igraph's VertexClustering.  Deviations, all deliberate:
      * the clustering is deterministic (no RNG), so CPHASING_RANDOM_SEED has no effect;
      * by default the caller's A is NOT modified.  The reference re-weights the caller's matrix in place (``A.data[...]
        *= allelic_factor``, ``A.data[weak] = 0``, ``setdiag(0)``, :1157-1172), and its resolution auto-search hands the
        same A to every call (hyperpartition.py:1049-1104): with allelic_factor = -1 the allelic entries change sign on
        every call (pinned by three consecutive calls of the reference, tests/golden/ref_irmm_rounds.npz ``inplace_*``).
        ``REFERENCE_INPLACE_A = True`` (or env CPHASING_B200_INPLACE_IRMM=1) reproduces that: the caller's matrix then
        carries the re-weighted values after the call.  Identical either way at the CLI default ``-af 0``;
      * ``max_round > 1`` (iterative hyperedge re-weighting, :1258-1397; CLI default 1) keeps W, the suppressed
        monomers and A on the device between rounds (csrc/irmm.cu); the matrix arithmetic of rounds >= 2 is fp64, as
        in the reference (its A_new is a float64 product).
    """
    from .. import community, graphops
    if A is None or max_round > 1:        # :1120 -- with max_round > 1 the reference rebuilds A from H as well
        # A = H W D_e^-1 H^T, threshold, normalise, zero diagonal (:1120-1150)
        A = HyperGraph.clique_expansion_init(H, NW=NW, min_weight=float(min_weight) if min_weight > 0 else 0.0)
    raw_A = A
    inplace = REFERENCE_INPLACE_A or os.environ.get("CPHASING_B200_INPLACE_IRMM", "0") == "1"
    if inplace and max_round <= 1:
        raw_A = wrap_adjacency(graphops.clone_adjacency(graphops.upload_adjacency(A)),
                               dtype=A.dtype if A.dtype in (np.float32, np.float64) else np.float32)     # :1157 raw_A = A.copy()
    a = graphops.clone_adjacency(graphops.upload_adjacency(A))
    if P_allelic_idx or P_weak_idx:
        dev = a.indptr.device
        Pa = _upper_pairs_device(P_allelic_idx, dev) if P_allelic_idx else None
        Pw = _upper_pairs_device(P_weak_idx, dev) if P_weak_idx else None
        graphops.reweight_device(a, Pa, Pw, allelic_factor, 0.0)          # :1158-1170
    graphops.zero_diagonal(a)                                             # :1172 (kept as explicit zeros)
    use_leiden = (isinstance(allelic_factor, (int, float)) and allelic_factor < 0) or method == "leiden"
    mem, q, stats = community.community_device(a, float(resolution), "leiden" if use_leiden else "louvain",
                                               n_iterations=2, use32=False)
    cluster_assignments = community.VertexClustering(mem.cpu().numpy(), q, stats)
    cluster_results = [set(c) for c in cluster_assignments]
    if max_round > 1:
        return _irmm_rounds(H, raw_A, a, NW, P_allelic_idx, P_weak_idx, allelic_factor, resolution, min_weight, threshold,
                            max_round, use_leiden, cluster_assignments)
    A_out = wrap_adjacency(a, dtype=A.dtype if A.dtype in (np.float32, np.float64) else np.float32)
    if inplace:
        _adopt(A, a)
    return raw_A, A_out, cluster_assignments, cluster_results


REFERENCE_INPLACE_A = False      # see IRMM: reproduce the reference's in-place re-weighting of the caller's A


def _adopt(A, a: D.DeviceCSR):
    """make the caller's matrix A carry the values of the device matrix `a` (same sparsity pattern): what the reference's
    in-place edits of algorithms/hypergraph.py:1157-1172 leave behind in the caller's object"""
    dt = A.dtype if A.dtype in (np.float32, np.float64) else np.float32
    if isinstance(A, DeviceBackedCSR):
        fresh = wrap_adjacency(a, dtype=dt)
        A._loader, A._cph, A._parts = fresh._loader, fresh._cph, [None, None, None]
    else:
        if not A.has_sorted_indices:
            A.sort_indices()          # the device copy was built from the sorted form: same entry order
        d = a.data32 if (dt == np.float32 and a.data32 is not None) else a.data64
        A.data[:] = D.to_host(d).astype(dt, copy=False)


def _irmm_rounds(H, raw_A, a, NW, P_allelic_idx, P_weak_idx, allelic_factor, resolution, min_weight, threshold, max_round,
                 use_leiden, cluster_assignments):
    """Rounds 2.. of IRMM (:1258-1397): per-hyperedge re-weighting from the current clusters, damped update of A,
    re-clustering (Leiden starts from the previous membership, :1368-1373), until the partition is stable (ARI > 0.99),
    the weights stop moving or max_round is reached.  All matrix work in csrc/irmm.cu; n-sized decisions here."""
    from .. import community, irmm
    dH = device_handle(H)
    if not isinstance(dH, D.DeviceH):
        dH = upload_incidence(H)
    R = irmm.Rounds(dH, a, node_w=_node_weights_from_NW(NW) if NW is not None else None, P_allelic=P_allelic_idx,
                    P_weak=P_weak_idx, allelic_factor=allelic_factor, min_weight=float(min_weight) if min_weight > 0 else 0.0)
    membership = np.asarray(cluster_assignments.membership)
    iter_round, diff_value, previous = 1, 1.0, None
    history = []
    while diff_value > threshold:
        if iter_round >= max_round:
            break
        c = len(cluster_assignments)
        diff_value = R.reweight(membership, c)
        R.blend(0.5)
        g = R.to_csr(1)
        mem, q, stats = community.community_device(g, float(resolution), "leiden" if use_leiden else "louvain",
                                                   n_iterations=2, initial_membership=membership if use_leiden else None,
                                                   use32=False)
        cluster_assignments = community.VertexClustering(mem.cpu().numpy(), q, stats)
        membership = np.asarray(cluster_assignments.membership)
        score = None
        if previous is not None:
            score = irmm.ari(previous, membership)
        history.append({"round": iter_round, "max_dW": diff_value, "masked": R.masked_entries(), "ari": score,
                        "groups": stats["groups"]})
        if score is not None and score > 0.99:
            logger.info("[IRMM] Converged early due to stable partitioning.")
            break
        previous = membership.copy()
        iter_round += 1
    cluster_results = [set(c) for c in cluster_assignments]
    A_out = wrap_adjacency(R.to_csr(0), dtype=np.float64)
    A_out._irmm_history = history
    A_out._irmm_state = R
    return raw_A, A_out, cluster_assignments, cluster_results
