"""Drop-in for ``cphasing.kprune.KPruneHyperGraph`` (reference: cphasing/kprune.py:385-552): allelic and
cross-allelic ("weak") contig pairs read off the clique-expansion matrix A, on the device
(csrc/kprune.cu).  ``kprune_pairs`` is the array-level entry the partition driver uses.
"""
from __future__ import annotations

import logging

from pathlib import Path

import numpy as np
import pandas as pd
import torch

from . import device as D
from . import graphops
from ._lib import nvtx, check, lib, ptr, stream

logger = logging.getLogger(__name__)


def _pairs_array(allele_pairs):
    if isinstance(allele_pairs, np.ndarray):
        return allele_pairs.astype(np.int64, copy=False).reshape(-1, 2)
    return np.asarray(list(allele_pairs), dtype=np.int64).reshape(-1, 2)


def allele_groups(allele_pairs, n, max_alleles=1):
    """CSR-like allele lists per contig, in allele-table order (kprune.py:437-447 keeps the first): for every contig a
    the first `max_alleles` partners b of the rows (a, b), duplicates included."""
    p = _pairs_array(allele_pairs)
    a, b = p[:, 0], p[:, 1]
    order = np.argsort(a, kind="stable")                     # table order survives inside each contig
    a_s = a[order]
    start = np.flatnonzero(np.r_[True, a_s[1:] != a_s[:-1]]) if a_s.size else np.zeros(0, dtype=np.int64)
    rank = np.arange(a_s.size) - np.repeat(start, np.diff(np.r_[start, a_s.size]))
    keep = (rank < max_alleles) & (a_s >= 0) & (a_s < n)
    ptr_ = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(np.bincount(a_s[keep], minlength=n)[:n], out=ptr_[1:])
    idx = np.zeros(max(int(ptr_[-1]), 1), dtype=np.int32)
    idx[:int(ptr_[-1])] = b[order][keep]
    return ptr_, idx


def groups_from_first(first, n):
    """{contig: its one allele} -> the CSR-like (ptr, idx) form of `allele_groups`"""
    ptr_ = np.zeros(n + 1, dtype=np.int64)
    ks = np.array(sorted(k for k in first if 0 <= k < n), dtype=np.int64)
    cnt = np.zeros(n, dtype=np.int64)
    cnt[ks] = 1
    np.cumsum(cnt, out=ptr_[1:])
    idx = np.zeros(max(len(ks), 1), dtype=np.int32)
    idx[:len(ks)] = [first[int(k)] for k in ks]
    return ptr_, idx


@nvtx("cph:kprune")
def kprune_device(a: D.DeviceCSR, allele_pairs, max_alleles=1, want_scores=False, first=None):
    """-> (flag uint8[nnz] device, row_of int32[nnz] device[, best_with, best_all]).
    first: {contig: allele} overriding the allele groups derived from `allele_pairs` (max_alleles = 1 only)."""
    dev = a.indptr.device
    n = a.n
    pairs = _pairs_array(allele_pairs)
    ag_ptr, ag_idx = groups_from_first(first, n) if (first is not None and max_alleles == 1) else \
        allele_groups(pairs, n, max_alleles)
    agp = D.to_dev(ag_ptr, torch.int64, dev); agi = D.to_dev(ag_idx, torch.int32, dev)
    alr = D.to_dev(pairs[:, 0].copy(), torch.int64, dev); alc = D.to_dev(pairs[:, 1].copy(), torch.int64, dev)
    row_of = torch.empty(max(a.nnz, 1), dtype=torch.int32, device=dev)
    check(lib().cph_csr_row_ids(ptr(a.indptr), n, ptr(row_of), stream()))
    flag = torch.empty(max(a.nnz, 1), dtype=torch.uint8, device=dev)
    bw = torch.zeros(max(a.nnz, 1), dtype=torch.float64, device=dev) if want_scores else None
    ba = torch.zeros(max(a.nnz, 1), dtype=torch.float64, device=dev) if want_scores else None
    d64 = a.data64
    d32 = a.data32 if d64 is None else None
    check(lib().cph_kprune(ptr(a.indptr), ptr(a.indices), ptr(d32), ptr(d64), n, a.nnz, ptr(row_of), ptr(agp), ptr(agi),
                           int(max_alleles), ptr(alr), ptr(alc), len(pairs), ptr(flag), ptr(bw), ptr(ba), stream()))
    out = (flag[:a.nnz], row_of[:a.nnz])
    return out + (bw[:a.nnz], ba[:a.nnz]) if want_scores else out


def kprune_device_pairs(A, allele_pairs, max_alleles=1, first=None):
    """A: scipy matrix / DeviceBackedCSR; allele_pairs: (i, j) contig-index pairs in allele-table order.
    Returns (allelic int64[ka, 2], weak int64[kw, 2]) as DEVICE tensors, rows sorted by (i, j), i < j -- plus the (c, c)
    rows the reference's weak list carries (see `_self_pairs`).  Nothing of the size of the pair lists touches the host."""
    a = graphops.upload_adjacency(A)
    flag, row_of = kprune_device(a, allele_pairs, max_alleles, first=first)
    cols = a.indices

    def pick(code):             # CSR order = sorted by (row, col) already
        ix = torch.nonzero(flag == code).flatten()
        return torch.stack([row_of[ix].long(), cols[ix].long()], dim=1)
    allelic, weak = pick(1), pick(2)
    if max_alleles == 1:
        selfp = np.asarray(sorted(_self_pairs(a, allele_pairs, first)), dtype=np.int64).reshape(-1, 2)
        if selfp.size:
            n = max(a.n, 1)
            sp = D.to_dev(selfp, torch.int64, weak.device)
            both = torch.cat([weak, sp])
            order = torch.argsort(both[:, 0] * n + both[:, 1], stable=True)        # (c, c) keys never collide with i < j rows
            weak = both[order]
    return allelic, weak


def kprune_arrays(A, allele_pairs, max_alleles=1, first=None):
    """host (numpy) form of `kprune_device_pairs`"""
    al, wk = kprune_device_pairs(A, allele_pairs, max_alleles, first)
    return D.to_host(al).reshape(-1, 2), D.to_host(wk).reshape(-1, 2)


def kprune_pairs(A, allele_pairs, max_alleles=1):
    """list form of `kprune_arrays`: (allelic list[(i, j)], weak list[(i, j)]), the reference's attribute types."""
    allelic, weak = kprune_arrays(A, allele_pairs, max_alleles)
    return list(map(tuple, allelic.tolist())), list(map(tuple, weak.tolist()))


def _self_pairs(a: D.DeviceCSR, allele_pairs, first=None):
    """The reference's `A.setdiag(0)` (kprune.py:419) stores an explicit zero on every diagonal cell and `triu(A)` keeps
    stored zeros (:427), so every (c, c) is a candidate pair as well.  Its 2x2 block is [c, al] x [c, al]: the straight
    matching scores 0, the crossed one A[c, al] + A[al, c] (minus the allelic zeroing of the upper triangle, :466), so
    (c, c) lands in the weak list exactly when c has an allele group and a stored non-zero contact with it.  n look-ups."""
    if first is None:
        first = {}
        for x, y in allele_pairs:
            x = int(x); y = int(y)
            if x not in first:
                first[x] = y
    cs = np.array([c for c, al in first.items() if 0 <= c < a.n and 0 <= al < a.n and c != al], dtype=np.int64)
    if cs.size == 0:
        return []
    als = np.array([first[int(c)] for c in cs], dtype=np.int64)
    lo, hi = np.minimum(cs, als), np.maximum(cs, als)
    off = graphops.find_pairs(a, hi, lo)                      # the LOWER entry is never zeroed by remove_allelic
    d = a.data64 if a.data64 is not None else a.data32
    ok = off >= 0
    val = torch.zeros(off.numel(), dtype=d.dtype, device=d.device)
    val[ok] = d[off[ok]]
    hit = (val != 0).cpu().numpy()
    return [(int(c), int(c)) for c in cs[hit]]


def allele_frame(alleletable):
    """allele2 table (core.py:289-300) -> DataFrame with contig names in columns 1, 2 and named mz / similarity columns.
    Accepts the reference's AlleleTable (`.data`), a DataFrame, a path, or a sequence of (name1, name2)."""
    if hasattr(alleletable, "data"):
        df = alleletable.data.copy()
    elif isinstance(alleletable, pd.DataFrame):
        df = alleletable.copy()
    elif isinstance(alleletable, (str, Path)):
        df = pd.read_csv(alleletable, sep='\t', header=None, comment="#", index_col=0,
                         names=[-1, 0, 1, 2, 'mz1', 'mz2', 'mzShared', 'similarity', 'strand'])
    else:
        df = pd.DataFrame([(0, a, b, 0, 0, 0, 1.0, 1) for a, b in alleletable],
                          columns=[0, 1, 2, 'mz1', 'mz2', 'mzShared', 'similarity', 'strand'])
    ren = {3: 'mz1', 4: 'mz2', 5: 'mzShared', 6: 'similarity', 7: 'strand'}
    df = df.rename(columns={c: ren[c] for c in df.columns if c in ren})
    for c, d in (('mzShared', 0), ('similarity', 1.0)):
        if c not in df.columns:
            df[c] = d
    return df


class KPruneHyperGraph:
    """kprune.py:385-552.  `alleletable` is an object with a ``.data`` DataFrame whose columns 1 and 2 hold the
    contig names of each allele pair (the reference's AlleleTable), or a plain sequence of (name1, name2)."""
    AlleleHeader2 = ['idx1', 'idx2', 'contig1', 'contig2', 'mz1', 'mz2', 'mzShared', 'similarity', 'strand']

    def __init__(self, alleletable, A, contig2idx, max_alleles=1):
        self.alleletable = alleletable
        df = allele_frame(alleletable)
        # allele group = the FIRST partner listed for a contig in the UNFILTERED table; the contig has a group only if that
        # partner (and the contig itself) is in contig2idx -- a later, surviving partner does not step in (kprune.py:437-447)
        g = df.groupby(1, sort=False)[2].first()
        g = g[g.index.isin(contig2idx) & g.isin(contig2idx)]
        self.allele_group_db = {int(contig2idx[k]): int(contig2idx[v]) for k, v in g.items()}
        df[1] = df[1].map(contig2idx.get)
        df[2] = df[2].map(contig2idx.get)
        df = df.dropna(subset=[1, 2])
        df[1] = df[1].astype(int); df[2] = df[2].astype(int)
        self.alleletable_idx = df
        self.A = A
        self.contig2idx = contig2idx
        self.max_alleles = max_alleles
        self._al_d = self._wk_d = None          # device forms of the two lists (int64 [k, 2]); host arrays on first use
        self._al_h = np.zeros((0, 2), dtype=np.int64)
        self._wk_h = np.zeros((0, 2), dtype=np.int64)
        self._lists = None

    def run(self):
        pairs = self.alleletable_idx[[1, 2]].to_numpy()
        al, wk = kprune_device_pairs(self.A, pairs, self.max_alleles, first=self.allele_group_db)
        logger.info(f"Removed {al.shape[0]} allelic contacts.")
        logger.info(f"Removed {wk.shape[0]} weak contacts with allelic.")
        if self._al_d is not None:              # a second run appends, like the reference's list.extend
            al = torch.cat([self._al_d, al])
            wk = torch.cat([self._wk_d, wk])
        self._al_d, self._wk_d = al, wk
        self._al_h = self._wk_h = None
        self._lists = None

    @property
    def allelic_pairs(self):                # int64 [ka, 2] on the host
        if self._al_h is None:
            self._al_h = D.to_host(self._al_d).reshape(-1, 2)
        return self._al_h

    @property
    def weak_pairs(self):
        if self._wk_h is None:
            self._wk_h = D.to_host(self._wk_d).reshape(-1, 2)
        return self._wk_h

    def device_pairs(self):
        """(allelic, weak) int64 [k, 2] device tensors, or (None, None) before run()"""
        return self._al_d, self._wk_d

    def _as_lists(self):
        if self._lists is None:
            self._lists = (list(map(tuple, self.allelic_pairs.tolist())), list(map(tuple, self.weak_pairs.tolist())))
        return self._lists

    @property
    def allelic_prune_list(self):       # kprune.py:412 (list of (i, j)); built on first use
        return self._as_lists()[0]

    @property
    def weak_prune_list(self):          # kprune.py:413
        return self._as_lists()[1]

    def allelic_meta(self):
        """(mz1, mz2, mzShared, similarity) float64[ka, 4] of the allelic pairs: first allele-table row of each (i, j)."""
        df = self.alleletable_idx
        meta = {}
        cols = list(df.columns)
        for row in df.itertuples(index=False):
            d = dict(zip(cols, row))
            meta.setdefault((d[1], d[2]), tuple(d.get(k, d.get(nm, 0)) for k, nm in
                                                ((3, 'mz1'), (4, 'mz2'), (5, 'mzShared'), (6, 'similarity'))))
        return np.array([meta.get((int(a), int(b)), (0, 0, 0, 0)) for a, b in self.allelic_pairs],
                        dtype=np.float64).reshape(-1, 4)

    def get_prune_table(self, symmetric=False):
        """7-column prune table: contig1 contig2 mz1 mz2 mzShared similarity type (core.py:711-723)."""
        names = np.empty(max(self.contig2idx.values(), default=-1) + 1, dtype=object)
        for c, i in self.contig2idx.items():
            names[i] = c
        al, wk = self.allelic_pairs, self.weak_pairs
        m = self.allelic_meta()
        zeros = np.zeros(len(wk))

        def col(x_al, x_wk, integral):
            v = np.concatenate([x_al, x_wk])
            return v.astype(np.int64) if integral and np.all(v == np.floor(v)) else v
        res = pd.DataFrame({'contig1': np.concatenate([names[al[:, 0]], names[wk[:, 0]]]),
                            'contig2': np.concatenate([names[al[:, 1]], names[wk[:, 1]]]),
                            'mz1': col(m[:, 0], zeros, True), 'mz2': col(m[:, 1], zeros, True),
                            'mzShared': col(m[:, 2], zeros, True), 'similarity': col(m[:, 3], zeros, False),
                            'type': np.concatenate([np.zeros(len(al), dtype=np.int64), np.ones(len(wk), dtype=np.int64)])})
        if symmetric:
            res2 = res.rename(columns={'contig1': 'contig2', 'contig2': 'contig1', 'mz1': 'mz2', 'mz2': 'mz1'})
            res = pd.concat([res, res2[res.columns]], axis=0)
        return res


# ------------------------------------------------------------------------------------------------------------------
# File contract of `cphasing-rs kprune` (kprune.py:51-85, cli.py:3525-3670; call sites hyperpartition.py:1573-1586,
# 1704-1718):   kprune <AlleleTable> <contacts.tsv> <out.prune.table> -n NORM -f FIRST_CLUSTER -w WHITELIST -t T -m METHOD
# The Rust source is not in the reference tree, so its arithmetic is restated from what the tree does say:
# the in-tree Python scorers (KPruneHyperGraph :385-552 = first listed allele, KPruner :269-335 = every listed allele) and
# the option help texts.  PARITY UNPINNED for everything beyond `-n none -m fast|precise`: options whose meaning only the
# binary defines raise instead of guessing silently.
# ------------------------------------------------------------------------------------------------------------------
UNPINNED = "parity unpinned (cphasing-rs absent)"


def read_contacts(path):
    """3-column contacts TSV `contig1<TAB>contig2<TAB>weight` written by HyperGraph.to_contacts
    (algorithms/hypergraph.py:712-753: rows i <= j of A) -> (names list, scipy CSR symmetric float64)."""
    from scipy.sparse import coo_matrix
    df = pd.read_csv(path, sep="\t", header=None, names=["c1", "c2", "w"], comment="#",
                     dtype={"c1": str, "c2": str, "w": np.float64})
    cat = pd.Categorical(pd.concat([df["c1"], df["c2"]], ignore_index=True))
    names = list(cat.categories)
    k = len(df)
    i, j = np.asarray(cat.codes[:k], dtype=np.int64), np.asarray(cat.codes[k:], dtype=np.int64)
    w = df["w"].to_numpy()
    off = i != j
    n = len(names)
    A = coo_matrix((np.concatenate([w, w[off]]), (np.concatenate([i, j[off]]), np.concatenate([j, i[off]]))), shape=(n, n))
    A = A.tocsr()
    A.sum_duplicates()
    A.sort_indices()
    return names, A


def read_first_cluster(path):
    """cluster table `group<TAB>n<TAB>ctg ctg ...` (core.py:1060-) -> dict contig -> group number"""
    out = {}
    with open(path) as fh:
        for gi, line in enumerate(fh):
            f = line.rstrip("\n").split("\t")
            if len(f) >= 3:
                for c in f[2].split(" "):
                    if c:
                        out[c] = gi
    return out


class KPrunerRust:
    """Same constructor and `run()` as the reference's wrapper of the `cphasing-rs kprune` binary (kprune.py:51-85), but the
    scoring runs in-process on the device (csrc/kprune.cu) and no subprocess is started."""
    METHODS = {"fast": 1, "precise": 5}          # alleles per contig that enter the score block (KP_MAXL - 1 = 5)
    # DEVIATION, on purpose: the reference's constructor defaults to method="greedy" (kprune.py:57), which only the
    # cphasing-rs binary defines (PARITY UNPINNED).  Defaulting to it would make `KPrunerRust(at, contacts)`
    # -- how reference-style callers construct it -- raise; the default here is `precise`, the command line's own default
    # (cli.py:3601), and an explicit method="greedy" still refuses loudly.

    def __init__(self, alleletable, contacts, output="prune.contig.table", method="precise", first_cluster=None, count_re=None,
                 whitelist=None, norm_method="none", threads=4, log_dir="logs"):
        self.alleletable = alleletable
        self.contacts = contacts
        self.output = output
        self.method = method
        self.first_cluster = None if (not first_cluster or first_cluster == "none") else first_cluster
        self.count_re = None if (not count_re or count_re == "none") else count_re
        self.whitelist = None if (not whitelist or whitelist == "none") else whitelist
        self.norm_method = norm_method
        self.threads = threads
        if self.method not in self.METHODS:
            raise NotImplementedError(f"kprune -m {self.method}: only the binary defines this method; {UNPINNED}.  "
                                      "`fast` scores the first listed allele (KPruneHyperGraph), `precise` every listed "
                                      "allele (KPruner)")
        if self.norm_method not in ("none", "cis"):
            raise NotImplementedError(f"kprune -n {self.norm_method}: only the binary defines this normalisation; {UNPINNED}")
        if self.count_re is not None:
            raise NotImplementedError(f"kprune --count-re: restriction-site normalisation is defined by the binary only; {UNPINNED}")

    def run(self):
        names, A = read_contacts(self.contacts)
        if self.norm_method == "cis":
            # contacts scaled by the cis (self) contacts of both contigs: w_ij / sqrt(w_ii w_jj); contigs without a cis
            # contact keep their raw weights.  The binary's exact formula is not in the tree.
            logger.warning(f"kprune -n cis: w_ij / sqrt(w_ii * w_jj) is this implementation's reading of the option; {UNPINNED}")
            d = A.diagonal().astype(np.float64)
            s = np.where(d > 0, 1.0 / np.sqrt(np.where(d > 0, d, 1.0)), 1.0)
            C = A.tocoo()
            A = type(A)((C.data * s[C.row] * s[C.col], (C.row, C.col)), shape=A.shape).tocsr()
            A.sort_indices()
        idx = {c: i for i, c in enumerate(names)}
        keep = None
        if self.whitelist is not None:
            wl = {l.strip().split()[0] for l in open(self.whitelist) if l.strip()}
            keep = np.array([c in wl for c in names])
        if self.first_cluster is not None:
            # "Use the first cluster results to execute kprune" (cli.py:3594-3601): only contacts inside one first-round
            # group are candidates; contigs outside every group take no part
            grp = read_first_cluster(self.first_cluster)
            g = np.array([grp.get(c, -1) for c in names], dtype=np.int64)
            C = A.tocoo()
            ok = (g[C.row] >= 0) & (g[C.row] == g[C.col])
            A = type(A)((C.data[ok], (C.row[ok], C.col[ok])), shape=A.shape).tocsr()
            A.sort_indices()
        if keep is not None:
            C = A.tocoo()
            ok = keep[C.row] & keep[C.col]
            A = type(A)((C.data[ok], (C.row[ok], C.col[ok])), shape=A.shape).tocsr()
            A.sort_indices()
        kp = KPruneHyperGraph(self.alleletable, A, idx, max_alleles=self.METHODS[self.method])
        kp.run()
        kp.get_prune_table().to_csv(self.output, sep="\t", header=None, index=None)          # core.py:711-723
        logger.info(f"Successful output the allelic and cross-allelic information into `{self.output}`")
        self.kp = kp
        return kp
