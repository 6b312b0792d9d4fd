"""CPU oracle for the hyperpartition hot path -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference``
legs may import this module.  The product (``cphasing_b200``) never does.

A plain numpy/scipy restatement of the reference's algorithm, each function citing the
reference file:line it follows (all paths under /root/reference/cphasing/).

Parity status of this oracle:
  * PINNED  against outputs of the reference's own code (stub-imported, see
    tests/golden/make_golden.py -> tests/golden/ref_*.npz):
      hypergraph_get_data, incidence_matrix, clique_expansion (fp32 literal + fp64),
      normalize_weight_svc, normalize_adjacency, filter_adjacency, extract_incidence_matrix2,
      find_csr_data_offsets-based re-weighting.
  * UNPINNED (the reference routes these through polars / igraph / cphasing-rs, none of which
    is installable here; restated from the call sites and the published algorithm):
      porec_edges, pairs_edges (polars group_by/sort), kprune_hypergraph (igraph
      maximum_bipartite_matching on a 2x2), Louvain/Leiden (oracle/community.c).
"""
from __future__ import annotations

import numpy as np
from scipy.sparse import csr_matrix, triu

ORDER_DTYPE = np.int8          # _config.py:17  HYPERGRAPH_ORDER_DTYPE
COL_DTYPE = np.int32           # _config.py:18  HYPERGRAPH_COL_DTYPE


# --------------------------------------------------------------------------------------------
# extractor pre-filters: HCR BED and split-contig index
# --------------------------------------------------------------------------------------------
def hcr_hits(names, pos, bed_names, bed_start, bed_end):
    """hypergraph.py:1721-1752 / :786-893 restated without polars: ``join_asof(strategy="backward", by=chrom, on=pos)``
    against the BED sorted by start picks, per record, the LAST interval of the same contig whose start <= pos; the record
    is inside the HCR iff ``pos < end`` of that interval (no match -> outside).  Plain per-record loop over a per-contig
    list: deliberately the slow, obviously-correct form.  Returns a bool array."""
    names = np.asarray(names); pos = np.asarray(pos).astype(np.int64)
    by = {}
    for c, a, b in zip(np.asarray(bed_names).tolist(), np.asarray(bed_start).tolist(), np.asarray(bed_end).tolist()):
        by.setdefault(c, []).append((a, b))
    for c in by:
        by[c].sort(key=lambda t: t[0])          # python's sort is stable, like the sort("start_bed") it restates
    out = np.zeros(len(names), dtype=bool)
    for i, (c, p) in enumerate(zip(names.tolist(), pos.tolist())):
        hit = None
        for a, b in by.get(c, ()):
            if a <= p:
                hit = b
            else:
                break
        out[i] = hit is not None and p < hit
    return out


def split_index(names, pos, split_init, split_length):
    """hypergraph.py:1406-1424 / :994-1023: row = init_idx[name] + pos // split_length; names outside the map drop
    (inner join).  split_init: dict name code -> first bin (or array with -1 for missing).  Returns (row, ok)."""
    names = np.asarray(names).astype(np.int64); pos = np.asarray(pos).astype(np.int64)
    init = np.asarray(split_init).astype(np.int64)
    ok = (names >= 0) & (names < len(init))
    i0 = np.where(ok, init[np.where(ok, names, 0)], -1)
    ok &= i0 >= 0
    return i0 + pos // int(split_length), ok


# --------------------------------------------------------------------------------------------
# stage 1a: Pore-C records -> hyperedge triplets
# --------------------------------------------------------------------------------------------
def porec_edges(read_idx, contig_idx, mapq, start=None, end=None, contig_len=None,
                min_mapq=1, min_order=2, max_order=50, edge_length=0, n_contigs=None, hcr=None, split=None):
    """hypergraph.py:1774-1775 (mapq >= min_quality), :1438-1457 (edge filter),
    :1461-1468 (distinct-contig order filter), :1470-1477 (group_by read,contig -> count, max mapq),
    :1877-1889 (sort by contig; cast).  Records whose contig is not in the index (inner join
    :1435) are marked with contig_idx >= n_contigs and dropped.

    Returns (row i32, col i32, count u32, mapq i8) canonically sorted by (row, col).
    """
    read_idx = np.asarray(read_idx).astype(np.int64)
    contig_idx = np.asarray(contig_idx).astype(np.int64)
    mapq = np.asarray(mapq).astype(np.int8)
    keep = np.ones(read_idx.shape, dtype=bool)
    if n_contigs is not None and split is None and hcr is None:
        keep &= contig_idx < n_contigs
    if min_mapq > 0:                                        # :1774
        keep &= mapq >= min_mapq
    if hcr is not None or split is not None:
        # hcr = (bed_names, bed_start, bed_end, invert); split = (split_init per name code, split_length).  The contig
        # column holds codes of the ORIGINAL names here.  UInt32 midpoint as in :1738 / :1419.
        mid = np.asarray(start).astype(np.int64) + (np.asarray(end).astype(np.int64) - np.asarray(start).astype(np.int64)) // 2
        if hcr is not None:
            hit = hcr_hits(contig_idx, mid, hcr[0], hcr[1], hcr[2])
            keep &= ~hit if hcr[3] else hit
        if split is not None:
            contig_idx, ok = split_index(contig_idx, mid, split[0], split[1])
            keep &= ok
            edge_length = 0                                  # the split branch of :1406-1457 has no edge filter
        if n_contigs is not None:
            keep &= (contig_idx >= 0) & (contig_idx < n_contigs)
    if edge_length:
        # pos = start + (end-start)//2 ; keep pos < L_e or (length - pos) < L_e   (:1452-1456)
        st = np.asarray(start).astype(np.int64)
        en = np.asarray(end).astype(np.int64)
        ci = np.where(keep, contig_idx, 0)
        length = np.asarray(contig_len).astype(np.int64)[ci]
        pos = st + (en - st) // 2
        keep &= (pos < edge_length) | ((length - pos) < edge_length)
    r = read_idx[keep]
    c = contig_idx[keep]
    q = mapq[keep]
    if r.size == 0:
        z = np.zeros(0, np.int32)
        return z, z.copy(), np.zeros(0, np.uint32), np.zeros(0, np.int8)
    nc = int(c.max()) + 1
    key = r * nc + c
    order = np.argsort(key, kind="stable")
    key = key[order]; q = q[order]
    head = np.concatenate(([True], key[1:] != key[:-1]))
    starts = np.nonzero(head)[0]
    ukey = key[starts]
    count = np.diff(np.concatenate((starts, [key.size])))
    qmax = np.maximum.reduceat(q.astype(np.int16), starts).astype(np.int8)
    ur = ukey // nc
    uc = ukey % nc
    # n_unique(chrom_idx) over read_idx  (:1461-1463)
    rhead = np.concatenate(([True], ur[1:] != ur[:-1]))
    rstart = np.nonzero(rhead)[0]
    rorder = np.diff(np.concatenate((rstart, [ur.size])))
    order_of = np.repeat(rorder, rorder)
    ok = (order_of >= min_order) & (order_of <= max_order)   # :1465-1468
    ur, uc, count, qmax = ur[ok], uc[ok], count[ok], qmax[ok]
    o2 = np.lexsort((ur, uc))                                # sort by contig (:1877), canonical col order
    return (uc[o2].astype(np.int32), ur[o2].astype(np.int32),
            count[o2].astype(np.uint32), qmax[o2].astype(np.int8))


# --------------------------------------------------------------------------------------------
# stage 1b: Hi-C pairs -> triplets
# --------------------------------------------------------------------------------------------
def pairs_edges(chrom1, chrom2, mapq, pos1=None, pos2=None, contig_len=None,
                min_mapq=1, edge_length=0, n_contigs=None, hcr=None, split=None):
    """hypergraph.py:970-971 (mapq filter), :977-992 (both-ends edge filter, note the strict
    ``pos > length - edge_length`` form), :1031 (col = running row number of the surviving pairs),
    :1033-1044 (two triplets per pair, sorted by row), :1287-1292 (count = ones u32).
    Intra-contig pairs are kept and give a duplicated (row, col).
    """
    c1 = np.asarray(chrom1).astype(np.int64)
    c2 = np.asarray(chrom2).astype(np.int64)
    q = np.asarray(mapq).astype(np.int8)
    keep = np.ones(c1.shape, dtype=bool)
    if n_contigs is not None and hcr is None and split is None:
        keep &= (c1 < n_contigs) & (c2 < n_contigs)
    if min_mapq > 0:
        keep &= q >= min_mapq
    if hcr is not None:                                      # :786-893: a pair stays iff EITHER end is inside an HCR
        any_in = hcr_hits(c1, pos1, hcr[0], hcr[1], hcr[2]) | hcr_hits(c2, pos2, hcr[0], hcr[1], hcr[2])
        keep &= ~any_in if hcr[3] else any_in
    if split is not None:                                    # :994-1023
        c1, ok1 = split_index(c1, pos1, split[0], split[1])
        c2, ok2 = split_index(c2, pos2, split[0], split[1])
        keep &= ok1 & ok2
    if n_contigs is not None and (hcr is not None or split is not None):
        keep &= (c1 >= 0) & (c1 < n_contigs) & (c2 >= 0) & (c2 < n_contigs)
    if edge_length:
        L = np.asarray(contig_len).astype(np.int64)
        p1 = np.asarray(pos1).astype(np.int64); p2 = np.asarray(pos2).astype(np.int64)
        l1 = L[np.where(keep, c1, 0)]; l2 = L[np.where(keep, c2, 0)]
        keep &= ((p1 < edge_length) | (p1 > l1 - edge_length)) & ((p2 < edge_length) | (p2 > l2 - edge_length))
    c1, c2, q = c1[keep], c2[keep], q[keep]
    col = np.arange(c1.size, dtype=np.int64)
    row = np.concatenate((c1, c2))
    col = np.concatenate((col, col))
    qq = np.concatenate((q, q))
    o = np.lexsort((col, row))
    return (row[o].astype(np.int32), col[o].astype(np.int32),
            np.ones(row.size, np.uint32), qq[o].astype(np.int8))


# --------------------------------------------------------------------------------------------
# stage 1c: triplets -> filtered incidence matrix
# --------------------------------------------------------------------------------------------
def hypergraph_get_data(row, col, count, mapq, min_quality=1, mapq_filter=True):
    """algorithms/hypergraph.py:175-203."""
    row = np.asarray(row); col = np.asarray(col); count = np.asarray(count); mapq = np.asarray(mapq)
    if min_quality > 0 and len(mapq) > 1 and mapq_filter:
        k = mapq.astype(np.int8) >= min_quality
        row, col, count = row[k], col[k], count[k]
    return row, col, count, int(col.max()) + 1


def incidence_matrix(row, col, count, n_rows, n_cols, min_contacts=1):
    """algorithms/hypergraph.py:249-314.

    Returns (H csr float32, row_keep bool[n_rows], col_keep bool[n_cols] or None).
    Duplicated (row, col) entries are NOT merged (raw indptr construction, :277-282).
    """
    rows = np.asarray(row, dtype=np.int32)
    cols = np.asarray(col, dtype=COL_DTYPE)
    data = np.asarray(count, dtype=np.float32)
    if rows.size and not np.all(rows[:-1] <= rows[1:]):     # :269-275
        o = np.argsort(rows, kind="stable")
        rows, cols, data = rows[o], cols[o], data[o]
    indptr = np.concatenate(([0], np.cumsum(np.bincount(rows, minlength=n_rows))))
    H = csr_matrix((data, cols, indptr), shape=(n_rows, n_cols), dtype=np.float32)
    if not min_contacts:
        return H, np.ones(n_rows, bool), None
    rsum = np.zeros(n_rows, dtype=np.float64)
    np.add.at(rsum, rows, data.astype(np.float64))
    row_keep = rsum.astype(np.float32) >= min_contacts      # :287
    keep_e = row_keep[rows]
    csum = np.bincount(cols[keep_e], weights=data[keep_e].astype(np.float64), minlength=n_cols)
    col_keep = csum >= 2                                    # :305
    new_row = np.cumsum(row_keep) - 1
    new_col = np.cumsum(col_keep) - 1
    ke = keep_e & col_keep[cols]
    r2 = new_row[rows[ke]]; c2 = new_col[cols[ke]]; d2 = data[ke]
    n2 = int(row_keep.sum()); m2 = int(col_keep.sum())
    indptr2 = np.concatenate(([0], np.cumsum(np.bincount(r2, minlength=n2))))
    H2 = csr_matrix((d2, c2.astype(np.int32), indptr2), shape=(n2, m2), dtype=np.float32)
    return H2, row_keep, col_keep


def extract_incidence_matrix2(H, idx):
    """algorithms/hypergraph.py:938-954: sub_H = H[idx] minus columns whose summed count <= 1."""
    A = H[idx]
    if A.nnz > 0:
        s = np.bincount(A.indices, weights=A.data, minlength=A.shape[1])
        nz = s > 1
        keep = np.where(nz)[0]; rem = np.where(~nz)[0]
        if len(keep) < A.shape[1]:
            A = A.tocsc()[:, keep].tocsr()
    else:
        rem = np.array([]); keep = np.array([])
    return A, rem, keep


# --------------------------------------------------------------------------------------------
# stage 2: clique expansion
# --------------------------------------------------------------------------------------------
def edge_weights(H, use_high_order=True, W=None):
    """algorithms/hypergraph.py:385-395, 418, 427-430: d_e = 1/max(int8(nnz_e) - 1, 1) [* W]."""
    m = H.shape[1]
    if use_high_order:
        with np.errstate(over="ignore"):
            D = H.getnnz(axis=0).astype(ORDER_DTYPE) - np.int8(1)
    else:
        D = np.ones(m, dtype=ORDER_DTYPE)
    d = 1.0 / np.maximum(D, 1)
    if W is not None:
        d = np.asarray(W, dtype=np.float64) * d
    return d


def clique_expansion(H, use_high_order=True, W=None, min_weight=0.1, dtype=np.float64):
    """algorithms/hypergraph.py:372-455:  A = (H sqrt(d)) (H sqrt(d))^T, entries < min_weight dropped.

    dtype=float64 is the gate the GPU path is held to (north_star: fp64 accumulation, 1e-6 rel);
    dtype=float32 replays the reference literally (f32 scale, f32 sqrt, f32 SpGEMM).
    """
    d = edge_weights(H, use_high_order, W).astype(dtype)
    Hs = H.astype(dtype).tocsr().copy()
    Hs.data *= np.sqrt(d)[Hs.indices]
    A = (Hs @ Hs.T).tocsr()
    if min_weight > 0:                                      # :453-455
        A.data[A.data.astype(np.float32) < np.float32(min_weight)] = 0
        A.eliminate_zeros()
    A.sort_indices()
    return A


def clique_expansion_exact(H, use_high_order=True, W=None):
    """Same contraction without the sqrt split:  A_ij = sum_e d_e c_ie c_je  in float64
    (integer-exact when use_high_order=False and W is None)."""
    d = edge_weights(H, use_high_order, W)
    H64 = H.astype(np.float64).tocsr()
    Hd = H64.copy(); Hd.data *= d[Hd.indices]
    A = (Hd @ H64.T).tocsr()
    A.sort_indices()
    return A


def contact_counts(H):
    """The integer contact-count matrix  C = H H^T  in int64 (bit-exact gate)."""
    Hi = csr_matrix((H.data.astype(np.int64), H.indices, H.indptr), shape=H.shape)
    C = (Hi @ Hi.T).tocsr()
    C.sort_indices()
    return C


def normalize_weight_svc(H, lengths):
    """hyperpartition.py:530-533, 554-558: w_i = 1/sqrt(deg_i + 0.5 * a_i * sum(deg)/sum(a))."""
    a = np.asarray(lengths, dtype=np.float64).copy()
    a[a <= 0] = 1.0
    deg = np.asarray(H.sum(axis=1)).ravel().astype(np.float64)
    deg[deg <= 0] = 1.0
    rho = deg.sum() / a.sum()
    return 1.0 / np.sqrt(deg + 0.5 * a * rho)


def normalize_adjacency(A, w):
    """hyperpartition.py:568 (NW = outer(w,w) as f32) + algorithms/hypergraph.py:594-600."""
    coo = A.tocoo()
    nw = (w[coo.row] * w[coo.col]).astype(np.float32)
    return csr_matrix((coo.data * nw, (coo.row, coo.col)), shape=A.shape)


def filter_adjacency(A, lengths, high_q=0.999):
    """algorithms/hypergraph.py:540-565 (method='quantile')."""
    n = A.shape[0]
    if n == 0 or A.nnz == 0:
        return np.arange(n, dtype=np.int64)
    deg = np.asarray(np.abs(A).sum(axis=1)).ravel().astype(np.float64)
    sizes = np.asarray(lengths, dtype=np.float64).copy()
    sizes[~np.isfinite(sizes)] = np.nan
    sizes[sizes <= 0] = np.nan
    norm = deg / sizes * 10000.0
    x = norm[np.isfinite(norm)]
    if x.size == 0:
        return np.arange(n, dtype=np.int64)
    hi = float(np.nanquantile(x, float(high_q)))
    keep = ~(np.isfinite(norm) & (norm > hi))
    return np.where(keep)[0].astype(np.int64)


# --------------------------------------------------------------------------------------------
# stage 3: kprune on A
# --------------------------------------------------------------------------------------------
def allele_group_first(pairs):
    """kprune.py:437-447: allele group of a contig = the FIRST allele listed for it."""
    db = {}
    for a, b in pairs:
        if a not in db:
            db[a] = b
    return db


def kprune_hypergraph(A, allele_pairs, first_group=None):
    """kprune.py:407-517 (KPruneHyperGraph.run).

    A: csr adjacency (diagonal ignored, :419).  allele_pairs: list of (i, j) index pairs in
    allele-table order.  Returns (allelic list[(i,j)], weak list[(i,j)]) each sorted.

    allelic (:455-468) = allele pairs i<j that are stored non-zeros of triu(A); they are zeroed.
    weak (:470-517): for every remaining stored pair of triu(A) (the diagonal included, see below) whose
    two contigs both have an allele group, score the 2x2 block [c1,al1]x[c2,al2] on A and call
    the pair weak unless the max-weight perfect matching pairs c1 with c2.
    Tie rule (igraph's choice on equal sums is unspecified): a tie counts as strong.
    """
    A = A.tocsr().astype(np.float64).copy()
    A.setdiag(0)
    A.eliminate_zeros()
    U = triu(A).tocoo()
    pairs = set(zip(U.row.tolist(), U.col.tolist()))
    al = sorted(set((int(a), int(b)) for a, b in allele_pairs if a < b) & pairs)
    L = A.tolil()
    for a, b in al:
        L[a, b] = 0                       # :466 only the upper entry is zeroed
    D = L.todok()
    pairs -= set(al)
    # first_group: {contig: allele} taken from the UNFILTERED name table (get_allele_group, :437-447) when the contig
    # index does not cover the whole allele table; by default the first partner among the given index pairs
    grp = allele_group_first(allele_pairs) if first_group is None else dict(first_group)
    # `self.A.setdiag(0)` (:419) stores an explicit zero on EVERY diagonal cell and triu(A) keeps stored zeros, so every
    # (c, c) is a candidate pair too (:427).  Its 2x2 block is [c, al] x [c, al]: straight = 0, cross = A[c,al] + A[al,c]
    # after the allelic zeroing of the upper triangle -- the reference lists (c, c) as weak whenever c touches its allele.
    pairs |= {(i, i) for i in range(A.shape[0])}
    weak = []
    for c1, c2 in sorted(pairs):
        if c1 not in grp or c2 not in grp:
            continue
        a1, a2 = grp[c1], grp[c2]
        s = lambda i, j: float(D.get((i, j), 0.0))
        straight = s(c1, c2) + s(a1, a2)
        cross = s(c1, a2) + s(a1, c2)
        if cross > straight:
            weak.append((c1, c2))
    return al, weak


def kprune_general(score, allele_lists, cand_pairs):
    """kprune.py:269-335 (KPruner): l1 x l2 score block over [ctg,*alleles]; weak unless the
    maximum-weight matching assigns ctg1 to ctg2.  Brute-force assignment (small ploidy).
    score(i,j) -> float, allele_lists: dict ctg -> list of alleles."""
    from itertools import permutations
    weak = []
    for c1, c2 in cand_pairs:
        if c1 not in allele_lists or c2 not in allele_lists:
            continue
        g1 = [c1, *allele_lists[c1]]; g2 = [c2, *allele_lists[c2]]
        if len(g1) > len(g2):
            # transpose so rows are the shorter side
            S = np.array([[score(a, b) for a in g1] for b in g2]); tr = True
        else:
            S = np.array([[score(a, b) for b in g2] for a in g1]); tr = False
        r, c = S.shape
        best = -1.0; best_has = False
        for perm in permutations(range(c), r):
            tot = sum(S[i, perm[i]] for i in range(r))
            has = perm[0] == 0
            if tot > best + 1e-12 or (abs(tot - best) <= 1e-12 and has and not best_has):
                best, best_has = tot, has
        if not best_has:
            weak.append((c1, c2))
    return weak


# --------------------------------------------------------------------------------------------
# stage 4: graph preparation for clustering (IRMM)
# --------------------------------------------------------------------------------------------
def find_csr_data_offsets(A, rows, cols):
    """algorithms/hypergraph.py:81-97."""
    rows = np.asarray(rows); cols = np.asarray(cols)
    if len(rows) == 0:
        return np.array([], dtype=np.int64)
    coo = A.tocoo()
    nc = A.shape[1]
    ck = coo.row.astype(np.int64) * nc + coo.col.astype(np.int64)
    tk = rows.astype(np.int64) * nc + cols.astype(np.int64)
    s = np.argsort(ck)
    ii = np.clip(np.searchsorted(ck, tk, sorter=s), 0, len(ck) - 1)
    mi = s[ii]
    return mi[ck[mi] == tk]


def irmm_graph(A, P_allelic=None, P_weak=None, allelic_factor=-1.0):
    """algorithms/hypergraph.py:1157-1178, 1237-1238: re-weight, drop the diagonal, upper triangle.
    Returns (src i32, dst i32, weight f64) of the undirected graph handed to igraph."""
    A = A.tocsr().copy()
    if P_allelic is not None and len(P_allelic[0]):
        off = find_csr_data_offsets(A, np.asarray(P_allelic[0]), np.asarray(P_allelic[1]))
        A.data[off] *= allelic_factor
    if P_weak is not None and len(P_weak[0]):
        off = find_csr_data_offsets(A, np.asarray(P_weak[0]), np.asarray(P_weak[1]))
        A.data[off] = 0
    A.setdiag(0)
    U = triu(A, format="csr")
    U.eliminate_zeros()
    U.sort_indices()
    U = U.tocoo()
    return U.row.astype(np.int32), U.col.astype(np.int32), U.data.astype(np.float64)


def modularity(n, src, dst, w, membership, resolution=1.0):
    """Q = sum_c [ in_c/m - gamma (tot_c/2m)^2 ]  (undirected, weights w, no self loops needed)."""
    membership = np.asarray(membership)
    m = w.sum()
    if m == 0:
        return 0.0
    deg = np.bincount(src, weights=w, minlength=n) + np.bincount(dst, weights=w, minlength=n)
    nc = int(membership.max()) + 1
    tot = np.bincount(membership, weights=deg, minlength=nc)
    same = membership[src] == membership[dst]
    inn = np.bincount(membership[src][same], weights=w[same], minlength=nc)
    return float((inn / m - resolution * (tot / (2 * m)) ** 2).sum())


# --------------------------------------------------------------------------------------------
# stage 5: group post-processing of the partition driver
# --------------------------------------------------------------------------------------------
def merge_groups(A, K, size_db, k, allelic_idx_set=(), min_allelic_overlap=0.1, allele_use_method="conflict", method="mean"):
    """hyperpartition.py:2288-2471 (_merge, alleletable_df=None): greedy merge of the pair of groups with the largest
    Q_ij = V_ij - V_i. V_j. / sum(triu V), V = M^T A M (mean: divided by the size product), until k groups remain."""
    K = [list(g) for g in K]
    A = A.tocsr()
    it = 0
    while k and len(K) > k:
        g = len(K)
        if it > (g - k + 100):
            break
        Kt = [tuple(sorted(x)) for x in K]
        glen = np.array([sum(size_db.get(c, 0.0) for c in x) for x in Kt])
        c2g = {c: i for i, x in enumerate(Kt) for c in x}
        flag = np.ones((g, g)); overlap = np.zeros((g, g))
        if allelic_idx_set:
            s1, s2 = {}, {}
            for c1, c2 in allelic_idx_set:
                g1, g2 = c2g.get(c1), c2g.get(c2)
                if g1 is not None and g2 is not None and g1 != g2:
                    if g1 > g2:
                        g1, g2, c1, c2 = g2, g1, c2, c1
                    s1.setdefault((g1, g2), set()).add(c1)
                    s2.setdefault((g1, g2), set()).add(c2)
            for (i, j), a in s1.items():
                l1 = float(sum(size_db.get(c, 0.0) for c in a))
                l2 = float(sum(size_db.get(c, 0.0) for c in s2.get((i, j), set())))
                o1 = l1 / glen[i] if glen[i] > 0 else 0
                o2 = l2 / glen[j] if glen[j] > 0 else 0
                if o1 > min_allelic_overlap or o2 > min_allelic_overlap:
                    flag[i, j] = 0
                    overlap[i, j] = max(o1, o2)
        rows = [c for x in Kt for c in x]
        cols = [i for i, x in enumerate(Kt) for _ in x]
        M = csr_matrix((np.ones(len(rows), dtype=np.float32), (rows, cols)), shape=(A.shape[0], g))
        V = (M.T.dot(A).dot(M)).toarray().astype(np.float64)
        np.fill_diagonal(V, 0)
        if method == "mean":
            sz = np.array([len(x) for x in Kt], dtype=np.float32)
            so = np.outer(sz, sz); so[so == 0] = 1.0
            V = V / so
        total = max(np.triu(V, 1).sum(), 1e-9)
        best, bi = None, None
        for i in range(g):
            vi = V[i].sum()
            for j in range(i + 1, g):
                raw = V[i, j] - (V[j].sum() * vi) / total
                if flag[i, j]:
                    q = raw
                elif allele_use_method == "conflict":
                    q = -2 ** 64
                else:
                    q = max(overlap[i, j] * raw, 0)
                if best is None or q > best:
                    best, bi = q, (i, j)
        i1, i2 = bi
        K[i1].extend(K[i2])
        K.pop(i2)
        it += 1
    return K


def recluster_by_linkage(A, K, contig_lens, min_weight, allelic_idx_set=(), min_ratio=1.5):
    """hyperpartition.py:2860-2968 (length_norm = False branch).  Returns (new K, number of moved contigs)."""
    A = A.tocsr()
    n, g = A.shape[0], len(K)
    c2g = np.full(n, -1, dtype=int)
    rows, cols = [], []
    for gid, grp in enumerate(K):
        c2g[list(grp)] = gid
        rows.extend(grp); cols.extend([gid] * len(grp))
    look = {}
    for c1, c2 in allelic_idx_set:
        look.setdefault(c1, set()).add(c2)
        look.setdefault(c2, set()).add(c1)
    M = csr_matrix((np.ones(len(rows), dtype=np.float32), (rows, cols)), shape=(n, g))
    S = A.dot(M).toarray()
    best = np.argmax(S, axis=1)
    bS = S[np.arange(n), best]
    cur = np.zeros(n)
    valid = c2g != -1
    cur[valid] = S[np.where(valid)[0], c2g[valid]]
    tot = S.sum(axis=1)
    spec = np.divide(bS, tot, out=np.zeros(n), where=tot > 0)
    mask = (best != c2g) & (bS > cur * min_ratio) & (spec > 0.9) & (tot > min_weight)
    cand = sorted(np.where(mask)[0], key=lambda x: contig_lens[x], reverse=True)
    newK = [set(x) for x in K]
    moved = 0
    for c in cand:
        bg, cg = best[c], c2g[c]
        if S[c, bg] / (tot[c] + 1e-9) < 0.25:
            continue
        if look.get(c, set()).isdisjoint(newK[bg]):
            if cg != -1:
                newK[cg].discard(c)
            newK[bg].add(c)
            c2g[c] = bg
            moved += 1
    return ([sorted(int(x) for x in s) for s in newK] if moved else [list(x) for x in K]), moved


# --------------------------------------------------------------------------------------------
# IRMM rounds 2.. (algorithms/hypergraph.py:1258-1397) -- restated line by line; the clustering itself is passed in
# --------------------------------------------------------------------------------------------
def irmm_reweight_round(H, H_T, W, membership, n_clusters, alias):
    """One pass of :1270-1317.  H: current (re-weighted) csr float32; H_T: the transpose IRMM keeps from its set-up
    (:1135); W: float32 vector (W.data[0] of the dia matrix); membership: int[n].
    Aliasing, restated explicitly: `H_T = H.T` (:1135) and `H.tocoo()` (:1298) both SHARE H's data array, so the first
    time monomers are zeroed (:1307) the zeros land in H_T as well; `eliminate_zeros()` + `tocsr()` (:1308-1310) then give
    H fresh arrays and later rounds leave H_T alone.  `alias` says whether H still is the set-up object.
    Returns (H_new, H_T_new, W_new, alias_new); nothing is modified in place."""
    n_contigs, n_hyperedges = H.shape
    c = int(n_clusters)
    C = csr_matrix((np.ones(n_contigs, dtype=np.float32), (np.arange(n_contigs), np.asarray(membership))),
                   shape=(n_contigs, c))
    K = (H_T @ C).toarray()
    e_c = np.array(H.sum(axis=0)).flatten()
    w_prev = W
    inv_k = 1.0 / (1.0 + K)
    inv_k_sum = inv_k.sum(axis=1)
    denom = np.maximum(e_c * 2.0, 1.0)
    w_new = inv_k_sum * (e_c + c) / denom
    max_cluster_counts = K.max(axis=1)
    purity = max_cluster_counts / np.maximum(e_c, 1.0)
    noisy = (purity >= 0.75) & (e_c > max_cluster_counts) & (e_c >= 3)
    if np.any(noisy):
        pure = np.where(noisy)[0]
        maj = K[pure].argmax(axis=1)
        read_to_maj = np.full(n_hyperedges, -1, dtype=np.int32)
        read_to_maj[pure] = maj
        Hr = H.tocoo(copy=True)
        mono = np.asarray(membership, dtype=np.int32)[Hr.row]
        exp = read_to_maj[Hr.col]
        noise = (exp >= 0) & (mono != exp)
        if np.any(noise):
            if alias:        # the same zeros appear in H_T (stored explicitly there, which changes no product)
                T = H_T.tocoo(copy=True)
                exp_t = read_to_maj[T.row]
                T.data[(exp_t >= 0) & (np.asarray(membership, dtype=np.int32)[T.col] != exp_t)] = 0.0
                T.eliminate_zeros()
                H_T = T.tocsr()
            Hr.data[noise] = 0.0
            Hr.eliminate_zeros()
        H = Hr.tocsr()
        alias = False
    w_new = w_new * np.where((e_c >= 4) & (purity < 0.4), 0.1, 1.0)
    w_updated = 0.5 * (w_new + w_prev)
    W_out = np.where(e_c >= 2, w_updated, w_prev).astype(np.float32)
    return H, H_T, W_out, alias


def irmm_rounds(H, cluster_fn, NW=None, P_allelic=None, P_weak=None, allelic_factor=-1.0, resolution=1.0,
                min_weight=1.0, threshold=0.01, max_round=3):
    """IRMM with max_round > 1 (:1120-1150 setup, :1157-1256 first clustering, :1258-1397 rounds).
    cluster_fn(A_upper_symmetric csr, resolution, initial_membership or None) -> membership int[n].
    NW: the VECTOR w (NW matrix = float32 outer(w, w)).  Returns (raw_A, A, membership, history) where history holds
    per round (W, H.nnz, ari_to_previous)."""
    from scipy.sparse import dia_matrix, diags, triu as _triu
    H = H.tocsr().astype(np.float32)
    n, m = H.shape
    W = np.ones(m, dtype=np.float32)
    with np.errstate(over="ignore"):
        D_e_num = H.getnnz(axis=0).astype(ORDER_DTYPE) - ORDER_DTYPE(1)
    d = 1.0 / np.maximum(D_e_num, 1)

    def nw_scale(A):
        if NW is None:
            return A
        A = A.tocoo()
        nwm = np.outer(NW, NW).astype(np.float32)
        A.data = A.data * nwm[A.row, A.col]
        return A.tocsr()

    def prep(A):
        A = A.tocsr().copy()
        if P_allelic is not None and len(P_allelic[0]):
            off = np.unique(find_csr_data_offsets(A, np.asarray(P_allelic[0]), np.asarray(P_allelic[1])))
            A.data[off] *= allelic_factor
        if P_weak is not None and len(P_weak[0]):
            off = np.unique(find_csr_data_offsets(A, np.asarray(P_weak[0]), np.asarray(P_weak[1])))
            A.data[off] = 0
        A.setdiag(0)
        return A

    def graph(A):
        U = _triu(A, format="csr")
        U.eliminate_zeros()
        return (U + U.T).tocsr()

    from scipy.sparse import identity

    def product(Hc, Wv, HTc):
        # `H @ W @ D_e_inv @ H_T` exactly as written (:1136-1141, 1321): identity(float32) with W on its diagonal, a float64
        # dia_matrix, left to right -- the float64 sums then run in scipy's own order, which decides exact threshold ties
        Wm = identity(m, dtype=np.float32)
        Wm.data[0] = Wv
        return (Hc @ Wm @ dia_matrix((d, [0]), shape=(m, m)) @ HTc).tocsr()
    H_T = H.T                      # csc sharing H's arrays, as in the reference (never modified in place here)
    A = product(H, W, H_T)
    if min_weight > 0:
        A.data[A.data < min_weight] = 0
        A.eliminate_zeros()
    A = nw_scale(A)
    A.setdiag(0)
    A.eliminate_zeros()
    raw_A = A.copy()
    A = prep(A)
    membership = np.asarray(cluster_fn(graph(A), resolution, None))
    hist = []
    W_prev = W.copy()
    it, diff, prev_mem = 1, 1.0, None
    alias = True
    while diff > threshold:
        if it >= max_round:
            break
        c = int(membership.max()) + 1
        H, H_T, W, alias = irmm_reweight_round(H, H_T, W, membership, c, alias)
        # `H @ W @ D_e_inv @ H_T` (:1321) runs left to right: H @ W is a float32 product, D_e_inv is float64
        A_new = product(H, W, H_T)
        if min_weight > 0:
            A_new.data[A_new.data < min_weight] = 0
            A_new.eliminate_zeros()
        A = (A * 0.5 + A_new * 0.5).tocsr()
        A = nw_scale(A)
        A.setdiag(0)
        A.eliminate_zeros()
        A = prep(A)
        membership = np.asarray(cluster_fn(graph(A), resolution, membership))
        a = None
        if prev_mem is not None:
            from oracle.community_oracle import ari as _ari
            a = _ari(prev_mem, membership)
        hist.append((W.copy(), H.nnz, a))
        if a is not None and a > 0.99:
            break
        prev_mem = membership.copy()
        diff = float(np.max(np.abs(W - W_prev)))
        W_prev = W.copy()
        it += 1
    return raw_A, A, membership, hist
