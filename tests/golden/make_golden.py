#!/usr/bin/env python
"""Generate golden vectors by running the REFERENCE's own code (read-only, /root/reference).

Run in the build container only (the GPU box has no /root/reference):
    python tests/golden/make_golden.py

The reference imports polars / igraph / cooler / ... at module import time; none of them is
installable here, so they are replaced by inert stubs.  Only reference functions that never touch
those modules are executed:
    HyperEdges, HyperGraph.get_data / remove_rows / incidence_matrix / clique_expansion_init /
    normalize_adjacency_matrix / filter_adjacency_matrix, extract_incidence_matrix2,
    find_csr_data_offsets, HyperPartition.get_normalize_weight.
Inputs (hyperedge triplets) come from the seeded synthetic tables of cphasing_b200.synth pushed
through oracle.hypergraph_oracle.porec_edges / pairs_edges; inputs AND reference outputs are
stored, so the tests never need the reference.
"""
import os
import sys
import types

import numpy as np
import pandas as pd

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)


class _Stub(types.ModuleType):
    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        return _Stub(self.__name__ + "." + name)

    def __call__(self, *a, **k):
        return self


def import_reference():
    for m in ["igraph", "polars", "cooler", "pytools", "natsort", "pyranges", "pandarallel", "Bio",
              "Bio.SeqIO", "Bio.Seq", "rich_click", "pyfaidx", "pysam", "dask", "dask.array", "hicmatrix",
              "portion", "line_profiler", "seaborn", "colormaps", "intervaltree", "ncls", "pypdf", "pyfastx"]:
        if m not in sys.modules:
            try:
                __import__(m)
            except Exception:
                sys.modules[m] = _Stub(m)
    sys.path.insert(0, "/root/reference")
    import cphasing.algorithms.hypergraph as ah
    return ah


def csr_parts(prefix, M):
    M = M.tocsr().copy()
    M.sort_indices()
    return {prefix + "_indptr": M.indptr.astype(np.int64), prefix + "_indices": M.indices.astype(np.int32),
            prefix + "_data": M.data, prefix + "_shape": np.array(M.shape, dtype=np.int64)}


def run_case(ah, name, row, col, count, mapq, names, lengths, min_quality, min_contacts, min_weight,
             remove=()):
    idx = {c: i for i, c in enumerate(names)}
    sizes = {c: int(l) for c, l in zip(names, lengths)}
    edges = ah.HyperEdges(idx=idx, row=row.copy(), col=col.copy(), count=count.copy(), contigsizes=sizes,
                          mapq=mapq.copy())
    HG = ah.HyperGraph(edges, min_quality=min_quality, mapq_filter=True)
    out = {"in_row": row, "in_col": col, "in_count": count, "in_mapq": mapq,
           "lengths": np.asarray(lengths, dtype=np.int64),
           "params": np.array([min_quality, min_contacts, min_weight], dtype=np.float64),
           "gd_row": HG.row.copy(), "gd_col": HG.col.copy(), "gd_count": HG.count.copy(),
           "gd_shape": np.array(HG.shape, dtype=np.int64)}
    if len(remove):
        HG.remove_rows([names[i] for i in remove])
        out["remove"] = np.asarray(remove, dtype=np.int64)
    H = HG.incidence_matrix(min_contacts=min_contacts)
    out.update(csr_parts("H", H))
    out["H_nodes_idx"] = np.array([idx[c] for c in HG.nodes], dtype=np.int64)
    out["H_remove_contig_idx"] = np.asarray(HG.remove_contig_idx, dtype=np.int64)
    # clique expansion: weighted (reference float32), unweighted, with W
    A = ah.HyperGraph.clique_expansion_init(H, min_weight=min_weight)
    out.update(csr_parts("A", A))
    A0 = ah.HyperGraph.clique_expansion_init(H, min_weight=0)
    out.update(csr_parts("A0", A0))
    Au = ah.HyperGraph.clique_expansion_init(H, min_weight=min_weight, use_high_order=False)
    out.update(csr_parts("Au", Au))
    from scipy.sparse import dia_matrix
    rng = np.random.default_rng(7)
    wv = rng.uniform(0.5, 1.5, size=H.shape[1]).astype(np.float32)
    W = dia_matrix((wv, np.array([0])), shape=(H.shape[1], H.shape[1]), dtype=np.float32)
    Aw = ah.HyperGraph.clique_expansion_init(H, min_weight=min_weight, W=W)
    out["W"] = wv
    out.update(csr_parts("Aw", Aw))
    # normalisation (HyperPartition.get_normalize_weight, hard-wired 'svc')
    import cphasing.hyperpartition as hp
    nodes = HG.nodes
    # pandas 3 (copy-on-write) hands out read-only ``.values``; the reference (pinned to pandas 2.3)
    # writes into it (hyperpartition.py:531), so give it a writable stand-in for
    # ``contigsizes.loc[vertices]['length'].astype(float64).values``.
    class _Vals:
        def __init__(self, v): self.values = v
        def astype(self, dt): return _Vals(self.values.astype(dt))
    class _Loc:
        def __getitem__(self, v): return {"length": _Vals(np.array([sizes[c] for c in v]))}
    fake = types.SimpleNamespace(contigsizes=types.SimpleNamespace(loc=_Loc()), vertices=nodes, H=H)
    NW = hp.HyperPartition.get_normalize_weight(fake)
    out["NW_diag_sqrt"] = np.sqrt(np.diag(NW).astype(np.float64))
    An = ah.HyperGraph.normalize_adjacency_matrix(A.copy(), NW)
    out.update(csr_parts("An", An))
    # degree-outlier filter
    keep = ah.HyperGraph.filter_adjacency_matrix(A, nodes, sizes)
    out["filter_keep"] = keep
    # sub-incidence extraction on a row subset
    sub_idx = np.arange(0, H.shape[0], 3)
    subH, rem, kept = ah.extract_incidence_matrix2(H, sub_idx)
    out["sub_idx"] = sub_idx
    out.update(csr_parts("subH", subH))
    out["sub_kept"] = np.asarray(kept, dtype=np.int64)
    # pair re-weighting through find_csr_data_offsets (IRMM :1161-1170)
    coo = A.tocoo()
    pick = rng.choice(coo.nnz, size=min(20, coo.nnz), replace=False)
    pr, pc = coo.row[pick], coo.col[pick]
    out["pair_r"] = pr.astype(np.int64); out["pair_c"] = pc.astype(np.int64)
    out["pair_offsets"] = np.sort(ah.find_csr_data_offsets(A.tocsr(), pr, pc))
    np.savez_compressed(os.path.join(HERE, f"ref_{name}.npz"), **out)
    print(f"{name}: n={len(names)} nnz_in={len(row)} H={H.shape} nnz(H)={H.nnz} nnz(A)={A.nnz}")


def main():
    ah = import_reference()
    from cphasing_b200 import synth
    from oracle import hypergraph_oracle as orc

    # 1. hand-checkable Hi-C toy of SURVEY.md §4: pairs (0,1),(0,0),(1,2),(0,1) -> A=[[6,2,0],[2,3,1],[0,1,1]]
    r, c, cnt, q = orc.pairs_edges([0, 0, 1, 0], [1, 0, 2, 1], [60, 60, 60, 60], min_mapq=1)
    run_case(ah, "toy_hic", r, c, cnt, q, ["a", "b", "c"], [100000, 200000, 300000], 1, 0, 0.0)

    # 2. 3-contig Pore-C toy with counts
    r = np.array([0, 0, 1, 2], dtype=np.int32); c = np.array([0, 1, 0, 1], dtype=np.int32)
    cnt = np.array([1, 1, 1, 3], dtype=np.uint32); q = np.array([60, 60, 60, 60], dtype=np.int8)
    run_case(ah, "toy_porec", r, c, cnt, q, ["a", "b", "c"], [100000, 200000, 300000], 1, 0, 0.0)

    # 3. synthetic Pore-C, planted 3 chr x 2 hap, with filters active
    g = synth.make_genome(60, 3, 2, seed=11)
    t = synth.porec_table(g, 4000, seed=12)
    r, c, cnt, q = orc.porec_edges(t["read_idx"], t["contig_idx"], t["mapq"], min_mapq=1, min_order=2,
                                   max_order=50, n_contigs=g.n)
    run_case(ah, "porec_small", r, c, cnt, q, g.names, g.length, 1, 25, 2.0, remove=(3, 17))

    # 4. higher order (exp law), mapq filter 10
    g = synth.make_genome(120, 5, 4, seed=21)
    t = synth.porec_table(g, 6000, seed=22, law="exp")
    r, c, cnt, q = orc.porec_edges(t["read_idx"], t["contig_idx"], t["mapq"], min_mapq=0, min_order=2,
                                   max_order=50, n_contigs=g.n)
    run_case(ah, "porec_highorder", r, c, cnt, q, g.names, g.length, 10, 25, 0.1)

    # 5. Hi-C pairs with intra-contig duplicates
    g = synth.make_genome(40, 4, 2, seed=31)
    t = synth.pairs_table(g, 20000, seed=32)
    r, c, cnt, q = orc.pairs_edges(t["chrom1"], t["chrom2"], t["mapq"], min_mapq=1, n_contigs=g.n)
    run_case(ah, "pairs_small", r, c, cnt, q, g.names, g.length, 1, 25, 2.0)


if __name__ == "__main__":
    main()
