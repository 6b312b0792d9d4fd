#!/usr/bin/env python
"""Golden vectors for the partition driver's post-processing, produced by the REFERENCE's own
``HyperPartition._merge`` (hyperpartition.py:2288-2471), ``recluster_by_linkage`` (:2860-2968) and
``filter_cluster`` (:4012-4044), stub-imported read-only from /root/reference (see make_golden.py).

    python tests/golden/make_golden_partition.py        # build container only

Input: A = the reference's clique expansion of a seeded synthetic Pore-C table (5 chromosomes x 4 haplotypes,
120 contigs); K = the planted groups, each split in two (so _merge has work), plus a few contigs deliberately put in a
wrong group (so recluster_by_linkage has work).  Inputs and outputs are stored; tests never need the reference.
"""
import os
import sys
import types

import numpy as np
import pandas as pd

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
import make_golden as mg          # noqa: E402


def labels_of(K, n):
    lab = np.full(n, -1, dtype=np.int64)
    for i, g in enumerate(K):
        lab[list(g)] = i
    return lab


def main():
    ah = mg.import_reference()
    import cphasing.hyperpartition as hp
    from cphasing_b200 import synth
    from oracle import hypergraph_oracle as orc
    g = synth.make_genome(120, 5, 4, seed=41)
    t = synth.porec_table(g, 30000, seed=42)
    r, c, cnt, q = orc.porec_edges(t["read_idx"], t["contig_idx"], t["mapq"], min_mapq=1, n_contigs=g.n)
    idx = {nm: i for i, nm in enumerate(g.names)}
    sizes = {nm: int(l) for nm, l in zip(g.names, g.length)}
    edges = ah.HyperEdges(idx=idx, row=r.copy(), col=c.copy(), count=cnt.copy(), contigsizes=sizes, mapq=q.copy())
    HG = ah.HyperGraph(edges, min_quality=1, mapq_filter=False)
    H = HG.incidence_matrix(min_contacts=5)
    assert H.shape[0] == g.n, "the case keeps every contig so vertex ids are contig ids"
    A = ah.HyperGraph.clique_expansion_init(H, min_weight=1.0)
    n = A.shape[0]
    out = {"in_row": r, "in_col": c, "in_count": cnt, "in_mapq": q, "lengths": g.length.astype(np.int64),
           "group": g.group.astype(np.int64)}
    out.update(mg.csr_parts("A", A))
    # over-split planted groups
    K0 = []
    for gi in range(g.G):
        mem = np.where(g.group == gi)[0]
        K0.append([int(x) for x in mem[::2]])
        K0.append([int(x) for x in mem[1::2]])
    out["K0_labels"] = labels_of(K0, n)
    vsz = pd.DataFrame({"length": {i: int(g.length[i]) for i in range(n)}})
    for method in ("mean", "sum"):
        K1 = hp.HyperPartition._merge(A, [list(x) for x in K0], vsz, k=g.G, allelic_idx_set=set(), method=method)
        out[f"merge_{method}_labels"] = labels_of(K1, n)
        out[f"merge_{method}_order"] = np.array([x[0] for x in K1], dtype=np.int64)     # first member: list order
    # allelic conflicts veto merges between homologous haplotypes (type-0 pairs of a prune table)
    al = synth.allele_pairs(g, seed=43)
    aset = set((int(a), int(b)) for a, b, *_ in al)
    K2 = hp.HyperPartition._merge(A, [list(x) for x in K0], vsz, k=g.G, allelic_idx_set=aset, min_allelic_overlap=0.1,
                                  allele_use_method="conflict", method="mean")
    out["allelic_pairs"] = np.array(sorted(aset), dtype=np.int64)
    out["merge_conflict_labels"] = labels_of(K2, n)
    # recluster_by_linkage: planted groups with 6 contigs misplaced, two of them left in no group at all
    rng = np.random.default_rng(44)
    K3 = [[int(x) for x in np.where(g.group == gi)[0]] for gi in range(g.G)]
    moved = rng.choice(n, size=6, replace=False)
    for j, v in enumerate(moved):
        src = int(g.group[v])
        K3[src].remove(int(v))
        if j >= 4:
            continue                                     # in no group
        K3[(src + 5) % g.G].append(int(v))
    out["K3_labels"] = labels_of(K3, n)
    fake = types.SimpleNamespace(K=[list(x) for x in K3], allelic_idx_set=set(), min_weight=1.0, H=None,
                                 idx_to_vertices={i: nm for i, nm in enumerate(g.names)}, contig_sizes=sizes)
    # (specificity > 0.9 needs a clean signal: use the expansion thresholded at 40, as a high --min-weight run would)
    Ar = ah.HyperGraph.clique_expansion_init(H, min_weight=40.0)
    out.update(mg.csr_parts("Ar", Ar))
    cnt_moved = hp.HyperPartition.recluster_by_linkage(fake, A=Ar, min_ratio=1.5, iter_count=1)
    out["recluster_count"] = np.array([cnt_moved], dtype=np.int64)
    out["recluster_labels"] = labels_of(fake.K, n)
    # filter_cluster at 8 Mb
    fake2 = types.SimpleNamespace(K=[list(x) for x in K0], vertices_idx_sizes={i: int(g.length[i]) for i in range(n)},
                                  min_scaffold_length=8_000_000)
    Kf = hp.HyperPartition.filter_cluster(fake2, verbose=0)
    out["filter_cluster_kept"] = np.array([K0.index(x) for x in Kf], dtype=np.int64)
    np.savez_compressed(os.path.join(HERE, "ref_partition_post.npz"), **out)
    print(f"partition_post: n={n} nnz(A)={A.nnz} merge mean groups={len(set(out['merge_mean_labels']))} "
          f"conflict groups={len(set(out['merge_conflict_labels']))} reclustered={cnt_moved} filter kept={len(Kf)}")


if __name__ == "__main__":
    main()
