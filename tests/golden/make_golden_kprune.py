#!/usr/bin/env python
"""Golden vectors for KPruneHyperGraph (cphasing/kprune.py:385-552) produced by running the REFERENCE's own class
read-only from /root/reference (stub import as in make_golden.py).

    python tests/golden/make_golden_kprune.py        # build container only

python-igraph cannot be installed here; the only thing the class asks of it is the maximum-weight matching of a 2 x 2
bipartite graph (`is_strong_contact`, :473-487).  A stand-in `igraph.Graph.Bipartite` answers that by comparing the two
perfect matchings (equal sums count as "straight", i.e. a strong contact -- igraph's own choice on ties is not
specified).  The allele-group table, the candidate pairs from triu(A), the score look-ups, the allelic zeroing and the
prune table are the reference's own code.
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


class _Matching:
    def __init__(self, straight):
        self.straight = straight

    def match_of(self, v):
        return {0: 2, 1: 3, 2: 0, 3: 1}[v] if self.straight else {0: 3, 1: 2, 2: 1, 3: 0}[v]


class _Bip:
    def __init__(self, types_, edges):
        self.edges = list(edges)
        self.es = {}

    def maximum_bipartite_matching(self, weights="weight"):
        w = dict(zip(self.edges, self.es[weights]))
        straight = w[(0, 2)] + w[(1, 3)]
        cross = w[(0, 3)] + w[(1, 2)]
        return _Matching(straight >= cross)


def main():
    ig = types.ModuleType("igraph")
    ig.Graph = types.SimpleNamespace(Bipartite=_Bip)
    sys.modules["igraph"] = ig
    ah = mg.import_reference()
    import cphasing.kprune as kp
    kp.ig = ig
    from cphasing_b200 import synth
    from oracle import hypergraph_oracle as orc
    g = synth.make_genome(160, 4, 4, seed=61)
    t = synth.porec_table(g, 40000, seed=62)
    r, c, cnt, q = orc.porec_edges(t["read_idx"], t["contig_idx"], t["mapq"], min_mapq=1, n_contigs=g.n)
    idx = {nm: i for i, nm in enumerate(g.names)}
    sizes = {nm: int(l) for nm, l in zip(g.names, g.length)}
    edges = ah.HyperEdges(idx=idx, row=r.copy(), col=c.copy(), count=cnt.copy(), contigsizes=sizes, mapq=q.copy())
    HG = ah.HyperGraph(edges, min_quality=1, mapq_filter=False)
    H = HG.incidence_matrix(min_contacts=5)
    assert H.shape[0] == g.n
    A = ah.HyperGraph.clique_expansion_init(H, min_weight=1.0, use_high_order=False)
    out = {}
    out.update(mg.csr_parts("A", A))
    al = synth.allele_pairs(g, seed=63)
    rows = [(0, 0, g.names[a], g.names[b], m1, m2, sh, sim, 1) for a, b, m1, m2, sh, sim in al]
    df = pd.DataFrame(rows, columns=[-1, 0, 1, 2, 3, 4, 5, 6, 7]).set_index(-1)
    at = types.SimpleNamespace(data=df)
    out["allele_pairs"] = np.array([(a, b) for a, b, *_ in al], dtype=np.int64)
    out["allele_meta"] = np.array([(m1, m2, sh, sim) for _, _, m1, m2, sh, sim in al], dtype=np.float64)
    import io, contextlib
    with contextlib.redirect_stdout(io.StringIO()):          # the constructor prints the allele-group dict (:426)
        k = kp.KPruneHyperGraph(at, A.copy(), idx)
    k.run()
    out["allelic"] = np.array(sorted(map(tuple, k.allelic_prune_list)), dtype=np.int64).reshape(-1, 2)
    out["weak"] = np.array(sorted(map(tuple, k.weak_prune_list)), dtype=np.int64).reshape(-1, 2)
    grp = k.allele_group_db
    out["allele_group"] = np.array(sorted((int(a), int(b)) for a, b in grp.items()), dtype=np.int64)
    print("allelic", len(out["allelic"]), "weak", len(out["weak"]), "self pairs in weak",
          int((out["weak"][:, 0] == out["weak"][:, 1]).sum()))
    # second case: the contig index does NOT cover the whole allele table (what HyperPartition hands over after its
    # min_contacts / min_length / min_cis_weight filters).  A contig whose FIRST listed partner is missing gets no allele
    # group -- a later, surviving partner does not step in (kprune.py:437-447).
    rng = np.random.default_rng(64)
    first_partner = {}
    for a, b, *_ in al:
        first_partner.setdefault(int(a), int(b))
    multi = [a for a in first_partner if sum(1 for x, *_ in al if int(x) == a) > 1]
    drop = set(int(first_partner[a]) for a in rng.choice(multi, size=min(10, len(multi)), replace=False))
    drop |= set(int(x) for x in rng.choice(g.n, size=4, replace=False))
    keep = np.array([i for i in range(g.n) if i not in drop], dtype=np.int64)
    sub_idx = {g.names[i]: j for j, i in enumerate(keep)}
    A_sub = A[keep][:, keep].tocsr()
    with contextlib.redirect_stdout(io.StringIO()):
        k2 = kp.KPruneHyperGraph(at, A_sub.copy(), sub_idx)
    k2.run()
    out["sub_keep"] = keep
    out["sub_allelic"] = np.array(sorted(map(tuple, k2.allelic_prune_list)), dtype=np.int64).reshape(-1, 2)
    out["sub_weak"] = np.array(sorted(map(tuple, k2.weak_prune_list)), dtype=np.int64).reshape(-1, 2)
    out["sub_allele_group"] = np.array(sorted((int(a), int(b)) for a, b in k2.allele_group_db.items()), dtype=np.int64)
    n_lost = sum(1 for a in first_partner if a not in drop and first_partner[a] in drop)
    print("subset case: kept", len(keep), "contigs;", n_lost, "contigs lose their group; allelic", len(out["sub_allelic"]),
          "weak", len(out["sub_weak"]))
    try:
        pt = k.get_prune_table()
        out["table_shape"] = np.array(pt.shape, dtype=np.int64)
        out["table_types"] = pt["type"].to_numpy().astype(np.int64) if "type" in pt else np.zeros(0, np.int64)
        print("prune table", pt.shape, list(pt.columns))
    except Exception as e:                                   # recorded, not fatal: the lists are the contract
        print("get_prune_table failed in the reference:", repr(e))
    path = os.path.join(HERE, "ref_kprune.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
