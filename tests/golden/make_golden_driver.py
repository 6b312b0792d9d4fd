#!/usr/bin/env python
"""Golden groups from the REFERENCE's own partition driver: ``HyperPartition.single_partition`` (hyperpartition.py:911-1237)
and ``incremental_partition`` / ``_incremental_partition`` (:1726-2286, :1239-1560), run read-only from /root/reference
(or the install under baseline/_ref) on the planted 5 x 4 genome.

    python tests/golden/make_golden_driver.py        # build container only

Everything the driver does -- filters, the resolution auto-search with its step sizes, caps and back-off, filter_cluster,
_merge, recluster_by_linkage, the mapping of prune pairs into sub-graphs, the second round over the first-round groups
(joblib runs it in-process at threads = 1) -- is the reference's code on the reference's HyperGraph / IRMM.  What cannot
run here is python-igraph: as in make_golden_irmm.py a stand-in ``igraph.Graph`` answers community_multilevel /
community_leiden with the oracle's sequential C clustering.  The prune table fed to the `prunetable` cases is written by
the oracle's restatement of KPruneHyperGraph (itself pinned to a reference run, ref_kprune.npz).

Two compatibility shims, neither touching the algorithm: pandas 3 hands out read-only ``to_numpy()`` views that the
reference (pinned to pandas 2.3) writes into (algorithms/hypergraph.py:549), and ``ultra_complex=0`` because the
reference's ultra-complex branch reads an attribute it only sets with a prune table and an undefined ``_results2``
(:1129-1146).
"""
import contextlib
import io
import os
import sys
import tempfile

import numpy as np
import pandas as pd

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(ROOT, "baseline"))

N_CONTIGS, CHROMS, PLOIDY, M, SEED = 200, 5, 4, 60000, 41
KW = dict(min_contacts=5, min_weight=1.0, min_cis_weight=5, resolution1=-1, init_resolution1=1.0, resolution2=-1,
          init_resolution2=1.0, min_scaffold_length=10000, threads=1, min_quality1=1, min_quality2=1)
CASES = {
    # name: (method, extra constructor kwargs, uses the prune table)
    "single": ("single", dict(allelic_factor=0, cross_allelic_factor=0), False),
    "single_prune_leiden": ("single", dict(allelic_factor=-1, cross_allelic_factor=0.3), True),
    "incremental": ("incremental", dict(allelic_factor=0, cross_allelic_factor=0), False),
    "incremental_prune": ("incremental", dict(allelic_factor=-1, cross_allelic_factor=0.3), True),
}


def pandas_writable_shim():
    for cls in (pd.Series, pd.Index):
        orig = cls.to_numpy

        def to_numpy(self, *a, _o=orig, **k):
            x = _o(self, *a, **k)
            return x if x.flags.writeable else x.copy()
        cls.to_numpy = to_numpy


def inputs():
    """-> genome, triplets (row, col, count, mapq), prune-table DataFrame (names)"""
    from cphasing_b200 import synth
    from oracle import hypergraph_oracle as orc
    g = synth.make_genome(N_CONTIGS, CHROMS, PLOIDY, seed=SEED)
    t = synth.porec_table(g, M, seed=SEED + 1)
    trip = orc.porec_edges(t["read_idx"], t["contig_idx"], t["mapq"], min_mapq=1, n_contigs=g.n)
    return g, trip


def prune_table(g, trip):
    """7-column prune table (core.py:711-723) of the oracle's KPruneHyperGraph restatement on the unweighted A"""
    from cphasing_b200 import synth
    from oracle import hypergraph_oracle as orc
    rr, cc, kk, ncol = orc.hypergraph_get_data(*trip, 1, mapq_filter=False)
    H, _, keep = orc.incidence_matrix(rr, cc, kk, g.n, ncol, KW["min_contacts"])
    assert H.shape[0] == g.n
    Au = orc.clique_expansion(H, use_high_order=False, min_weight=KW["min_weight"])
    al = synth.allele_pairs(g, seed=SEED + 2)
    meta = {(int(a), int(b)): (m1, m2, sh, sim) for a, b, m1, m2, sh, sim in al}
    allelic, weak = orc.kprune_hypergraph(Au, [(int(a), int(b)) for a, b, *_ in al])
    rows = []
    for a, b in allelic:
        m1, m2, sh, sim = meta.get((a, b), meta.get((b, a), (0, 0, 0, 0)))
        rows.append((g.names[a], g.names[b], m1, m2, sh, sim, 0))
    for a, b in weak:
        rows.append((g.names[a], g.names[b], 0, 0, 0, 0, 1))
    return pd.DataFrame(rows, columns=['contig1', 'contig2', 'mz1', 'mz2', 'mzShared', 'similarity', 'type'])


def labels(hp, g):
    lab = np.full(g.n, -1, dtype=np.int64)
    idx = {nm: i for i, nm in enumerate(g.names)}
    for gi, grp in enumerate(hp.K):
        for v in grp:
            lab[idx[hp.vertices[v]]] = gi
    return lab


def main():
    pandas_writable_shim()
    import ref_loader
    ah = ref_loader.load()
    import make_golden_irmm as mgi
    import igraph
    igraph.Graph = mgi._Graph
    import cphasing.algorithms.hypergraph as rah
    import cphasing.hyperpartition as rhp
    rah.ig = igraph
    g, trip = inputs()
    sizes = pd.DataFrame({"length": pd.Series(g.contigsizes)})
    pt = prune_table(g, trip)
    out = {"planted": g.group.astype(np.int64), "n_allelic": np.int64((pt["type"] == 0).sum()),
           "n_weak": np.int64((pt["type"] == 1).sum())}
    with tempfile.TemporaryDirectory() as wd:
        os.chdir(wd)
        ptf = os.path.join(wd, "prune.table")
        pt.to_csv(ptf, sep="\t", header=False, index=False)
        for name, (method, extra, use_pt) in CASES.items():
            edges = ah.HyperEdges(idx=g.contig_idx, row=trip[0].copy(), col=trip[1].copy(), count=trip[2].copy(),
                                  contigsizes=g.contigsizes, mapq=trip[3].copy())
            with contextlib.redirect_stdout(io.StringIO()):
                hp = rhp.HyperPartition(edges, sizes, k=[CHROMS, PLOIDY], prunetable=ptf if use_pt else None,
                                        ultra_complex=0, **KW, **extra)
                if method == "single":
                    hp.single_partition(k=CHROMS * PLOIDY)
                else:
                    hp.incremental_partition([CHROMS, PLOIDY])
            lab = labels(hp, g)
            out[f"{name}_labels"] = lab
            out[f"{name}_sizes"] = np.array([len(x) for x in hp.K], dtype=np.int64)
            if method == "incremental":
                out[f"{name}_inc_chr_idx"] = np.array(hp.inc_chr_idx, dtype=np.int64)
            print(name, len(hp.K), "groups; sizes", sorted(len(x) for x in hp.K))
        os.chdir(ROOT)
    np.savez_compressed(os.path.join(HERE, "ref_driver.npz"), **out)


if __name__ == "__main__":
    main()
