#!/usr/bin/env python
"""Golden vectors for IRMM (algorithms/hypergraph.py:1067-1399), max_round = 1 and max_round > 1, produced by running
the REFERENCE's own IRMM read-only from /root/reference (see make_golden.py for the stub import).

    python tests/golden/make_golden_irmm.py        # build container only

python-igraph cannot be installed here, so `igraph` is replaced by a small stand-in whose Graph records every graph
the reference hands to it (vertex count, edge list, weights -- i.e. the output of the reference's graph preparation and
of every re-weighting round) and answers community_multilevel / community_leiden with the oracle's sequential C
clustering (oracle/community.c).  Everything between two clustering calls -- K = H^T C, purity rules, W update,
A_new = H' W D_e^-1 H^T, damping, normalisation, pair re-weighting, triu -- is the reference's own arithmetic.
"""
import os
import sys
import types

import numpy as np
from scipy.sparse import csr_matrix

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
import make_golden as mg          # noqa: E402

CALLS = []


class _Clustering:
    def __init__(self, membership):
        self.membership = [int(x) for x in membership]

    def __iter__(self):
        m = np.asarray(self.membership)
        return iter([np.where(m == c)[0].tolist() for c in range(m.max() + 1)])

    def __len__(self):
        return int(max(self.membership)) + 1


class _Graph:
    def __init__(self, n=0, edges=None, edge_attrs=None, directed=False):
        self.n = int(n)
        self.edges = np.asarray(edges, dtype=np.int64).reshape(-1, 2)
        self.w = np.asarray(edge_attrs["weight"], dtype=np.float64)

    def _sym(self):
        r, c = self.edges[:, 0], self.edges[:, 1]
        return csr_matrix((np.concatenate([self.w, self.w]), (np.concatenate([r, c]), np.concatenate([c, r]))),
                          shape=(self.n, self.n))

    def _run(self, resolution, method, initial):
        from oracle import community_oracle as co
        mem, _ = co.community(self._sym(), resolution, method, initial=initial)
        CALLS.append({"n": self.n, "edges": self.edges.copy(), "w": self.w.copy(), "membership": mem.copy(),
                      "method": method, "resolution": float(resolution)})
        return _Clustering(mem)

    def community_multilevel(self, weights=None, resolution=1.0, **kw):
        return self._run(resolution, "louvain", None)

    def community_leiden(self, weights=None, resolution_parameter=1.0, objective_function="modularity",
                         initial_membership=None, **kw):
        return self._run(resolution_parameter, "leiden", initial_membership)


def fake_igraph():
    m = types.ModuleType("igraph")
    m.Graph = _Graph
    m.set_random_number_generator = lambda r: None
    return m


def main():
    sys.modules["igraph"] = fake_igraph()
    ah = mg.import_reference()
    ah.ig = sys.modules["igraph"]
    # the reference calls sklearn's adjusted_rand_score (:1381); sklearn's input checks probe sys.modules["polars"], which
    # is a stub here -- take the stub out again now that the reference is imported
    if type(sys.modules.get("polars")).__name__ == "_Stub":
        sys.modules.pop("polars")
    from cphasing_b200 import synth
    from oracle import hypergraph_oracle as orc
    g = synth.make_genome(160, 4, 4, seed=51)
    t = synth.porec_table(g, 50000, seed=52)
    r, c, cnt, q = orc.porec_edges(t["read_idx"], t["contig_idx"], t["mapq"], min_mapq=1, n_contigs=g.n)
    idx = {nm: i for i, nm in enumerate(g.names)}
    sizes = {nm: int(l) for nm, l in zip(g.names, g.length)}
    edges = ah.HyperEdges(idx=idx, row=r.copy(), col=c.copy(), count=cnt.copy(), contigsizes=sizes, mapq=q.copy())
    HG = ah.HyperGraph(edges, min_quality=1, mapq_filter=False)
    H = HG.incidence_matrix(min_contacts=5)
    assert H.shape[0] == g.n
    out = {"in_row": r, "in_col": c, "in_count": cnt, "in_mapq": q, "lengths": g.length.astype(np.int64),
           "group": g.group.astype(np.int64)}
    out.update(mg.csr_parts("H", H))
    al = synth.allele_pairs(g, seed=53)
    pa = np.array([(a, b) for a, b, *_ in al] + [(b, a) for a, b, *_ in al], dtype=np.int64)
    out["allelic_pairs"] = pa
    import pandas as pd
    cases = {"louvain": dict(method="louvain", allelic_factor=0, P=None),
             "leiden": dict(method="leiden", allelic_factor=-1, P=[pd.Series(pa[:, 0]), pd.Series(pa[:, 1])])}
    for name, kw in cases.items():
        for max_round in (1, 4):
            CALLS.clear()
            raw_A, A, ca, K = ah.IRMM(H.copy(), None, None, kw["P"], None, kw["allelic_factor"], 0.3, 1.6, 1.0, 0.01,
                                      max_round, method=kw["method"])
            tag = f"{name}_r{max_round}"
            out.update(mg.csr_parts(tag + "_rawA", csr_matrix(raw_A)))
            out.update(mg.csr_parts(tag + "_A", csr_matrix(A)))
            out[tag + "_membership"] = np.asarray(ca.membership, dtype=np.int64)
            out[tag + "_ncalls"] = np.array([len(CALLS)], dtype=np.int64)
            for i, call in enumerate(CALLS):
                out[f"{tag}_call{i}_edges"] = call["edges"].astype(np.int32)
                out[f"{tag}_call{i}_w"] = call["w"]
                out[f"{tag}_call{i}_membership"] = call["membership"].astype(np.int64)
            print(tag, "clustering calls:", len(CALLS), "groups:", len(K), "nnz(A):", csr_matrix(A).nnz)
    # ---- what repeated calls on the SAME A do (the resolution auto-search, hyperpartition.py:1049-1104, passes one A to
    # every IRMM call): the reference re-weights the caller's matrix IN PLACE (:1157-1172), so with allelic_factor = -1 the
    # allelic entries change sign on every call and weak entries stay zero.  Three consecutive calls, one matrix.
    A_shared = ah.HyperGraph.clique_expansion_init(H.copy(), min_weight=1.0)
    out.update(mg.csr_parts("inplace_A0", csr_matrix(A_shared).copy()))
    wk = pa[::7][:40]                                                   # a few weak pairs as well
    out["inplace_weak_pairs"] = wk
    P_al = [pd.Series(pa[:, 0]), pd.Series(pa[:, 1])]
    P_wk = [pd.Series(np.concatenate([wk[:, 0], wk[:, 1]])), pd.Series(np.concatenate([wk[:, 1], wk[:, 0]]))]
    for i, res in enumerate((1.0, 1.2, 1.4)):
        CALLS.clear()
        raw_A, A_ret, ca, K = ah.IRMM(None, A_shared, None, P_al, P_wk, -1, 0.3, res, 1.0, 0.01, 1, method="leiden")
        assert len(CALLS) == 1
        out[f"inplace_call{i}_edges"] = CALLS[0]["edges"].astype(np.int32)
        out[f"inplace_call{i}_w"] = CALLS[0]["w"]
        out.update(mg.csr_parts(f"inplace_call{i}_rawA", csr_matrix(raw_A)))
        out.update(mg.csr_parts(f"inplace_call{i}_callerA", csr_matrix(A_shared).copy()))
        print("inplace call", i, "edges", len(CALLS[0]["w"]), "negative weights", int((CALLS[0]["w"] < 0).sum()))
    out["params"] = np.array([1.6, 1.0, 0.01], dtype=np.float64)      # resolution, min_weight, threshold
    path = os.path.join(HERE, "ref_irmm_rounds.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
