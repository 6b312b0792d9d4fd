"""The oracle's restatement of IRMM's iterative re-weighting (oracle/hypergraph_oracle.py::irmm_rounds) against the
REFERENCE's own IRMM run (tests/golden/ref_irmm_rounds.npz, made by tests/golden/make_golden_irmm.py: the reference's
code with igraph's clustering calls answered by oracle/community.c).  CPU only."""
import numpy as np
import pytest
from scipy.sparse import csr_matrix

from helpers import load_golden, csr_from
from oracle import community_oracle as co
from oracle import hypergraph_oracle as orc


def graph_dict(edges, w):
    return {(int(a), int(b)): float(x) for (a, b), x in zip(edges, w)}


def assert_graphs_close(got, want, thr, rtol=2e-5, max_edge_cases=10):
    """same edges and weights.  A_new entries within rounding of the threshold may be kept on one side only: after the
    damped update such an entry differs by thr/2 (or is missing); at most `max_edge_cases` of those."""
    edge_cases = 0
    for k in set(got) | set(want):
        if k in got and k in want:
            dlt = abs(got[k] - want[k])
            if dlt <= rtol * max(abs(want[k]), 1e-12):
                continue
            assert abs(dlt - 0.5 * thr) <= 4 * rtol * thr, (k, got[k], want[k])
        else:
            v = got.get(k, want.get(k))
            assert abs(abs(v) - thr) <= 4 * rtol * thr or abs(abs(v) - 0.5 * thr) <= 4 * rtol * thr, (k, v)
        edge_cases += 1
    assert edge_cases <= max_edge_cases, edge_cases


@pytest.mark.parametrize("name", ["louvain", "leiden"])
@pytest.mark.parametrize("max_round", [1, 4])
def test_irmm_rounds_match_reference(name, max_round):
    g = load_golden("irmm_rounds")
    H = csr_from(g, "H")
    res, mw, thr = g["params"]
    tag = f"{name}_r{max_round}"
    leiden = name == "leiden"
    P = (g["allelic_pairs"][:, 0], g["allelic_pairs"][:, 1]) if leiden else None
    calls = []

    def cl(G, resolution, init):
        U = csr_matrix(np.triu(G.toarray(), 1)).tocoo()
        m, _ = co.community(G, resolution, "leiden" if leiden else "louvain", initial=init if leiden else None)
        calls.append((np.column_stack([U.row, U.col]), U.data.copy(), m.copy()))
        return m
    raw_A, A, mem, hist = orc.irmm_rounds(H, cl, P_allelic=P, allelic_factor=-1.0 if leiden else 0.0, resolution=res,
                                          min_weight=mw, threshold=thr, max_round=max_round)
    assert len(calls) == int(g[tag + "_ncalls"][0])
    for i, (e, w, m) in enumerate(calls):
        assert_graphs_close(graph_dict(e, w), graph_dict(g[f"{tag}_call{i}_edges"], g[f"{tag}_call{i}_w"]), mw)
        assert co.ari(m, g[f"{tag}_call{i}_membership"]) == 1.0
    assert co.ari(mem, g[tag + "_membership"]) == 1.0
    Aref = csr_from(g, tag + "_A").toarray()
    Ag = A.toarray()
    close = np.isclose(Ag, Aref, rtol=2e-5, atol=1e-9)
    # entries of A_new within rounding of the threshold may be kept on one side only (half weight after damping)
    assert (~close).sum() <= 20, (~close).sum()
