"""The clustering oracle (oracle/community.c) against networkx's independent modularity and Louvain
implementations, and against planted partitions."""
import numpy as np
import pytest
from scipy.sparse import csr_matrix

from oracle import community_oracle as co


def planted(n, g, p_in, p_out, seed, wmax=5.0):
    rng = np.random.default_rng(seed)
    truth = np.arange(n) % g
    P = np.where(truth[:, None] == truth[None, :], p_in, p_out)
    U = np.triu(rng.random((n, n)) < P, 1)
    W = np.triu(rng.uniform(0.5, wmax, size=(n, n)), 1) * U
    A = csr_matrix(W + W.T)
    return A, truth


@pytest.mark.parametrize("method", ["louvain", "leiden"])
def test_planted_partition_recovered(method):
    A, truth = planted(300, 6, 0.4, 0.01, seed=1)
    mem, g = co.community(A, 1.0, method)
    assert g == 6 and co.ari(mem, truth) == 1.0
    import networkx as nx
    G = nx.from_scipy_sparse_array(A)
    q_nx = nx.community.modularity(G, [set(np.where(mem == c)[0]) for c in range(g)], weight="weight", resolution=1.0)
    assert abs(co.modularity(A, mem, 1.0) - q_nx) < 1e-12


@pytest.mark.parametrize("gamma", [0.5, 1.0, 2.0])
def test_quality_matches_networkx_louvain(gamma):
    import networkx as nx
    A, truth = planted(400, 8, 0.25, 0.03, seed=2)
    G = nx.from_scipy_sparse_array(A)
    best = max(nx.community.modularity(G, nx.community.louvain_communities(G, weight="weight", resolution=gamma, seed=s),
                                       weight="weight", resolution=gamma) for s in range(3))
    for method in ("louvain", "leiden"):
        mem, g = co.community(A, gamma, method)
        q = co.modularity(A, mem, gamma)
        assert q >= best - 0.05 * abs(best), (method, q, best)


def test_leiden_communities_connected():
    """Leiden's guarantee: every community induces a connected subgraph."""
    from scipy.sparse.csgraph import connected_components
    A, _ = planted(500, 10, 0.05, 0.004, seed=3)
    mem, g = co.community(A, 1.0, "leiden")
    for c in range(g):
        ix = np.where(mem == c)[0]
        ncomp, _ = connected_components(A[ix][:, ix], directed=False)
        assert ncomp == 1


def test_initial_membership_and_resolution_monotone():
    A, truth = planted(240, 4, 0.5, 0.05, seed=4)
    m1, g1 = co.community(A, 0.3, "louvain")
    m2, g2 = co.community(A, 3.0, "louvain")
    assert g1 <= g2
    m3, g3 = co.community(A, 1.0, "leiden", initial=truth)
    assert co.ari(m3, truth) == 1.0
