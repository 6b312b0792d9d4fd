"""GPU parity of Louvain / Leiden (csrc/community.cu through the C-ABI) against the sequential oracle
(oracle/community.c) -- north star: modularity within 1e-5 relative and ARI >= 0.99 on inputs with a clear optimum;
on noisy inputs (no clear optimum: sequential and phased sweeps end in different local optima, as two igraph seeds
do) the GPU result must not fall more than 1 % of |Q| below the oracle's (measured on B200 over 30 noisy cases with the
adaptive phase count: 28 within -0.2 % .. +3.8 %, two at -1.8 % / -2.1 %, profiles/r2u_community_tiny.txt; with a fixed 8
phases it was -6.8 %).  The two violators (n = 2 000, gamma = 1) are in this file as strict expected failures.
With one vertex per phase the device sweep IS the oracle's sequential sweep and must reproduce its partition."""
import numpy as np
import pytest
from scipy.sparse import csr_matrix

from oracle import community_oracle as co

pytestmark = pytest.mark.gpu


def planted(n, g, p_in, p_out, seed, wmax=5.0):
    rng = np.random.default_rng(seed)
    truth = np.arange(n) % g
    P = np.where(truth[:, None] == truth[None, :], p_in, p_out)
    U = np.triu(rng.random((n, n)) < P, 1)
    W = np.triu(rng.uniform(0.5, wmax, size=(n, n)), 1) * U
    A = csr_matrix(W + W.T)
    return A, truth


def gpu_cluster(A, gamma, method, **kw):
    from cphasing_b200 import community
    vc = community.cluster(A, gamma, method, **kw)
    return np.asarray(vc.membership), vc


@pytest.mark.parametrize("method", ["louvain", "leiden"])
@pytest.mark.parametrize("n,g,p_in,p_out,seed", [(300, 6, 0.4, 0.01, 1), (1200, 12, 0.3, 0.005, 5), (64, 2, 0.9, 0.02, 7),
                                                  (800, 4, 0.9, 0.1, 13)])     # last case: mean degree > 128 -> CTA-per-vertex path
def test_planted_partition_matches_oracle(method, n, g, p_in, p_out, seed):
    A, truth = planted(n, g, p_in, p_out, seed)
    mem_o, g_o = co.community(A, 1.0, method)
    mem, vc = gpu_cluster(A, 1.0, method)
    assert co.ari(mem, truth) >= 0.99
    assert co.ari(mem, mem_o) >= 0.99
    q_o = co.modularity(A, mem_o, 1.0)
    q_g = co.modularity(A, mem, 1.0)
    assert abs(q_g - q_o) <= 1e-5 * abs(q_o)
    # the quality the kernel reports is the oracle's formula on the kernel's own partition (fixed point vs double)
    assert abs(vc.quality - q_g) <= 1e-9 * max(abs(q_g), 1.0)
    assert len(vc) == g_o == g
    # group numbering = order of first appearance, like the oracle's renumber()
    first = [int(np.where(mem == c)[0][0]) for c in range(len(vc))]
    assert first == sorted(first)


@pytest.mark.parametrize("method", ["louvain", "leiden"])
@pytest.mark.parametrize("gamma", [0.5, 1.0, 2.0])
def test_noisy_graph_quality_close_to_oracle(method, gamma):
    A, truth = planted(400, 8, 0.25, 0.03, seed=2)
    mem_o, _ = co.community(A, gamma, method)
    mem, vc = gpu_cluster(A, gamma, method)
    q_o = co.modularity(A, mem_o, gamma)
    q_g = co.modularity(A, mem, gamma)
    assert q_g >= q_o - 0.01 * abs(q_o), (q_g, q_o)


NOISY_4K = {}


def noisy_4k(case):
    if case not in NOISY_4K:
        NOISY_4K[case] = planted(*case)
    return NOISY_4K[case]


@pytest.mark.parametrize("method", ["louvain", "leiden"])
@pytest.mark.parametrize("gamma", [0.5, 1.0, 2.0])
@pytest.mark.parametrize("case", [(4000, 48, 0.3, 0.02, 4), (4000, 48, 0.2, 0.02, 5)])
def test_noisy_4000_vertices_within_one_percent(method, gamma, case):
    """the n = 4 000 graphs of profiles/r1c_community_phases.txt (in-group degree 25, out-group degree 78: Q ~ 0.2) and a
    noisier one (Q ~ 0.15), the size of workload C's contig graph"""
    A, truth = noisy_4k(case)
    mem_o, _ = co.community(A, gamma, method)
    mem, vc = gpu_cluster(A, gamma, method)
    q_o = co.modularity(A, mem_o, gamma)
    q_g = co.modularity(A, mem, gamma)
    assert q_g >= q_o - 0.01 * abs(q_o), (q_g, q_o)


KNOWN_GAP = {("louvain", 1.0): "-1.79 %", ("leiden", 1.0): "-2.08 %"}     # profiles/r2u_community_tiny.txt:8, 11


@pytest.mark.parametrize("method,gamma", [
    pytest.param(m, g_, marks=pytest.mark.xfail(strict=True, reason=f"known violator of the 1 % bound: {KNOWN_GAP[(m, g_)]} "
                                                                    "of the oracle's Q on B200 (phased sweeps end in a "
                                                                    "worse local optimum on this graph); kept in the gate "
                                                                    "so that a fix flips it"))
    if (m, g_) in KNOWN_GAP else pytest.param(m, g_)
    for m in ("louvain", "leiden") for g_ in (0.5, 1.0, 2.0)])
def test_noisy_2000_vertices_within_one_percent(method, gamma):
    """the n = 2 000 graph of scripts/exp_community2.py (20 groups, in-group degree 10, out-group degree 19: Q ~ 0.28 at
    gamma = 1).  Four of its six cases hold the 1 % bound; the two at gamma = 1 do NOT (measured -1.8 % / -2.1 %) and are
    listed as strict expected failures rather than left out: the tolerance of this file is not met everywhere."""
    A, truth = noisy_4k((2000, 20, 0.1, 0.01, 3))
    mem_o, _ = co.community(A, gamma, method)
    mem, vc = gpu_cluster(A, gamma, method)
    q_o = co.modularity(A, mem_o, gamma)
    q_g = co.modularity(A, mem, gamma)
    assert q_g >= q_o - 0.01 * abs(q_o), (q_g, q_o)


@pytest.mark.parametrize("case", [(400, 8, 0.25, 0.03, 2), (1500, 20, 0.2, 0.02, 9)])
def test_one_vertex_per_phase_is_the_sequential_sweep(case, monkeypatch):
    """n_phases >= n: every phase holds one vertex, so the device Louvain visits the vertices like the oracle does (index
    order, best gain, ties to the lowest community id) and must end in the SAME partition, noisy graph or not"""
    monkeypatch.setenv("CPH_COMMUNITY_PHASES", "1000000")
    A, truth = planted(*case)
    mem_o, _ = co.community(A, 1.0, "louvain")
    mem, vc = gpu_cluster(A, 1.0, "louvain")
    assert co.ari(mem, mem_o) == 1.0
    assert abs(co.modularity(A, mem, 1.0) - co.modularity(A, mem_o, 1.0)) <= 1e-12


def test_deterministic_and_float32_input():
    from cphasing_b200 import community
    A, truth = planted(500, 5, 0.2, 0.02, seed=11)
    m1, _ = gpu_cluster(A, 1.0, "leiden")
    m2, _ = gpu_cluster(A, 1.0, "leiden")
    assert np.array_equal(m1, m2)
    m3, _ = gpu_cluster(A.astype(np.float32), 1.0, "leiden")
    assert co.ari(m1, m3) >= 0.99


def test_leiden_communities_connected():
    from scipy.sparse.csgraph import connected_components
    A, _ = planted(500, 10, 0.05, 0.004, seed=3)
    mem, vc = gpu_cluster(A, 1.0, "leiden")
    for c in range(len(vc)):
        ix = np.where(mem == c)[0]
        ncomp, _ = connected_components(A[ix][:, ix], directed=False)
        assert ncomp == 1


def test_negative_weights_and_initial_membership():
    """allelic_factor < 0 makes allelic pairs repel (algorithms/hypergraph.py:1158-1170, 1247)."""
    A, truth = planted(240, 4, 0.5, 0.05, seed=4)
    A = A.tolil()
    rng = np.random.default_rng(0)
    for _ in range(200):                       # negative edges between different planted groups
        i, j = rng.integers(0, 240, 2)
        if truth[i] != truth[j]:
            A[i, j] = A[j, i] = -3.0
    A = A.tocsr()
    mem_o, _ = co.community(A, 1.0, "leiden")
    mem, vc = gpu_cluster(A, 1.0, "leiden")
    assert co.ari(mem, truth) >= 0.99 and co.ari(mem, mem_o) >= 0.99
    m3, _ = gpu_cluster(A, 1.0, "leiden", initial_membership=truth)
    assert co.ari(m3, truth) == 1.0
    m4, _ = gpu_cluster(A, 1.0, "louvain", initial_membership=truth)
    assert co.ari(m4, truth) == 1.0


def test_resolution_monotone_and_modularity_entry():
    from cphasing_b200 import community, graphops
    A, truth = planted(240, 4, 0.5, 0.05, seed=4)
    m1, v1 = gpu_cluster(A, 0.3, "louvain")
    m2, v2 = gpu_cluster(A, 3.0, "louvain")
    assert len(v1) <= len(v2)
    a = graphops.upload_adjacency(A)
    for gamma in (0.3, 1.0, 3.0):
        q = community.modularity_device(a, truth, gamma)
        assert abs(q - co.modularity(A, truth, gamma)) < 1e-9


def test_empty_and_isolated_vertices():
    A = csr_matrix((50, 50), dtype=np.float64)
    mem, vc = gpu_cluster(A, 1.0, "louvain")
    assert np.array_equal(mem, np.arange(50))
    B, truth = planted(100, 2, 0.6, 0.0, seed=9)
    B = B.tolil()
    B[10, :] = 0
    B[:, 10] = 0
    B = B.tocsr()
    mem, vc = gpu_cluster(B, 1.0, "leiden")
    mem_o, g_o = co.community(B, 1.0, "leiden")
    assert co.ari(mem, mem_o) >= 0.99 and len(vc) == g_o
