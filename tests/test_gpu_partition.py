"""GPU parity of the stages downstream of A: kprune (kprune.py:385-552), the adjacency filters
(algorithms/hypergraph.py:526-580), IRMM's graph preparation + clustering (:1157-1256), end to end on a
workload-A shaped synthetic table (5 chromosomes x 4 haplotypes)."""
import numpy as np
import pytest
from scipy.sparse import csr_matrix

from oracle import community_oracle as co
from oracle import hypergraph_oracle as orc

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def case():
    """table -> H -> A through the device classes, plus the oracle's H and A of the same table."""
    from cphasing_b200 import synth
    from cphasing_b200 import hypergraph as hg
    from cphasing_b200.algorithms import hypergraph as ah
    g = synth.make_genome(400, 5, 4, seed=20240)
    t = synth.porec_table(g, 120_000, seed=20240)
    ex = hg.HyperExtractor.from_arrays(t["read_idx"], t["contig_idx"], t["mapq"], g.contig_idx, g.contigsizes,
                                       min_quality=1, edge_length=0)
    HG = ah.HyperGraph(ex.edges, min_quality=1)
    H = HG.incidence_matrix(min_contacts=25)
    r, c, cnt, q = orc.porec_edges(t["read_idx"], t["contig_idx"], t["mapq"], min_mapq=1, n_contigs=g.n)
    rr, cc, kk, ncol = orc.hypergraph_get_data(r, c, cnt, q, 1)
    Ho, keep_rows, _ = orc.incidence_matrix(rr, cc, kk, g.n, ncol, 25)
    return dict(g=g, ah=ah, HG=HG, H=H, Ho=Ho, nodes=HG.nodes, alleles=synth.allele_pairs(g, seed=3))


def _kept_index(case):
    """allele pairs re-indexed to the rows that survived min_contacts."""
    name2new = {nm: i for i, nm in enumerate(case["nodes"])}
    names = sorted(case["g"].contig_idx, key=case["g"].contig_idx.get)
    out = []
    for a, b, *_ in case["alleles"]:
        na, nb = names[a], names[b]
        if na in name2new and nb in name2new:
            out.append((name2new[na], name2new[nb]))
    return out


def test_kprune_matches_oracle(case):
    from cphasing_b200 import kprune
    ah = case["ah"]
    A = ah.HyperGraph.clique_expansion_init(case["H"], min_weight=1.0, use_high_order=False)
    Ao = orc.clique_expansion(case["Ho"], use_high_order=False, min_weight=1.0)
    pairs = _kept_index(case)
    al, wk = kprune.kprune_pairs(A, pairs)
    al_o, wk_o = orc.kprune_hypergraph(Ao, pairs)
    assert al == al_o
    assert wk == wk_o
    assert len(al) > 50 and len(wk) > 50           # the case exercises both lists


@pytest.mark.parametrize("max_alleles", [2, 3])
def test_kprune_general_blocks_match_oracle(case, max_alleles):
    """max_alleles > 1: the l1 x l2 score block over [contig, *alleles] of KPruner (kprune.py:269-335) with an exact
    maximum-weight assignment, against the oracle's brute force over permutations."""
    from cphasing_b200 import kprune
    ah = case["ah"]
    A = ah.HyperGraph.clique_expansion_init(case["H"], min_weight=1.0, use_high_order=False)
    Ao = orc.clique_expansion(case["Ho"], use_high_order=False, min_weight=1.0).tocsr().astype(np.float64)
    g = case["g"]
    # every contig lists the contigs of the other haplotypes of its chromosome region: up to ploidy - 1 alleles each
    from cphasing_b200 import synth
    pairs = []
    name2new = {nm: i for i, nm in enumerate(case["nodes"])}
    names = sorted(g.contig_idx, key=g.contig_idx.get)
    for a, b, *_ in synth.allele_pairs(g, seed=77, per_contig=3):
        if names[a] in name2new and names[b] in name2new:
            pairs.append((name2new[names[a]], name2new[names[b]]))
    al, wk = kprune.kprune_pairs(A, pairs, max_alleles=max_alleles)
    # oracle: allelic pairs (stored upper entries that are allele pairs) are zeroed in the upper triangle, then every other
    # stored upper entry is scored on its block
    from scipy.sparse import triu
    Az = Ao.copy(); Az.setdiag(0); Az.eliminate_zeros()
    U = triu(Az).tocoo()
    stored = set(zip(U.row.tolist(), U.col.tolist()))
    al_o = sorted(set((a, b) for a, b in pairs if a < b) & stored)
    assert al == al_o
    D = Az.todok()
    for a, b in al_o:
        D[a, b] = 0.0
    lists = {}
    for a, b in pairs:
        if len(lists.setdefault(a, [])) < max_alleles:
            lists[a].append(b)
    score = lambda i, j: 0.0 if i == j else float(D.get((i, j), 0.0))
    wk_o = orc.kprune_general(score, lists, sorted(stored - set(al_o)))
    assert wk == sorted(wk_o)
    assert len(wk) > 50
    assert max(len(v) for v in lists.values()) == max_alleles


def test_kprune_hypergraph_class_and_table(case):
    import pandas as pd
    from cphasing_b200 import kprune
    ah = case["ah"]
    A = ah.HyperGraph.clique_expansion_init(case["H"], min_weight=1.0, use_high_order=False)
    names = list(case["nodes"])
    c2i = {nm: i for i, nm in enumerate(names)}
    allg = sorted(case["g"].contig_idx, key=case["g"].contig_idx.get)
    rows = [(0, 0, allg[a], allg[b], m1, m2, sh, sim, 1) for a, b, m1, m2, sh, sim in case["alleles"]]
    df = pd.DataFrame(rows, columns=[-1, 0, 1, 2, 3, 4, 5, 6, 7]).set_index(-1)

    class AT:      # the slice of core.AlleleTable the class reads (core.py:289-300)
        data = df
    kp = kprune.KPruneHyperGraph(AT(), A, c2i)
    kp.run()
    al_o, wk_o = orc.kprune_hypergraph(orc.clique_expansion(case["Ho"], use_high_order=False, min_weight=1.0), _kept_index(case))
    assert kp.allelic_prune_list == al_o and kp.weak_prune_list == wk_o
    pt = kp.get_prune_table()
    assert list(pt.columns) == ['contig1', 'contig2', 'mz1', 'mz2', 'mzShared', 'similarity', 'type']
    assert (pt["type"] == 0).sum() == len(al_o) and (pt["type"] == 1).sum() == len(wk_o)
    assert len(kp.get_prune_table(symmetric=True)) == 2 * len(pt)


def test_filter_adjacency_matrix_matches_oracle(case):
    ah = case["ah"]
    A = ah.HyperGraph.clique_expansion_init(case["H"], min_weight=2.0)
    Ao = orc.clique_expansion(case["Ho"], min_weight=2.0)
    sizes = case["g"].contigsizes
    keep = ah.HyperGraph.filter_adjacency_matrix(A, case["nodes"], sizes)
    lengths = np.array([sizes[nm] for nm in case["nodes"]], dtype=np.float64)
    keep_o = orc.filter_adjacency(Ao, lengths)
    assert np.array_equal(keep, keep_o)
    assert len(keep) < A.shape[0]
    import pandas as pd
    keep2 = ah.HyperGraph.filter_adjacency_matrix(A, case["nodes"], pd.DataFrame({"length": pd.Series(sizes)}))
    assert np.array_equal(keep2, keep_o)


@pytest.mark.parametrize("method,allelic_factor", [("louvain", 0), ("leiden", 0), ("louvain", -1)])
def test_irmm_matches_oracle(case, method, allelic_factor):
    """IRMM(H, A, ...) on the device vs oracle: graph preparation (irmm_graph) + sequential clustering.  The synthetic
    genome has a clear optimum per resolution (20 haplotype groups), so the north-star bar applies."""
    from cphasing_b200 import kprune
    ah = case["ah"]
    A = ah.HyperGraph.clique_expansion_init(case["H"], min_weight=2.0)
    Ao = orc.clique_expansion(case["Ho"], min_weight=2.0)
    pairs = _kept_index(case)
    al, wk = orc.kprune_hypergraph(orc.clique_expansion(case["Ho"], use_high_order=False, min_weight=1.0), pairs)
    sym = lambda L: [np.array([a for a, b in L] + [b for a, b in L]), np.array([b for a, b in L] + [a for a, b in L])]
    Pa, Pw = sym(al), sym(wk)
    raw_A, A2, ca, K = ah.IRMM(case["H"], A, None, Pa, Pw, allelic_factor=allelic_factor, resolution=1.0, method=method)
    src, dst, w = orc.irmm_graph(Ao, Pa, Pw, allelic_factor)
    n = Ao.shape[0]
    G = csr_matrix((np.concatenate([w, w]), (np.concatenate([src, dst]), np.concatenate([dst, src]))), shape=(n, n))
    use_leiden = method == "leiden" or allelic_factor < 0
    mem_o, g_o = co.community(G, 1.0, "leiden" if use_leiden else "louvain")
    mem = np.asarray(ca.membership)
    q_o = co.modularity(G, mem_o, 1.0)
    q_g = co.modularity(G, mem, 1.0)
    assert abs(q_g - q_o) <= 1e-5 * abs(q_o), (q_g, q_o)
    assert co.ari(mem, mem_o) >= 0.99
    assert [set(c) for c in ca] == K and sum(len(k) for k in K) == n
    # the returned A: re-weighted entries and a zero diagonal, raw_A untouched
    A2 = csr_matrix(A2)
    assert abs(A2.diagonal()).max() == 0
    assert abs(csr_matrix(raw_A) - csr_matrix(A)).nnz == 0
    if al:
        i, j = al[0]
        assert A2[i, j] == pytest.approx(allelic_factor * csr_matrix(A)[i, j])
    if wk:
        i, j = wk[0]
        assert A2[i, j] == 0 and A2[j, i] == 0
    # truth: every haplotype group of the synthetic genome is one cluster at resolution 1?  Not required -- but the
    # clusters must not mix chromosomes
    grp = case["g"].group
    names = sorted(case["g"].contig_idx, key=case["g"].contig_idx.get)
    name2old = {nm: i for i, nm in enumerate(names)}
    chrom = np.array([grp[name2old[nm]] // case["g"].ploidy for nm in case["nodes"]])
    for k in K:
        if len(k) > 5:
            vals, cnts = np.unique(chrom[list(k)], return_counts=True)
            assert cnts.max() >= 0.9 * len(k)


def test_irmm_builds_A_when_missing(case):
    ah = case["ah"]
    raw_A, A2, ca, K = ah.IRMM(case["H"], None, None, None, None, allelic_factor=0, resolution=1.0, min_weight=2.0)
    Ao = orc.clique_expansion(case["Ho"], min_weight=2.0)
    src, dst, w = orc.irmm_graph(Ao)
    n = Ao.shape[0]
    G = csr_matrix((np.concatenate([w, w]), (np.concatenate([src, dst]), np.concatenate([dst, src]))), shape=(n, n))
    mem_o, _ = co.community(G, 1.0, "louvain")
    assert co.ari(np.asarray(ca.membership), mem_o) >= 0.99


# ---- partition driver post-processing (hyperpartition.py:2288-2471, 2860-2968) against the reference's own outputs
def _groups(labels):
    labels = np.asarray(labels)
    return [sorted(np.where(labels == g)[0].tolist()) for g in range(int(labels.max()) + 1)]


def _part(K):
    return sorted(tuple(sorted(int(x) for x in g)) for g in K)


def test_group_sums_match_scipy():
    from helpers import csr_from, load_golden
    from cphasing_b200 import hyperpartition as hp
    g = load_golden("partition_post")
    A = csr_from(g, "A").astype(np.float64)
    K = _groups(g["K3_labels"])                      # some contigs are in no group (-1)
    S, V = hp.group_sums(A, K, want_S=True, want_V=True)
    n = A.shape[0]
    rows = [c for x in K for c in x]
    cols = [i for i, x in enumerate(K) for _ in x]
    M = csr_matrix((np.ones(len(rows)), (rows, cols)), shape=(n, len(K)))
    np.testing.assert_allclose(S, A.dot(M).toarray(), rtol=1e-9, atol=1e-9)
    np.testing.assert_allclose(V, M.T.dot(A).dot(M).toarray(), rtol=1e-9, atol=1e-9)


def test_merge_and_recluster_match_reference_golden():
    from helpers import csr_from, load_golden
    from cphasing_b200 import hyperpartition as hp
    import pandas as pd
    import types
    g = load_golden("partition_post")
    A = csr_from(g, "A")
    n = A.shape[0]
    vsz = pd.DataFrame({"length": {i: int(g["lengths"][i]) for i in range(n)}})
    K0 = _groups(g["K0_labels"])
    G = len(set(g["group"].tolist()))
    for method in ("mean", "sum"):
        K1 = hp.HyperPartition._merge(A, [list(x) for x in K0], vsz, k=G, method=method)
        assert _part(K1) == _part(_groups(g[f"merge_{method}_labels"]))
        assert [x[0] for x in K1] == g[f"merge_{method}_order"].tolist()        # same merge order too
    aset = set(map(tuple, g["allelic_pairs"].tolist()))
    K2 = hp.HyperPartition._merge(A, [list(x) for x in K0], vsz, k=G, allelic_idx_set=aset, min_allelic_overlap=0.1)
    assert _part(K2) == _part(_groups(g["merge_conflict_labels"]))
    Ar = csr_from(g, "Ar")
    fake = types.SimpleNamespace(K=_groups(g["K3_labels"]), allelic_idx_set=set(), min_weight=1.0, H=None,
                                 idx_to_vertices={i: f"c{i}" for i in range(n)},
                                 contig_sizes={f"c{i}": int(g["lengths"][i]) for i in range(n)})
    moved = hp.HyperPartition.recluster_by_linkage(fake, A=Ar, min_ratio=1.5, iter_count=1)
    assert moved == int(g["recluster_count"][0])
    assert _part(fake.K) == _part(_groups(g["recluster_labels"]))


@pytest.mark.parametrize("normalize", [False, True])
def test_single_partition_end_to_end(case, tmp_path, normalize):
    """table -> HyperPartition(...).single_partition(k) with the resolution auto-search: the groups are the planted
    haplotype groups of the synthetic genome, and the oracle pipeline (H, A, IRMM graph, sequential Louvain at the
    same resolutions, filter, merge) ends in the same partition."""
    from cphasing_b200 import hyperpartition as hp, synth
    from cphasing_b200 import hypergraph as hg
    g = case["g"]
    t = synth.porec_table(g, 120_000, seed=20240)
    ex = hg.HyperExtractor.from_arrays(t["read_idx"], t["contig_idx"], t["mapq"], g.contig_idx, g.contigsizes,
                                       min_quality=1, edge_length=0)
    P = hp.HyperPartition(ex.edges, g.contigsizes, k=[20], min_contacts=25, min_weight=2.0, min_cis_weight=20,
                          resolution1=-1, init_resolution1=1.0, min_scaffold_length=2_000_000, allelic_factor=0,
                          normalize=normalize, min_length=10000)
    K = P.single_partition(k=20)
    assert len(K) == 20
    names = list(P.vertices)
    old = {nm: i for i, nm in enumerate(sorted(g.contig_idx, key=g.contig_idx.get))}
    truth = np.array([g.group[old[nm]] for nm in names])
    mem = np.full(len(names), -1)
    for i, grp in enumerate(K):
        mem[grp] = i
    inside = mem >= 0
    assert inside.mean() > 0.95
    assert co.ari(mem[inside], truth[inside]) >= 0.99
    # sorted by total length, descending (:1210-1213)
    tot = [sum(P.vertices_idx_sizes[i] for i in grp) for grp in K]
    assert tot == sorted(tot, reverse=True)
    out = tmp_path / "out.clusters.txt"
    P.to_cluster(str(out))
    lines = out.read_text().strip().split("\n")
    assert len(lines) == 20 and lines[0].startswith("Chr01\t")
    f = lines[3].split("\t")
    assert int(f[1]) == len(f[2].split(" ")) == len(K[3])

    # oracle pipeline at the resolutions the search visited
    Ho = case["Ho"]
    Ao = orc.clique_expansion(Ho, min_weight=2.0)
    keep = np.where(Ao.diagonal() >= 20)[0]
    if len(keep) < Ao.shape[0]:
        Ho, _, _ = orc.extract_incidence_matrix2(Ho, keep)
        Ao = orc.clique_expansion(Ho, min_weight=2.0)
    assert Ao.shape[0] == len(names)
    if normalize:
        lens = np.array([g.contigsizes[nm] for nm in names], dtype=np.float64)
        Ao = orc.normalize_adjacency(Ao, orc.normalize_weight_svc(Ho, lens))
    src, dst, w = orc.irmm_graph(Ao)
    n = Ao.shape[0]
    Gm = csr_matrix((np.concatenate([w, w]), (np.concatenate([src, dst]), np.concatenate([dst, src]))), shape=(n, n))
    res, glen = P.resolutions_tried[-1]
    mem_o, _ = co.community(Gm, res, "louvain")
    sizes = P.vertices_idx_sizes
    Ko = [np.where(mem_o == c)[0].tolist() for c in range(mem_o.max() + 1)]
    Ko = [x for x in Ko if sum(sizes[i] for i in x) >= 2_000_000]
    assert len(Ko) == glen
    if len(Ko) > 20:
        Ko = orc.merge_groups(Ao, Ko, {i: float(s) for i, s in sizes.items()}, 20)
    mo = np.full(n, -1)
    for i, grp in enumerate(Ko):
        mo[grp] = i
    both = (mo >= 0) & (mem >= 0)
    assert co.ari(mem[both], mo[both]) >= 0.99


@pytest.mark.parametrize("use", ["alleletable", "prunetable", "none"])
def test_incremental_partition_two_rounds(case, use):
    """hyperpartition.py:1726-2286 + 1239-1560: round 1 -> 5 chromosome groups (auto-search down from the initial
    resolution, merge), round 2 -> 4 haplotype groups inside each (sub-H, cis filter, per-group kprune, IRMM with the
    resolution search, length filter, merge).  The result must be the planted 5 x 4 structure."""
    import pandas as pd
    from cphasing_b200 import hyperpartition as hp, synth, kprune
    from cphasing_b200 import hypergraph as hg
    g = case["g"]
    t = synth.porec_table(g, 120_000, seed=20240)
    ex = hg.HyperExtractor.from_arrays(t["read_idx"], t["contig_idx"], t["mapq"], g.contig_idx, g.contigsizes,
                                       min_quality=1, edge_length=0)
    allg = sorted(g.contig_idx, key=g.contig_idx.get)
    rows = [(0, 0, allg[a], allg[b], m1, m2, sh, sim, 1) for a, b, m1, m2, sh, sim in case["alleles"]]
    at = pd.DataFrame(rows, columns=[-1, 0, 1, 2, 3, 4, 5, 6, 7]).set_index(-1)
    kw = dict(min_contacts=25, min_weight=2.0, min_cis_weight=20, resolution1=-1, resolution2=-1, init_resolution1=0.8,
              init_resolution2=0.8, min_scaffold_length=2_000_000, allelic_factor=0, min_length=10000)
    if use == "alleletable":
        P = hp.HyperPartition(ex.edges, g.contigsizes, k=[5, 4], alleletable=at, **kw)
    elif use == "prunetable":
        P0 = hp.HyperPartition(ex.edges, g.contigsizes, k=[5, 4], alleletable=at, **kw)
        P0.kprune(at)
        table = P0.prunetable
        assert (table["type"] == 0).sum() > 50
        P = hp.HyperPartition(ex.edges, g.contigsizes, k=[5, 4], prunetable=table, **kw)
    else:
        P = hp.HyperPartition(ex.edges, g.contigsizes, k=[5, 4], **kw)
    K = P.incremental_partition([5, 4])
    names = list(P.vertices)
    old = {nm: i for i, nm in enumerate(allg)}
    truth = np.array([g.group[old[nm]] for nm in names])
    # round 1: chromosomes
    assert len(P.first_K) == 5
    m1 = np.full(len(names), -1)
    for i, grp in enumerate(P.first_K):
        m1[grp] = i
    in1 = m1 >= 0
    assert in1.mean() > 0.95 and co.ari(m1[in1], truth[in1] // g.ploidy) >= 0.99
    # round 2: haplotypes
    assert len(K) == 20
    assert P.second_group_info[:5] == ["1g1", "1g2", "1g3", "1g4", "2g1"]
    mem = np.full(len(names), -1)
    for i, grp in enumerate(K):
        assert mem[grp].max() == -1, "a contig landed in two groups"
        mem[grp] = i
    inside = mem >= 0
    assert inside.mean() > 0.95
    assert co.ari(mem[inside], truth[inside]) >= 0.99
    # a supplied first cluster (hyperpartition.py:2054-2063) gives the same second round
    first = [[names[i] for i in grp] for grp in P.first_K]
    P2 = hp.HyperPartition(ex.edges, g.contigsizes, k=[5, 4], alleletable=at if use == "alleletable" else None, **kw)
    K2 = P2.incremental_partition([5, 4], first_cluster=first)
    if use == "alleletable":
        assert sorted(map(sorted, K2)) == sorted(map(sorted, K))


def test_irmm_reweight_round_matches_oracle(case):
    """One round of IRMM's iterative re-weighting (algorithms/hypergraph.py:1270-1329) with the SAME membership on both
    sides: per-hyperedge weights W, the suppressed monomers of H_reweighted, and the damped, asymmetric A."""
    import torch
    from cphasing_b200 import community, graphops, irmm
    ah = case["ah"]
    H, Ho = case["H"], case["Ho"].astype(np.float32)
    A = ah.HyperGraph.clique_expansion_init(H, min_weight=2.0)
    a = graphops.clone_adjacency(graphops.upload_adjacency(A))
    graphops.zero_diagonal(a)
    mem, _, st = community.community_device(a, 1.0, "louvain")
    membership = mem.cpu().numpy()
    c = int(membership.max()) + 1
    assert 2 <= c < 100
    R = irmm.Rounds(ah.device_handle(H), a, min_weight=2.0)
    A0 = csr_matrix((a.data64.cpu().numpy(), a.indices.cpu().numpy(), a.indptr.cpu().numpy()), shape=A.shape).toarray()
    dW = R.reweight(membership, c)
    # oracle
    with np.errstate(over="ignore"):
        d = 1.0 / np.maximum(Ho.getnnz(axis=0).astype(np.int8) - np.int8(1), 1)
    H_T = Ho.T.tocsr()
    Hn, HTn, Wn, alias = orc.irmm_reweight_round(Ho, H_T, np.ones(Ho.shape[1], np.float32), membership, c, True)
    assert not alias and HTn.nnz == Hn.nnz          # first suppression round: the zeros reach H_T as well
    Wd = R.edge_weights().cpu().numpy()
    assert Wd.shape == Wn.shape
    np.testing.assert_allclose(Wd, Wn, rtol=2e-6)
    assert R.masked_entries() == Ho.nnz - Hn.nnz and R.masked_entries() > 0
    assert abs(dW - np.abs(Wn - 1.0).max()) < 1e-5
    R.blend(0.5)
    from scipy.sparse import diags
    A_new = ((Hn @ diags(Wn)).astype(np.float32) @ diags(d) @ HTn).toarray()
    edge = np.abs(A_new - 2.0) < 2e-5        # within float32 rounding of W of the threshold: may fall on either side
    assert edge.sum() < 1000                 # of 160 000
    A_new[A_new < 2.0] = 0
    want = 0.5 * A0 + 0.5 * A_new
    np.fill_diagonal(want, 0)
    got = R.dense.cpu().numpy()
    assert not R.alias and int(R.mask_t.sum().item()) == R.masked_entries()
    np.testing.assert_allclose(got[~edge], want[~edge], rtol=1e-6, atol=1e-9)
    # CSR views: mode 0 = A, mode 1 = mirrored upper triangle (what igraph clusters)
    g0, g1 = R.to_csr(0), R.to_csr(1)
    M0 = csr_matrix((g0.data64.cpu().numpy(), g0.indices.cpu().numpy(), g0.indptr.cpu().numpy()), shape=A.shape).toarray()
    M1 = csr_matrix((g1.data64.cpu().numpy(), g1.indices.cpu().numpy(), g1.indptr.cpu().numpy()), shape=A.shape).toarray()
    assert np.array_equal(M0, got)
    U = np.triu(got, 1)
    assert np.array_equal(M1, U + U.T)


@pytest.mark.parametrize("method,allelic_factor", [("louvain", 0), ("leiden", -1)])
def test_irmm_multi_round_end_to_end(case, method, allelic_factor):
    """IRMM(max_round=4) on the planted 5 x 4 genome: device loop against the oracle's restatement of the same loop
    driven by the sequential C clustering; both must land on the planted haplotype groups."""
    ah, g = case["ah"], case["g"]
    res = 2.0
    raw_A, A, ca, K = ah.IRMM(case["H"], None, None, None, None, allelic_factor, 0.3, res, 2.0, 0.01, 4, method=method)
    hist = A._irmm_history
    assert 1 <= len(hist) <= 3 and hist[0]["masked"] > 0
    mem = np.asarray(ca.membership)
    assert len(mem) == case["H"].shape[0] and sorted(map(sorted, K)) == sorted(sorted(c) for c in ca)
    names = list(case["nodes"])
    old = {nm: i for i, nm in enumerate(sorted(g.contig_idx, key=g.contig_idx.get))}
    truth = np.array([g.group[old[nm]] for nm in names])

    def cl(G, resolution, init):
        m, _ = co.community(G, resolution, "leiden" if (allelic_factor < 0 or method == "leiden") else "louvain",
                            initial=init if (allelic_factor < 0 or method == "leiden") else None)
        return m
    _, Ao, mo, ho = orc.irmm_rounds(case["Ho"], cl, allelic_factor=allelic_factor, resolution=res, min_weight=2.0,
                                    threshold=0.01, max_round=4)
    assert co.ari(mem, truth) >= 0.99 and co.ari(mo, truth) >= 0.99
    assert co.ari(mem, mo) >= 0.99
    # the first round's weights do not depend on later clustering differences: same number of suppressed monomers
    assert hist[0]["masked"] == case["Ho"].nnz - ho[0][1]


@pytest.mark.parametrize("name", ["louvain", "leiden"])
@pytest.mark.parametrize("max_round", [1, 4])
def test_irmm_against_reference_run(name, max_round):
    """Device IRMM against the REFERENCE's own IRMM run (tests/golden/ref_irmm_rounds.npz: the reference's code, igraph's
    clustering answered by oracle/community.c): same partition after every round, same returned A."""
    import cphasing_b200.algorithms.hypergraph as ah
    from helpers import load_golden, csr_from
    g = load_golden("irmm_rounds")
    H = csr_from(g, "H")
    res, mw, thr = g["params"]
    tag = f"{name}_r{max_round}"
    leiden = name == "leiden"
    P = [g["allelic_pairs"][:, 0], g["allelic_pairs"][:, 1]] if leiden else None
    raw_A, A, ca, K = ah.IRMM(H, None, None, P, None, -1 if leiden else 0, 0.3, float(res), float(mw), float(thr),
                              max_round, method=name)
    mem = np.asarray(ca.membership)
    assert co.ari(mem, g[tag + "_membership"]) >= 0.99
    if max_round > 1:
        hist = A._irmm_history
        assert len(hist) == int(g[tag + "_ncalls"][0]) - 1
    if co.ari(mem, g[tag + "_membership"]) == 1.0:
        Aref = csr_from(g, tag + "_A").toarray()
        Ag = csr_matrix(A).toarray()
        np.fill_diagonal(Ag, 0); np.fill_diagonal(Aref, 0)
        close = np.isclose(Ag, Aref, rtol=2e-5, atol=1e-9)
        assert (~close).sum() <= 40, (~close).sum()          # threshold ties of A_new (0.5 / 0.25 steps after damping)


@pytest.mark.parametrize("inplace", [True, False])
def test_irmm_repeated_calls_on_one_matrix(inplace, monkeypatch):
    """The resolution search hands ONE matrix to every IRMM call (hyperpartition.py:1049-1104) and the reference re-weights it
    in place (:1157-1172): at allelic_factor = -1 the allelic entries change sign on every call.  Pinned by three
    consecutive calls of the reference's IRMM (ref_irmm_rounds.npz `inplace_*`).  REFERENCE_INPLACE_A = True reproduces the
    graphs of all three calls and what the caller's matrix holds afterwards; the default leaves the caller's matrix alone,
    so every call sees the graph of the reference's FIRST call (deliberate deviations)."""
    import cphasing_b200.algorithms.hypergraph as ah
    from scipy.sparse import triu
    from helpers import load_golden, csr_from
    g = load_golden("irmm_rounds")
    monkeypatch.setattr(ah, "REFERENCE_INPLACE_A", inplace)
    pa, wk = g["allelic_pairs"], g["inplace_weak_pairs"]
    P_al = [pa[:, 0], pa[:, 1]]
    P_wk = [np.concatenate([wk[:, 0], wk[:, 1]]), np.concatenate([wk[:, 1], wk[:, 0]])]
    for plain in (True, False):                    # a plain scipy matrix and a device-backed one
        A0 = csr_from(g, "inplace_A0")
        A = A0.copy() if plain else ah.wrap_adjacency(__import__("cphasing_b200.graphops", fromlist=["x"]).upload_adjacency(A0))
        for i, res in enumerate((1.0, 1.2, 1.4)):
            raw_A, A_ret, ca, K = ah.IRMM(None, A, None, P_al, P_wk, -1, 0.3, res, 1.0, 0.01, 1, method="leiden")
            ref_call = i if inplace else 0
            n = A0.shape[0]
            want = csr_matrix((g[f"inplace_call{ref_call}_w"], (g[f"inplace_call{ref_call}_edges"][:, 0],
                                                                g[f"inplace_call{ref_call}_edges"][:, 1])), shape=(n, n))
            got = triu(csr_matrix(A_ret), k=1).tocsr()
            got.eliminate_zeros()
            d = abs(got - want)
            assert d.nnz == 0 or d.max() <= 1e-5 * abs(want).max()
            assert (got.data < 0).sum() == (want.data < 0).sum()
            caller = csr_matrix(A).toarray()
            ref_caller = csr_from(g, f"inplace_call{i}_callerA").toarray() if inplace else A0.toarray()
            assert np.allclose(caller, ref_caller, rtol=1e-5, atol=1e-7)
            assert np.allclose(csr_matrix(raw_A).toarray(), csr_from(g, f"inplace_call{ref_call}_rawA").toarray() if inplace
                               else A0.toarray(), rtol=1e-5, atol=1e-7)


def test_kprune_against_reference_run():
    """KPruneHyperGraph on the device against the lists of the REFERENCE's own KPruneHyperGraph.run()
    (tests/golden/ref_kprune.npz), identical including the (c, c) entries of the weak list."""
    import pandas as pd
    from cphasing_b200 import kprune
    from helpers import load_golden, csr_from
    g = load_golden("kprune")
    A = csr_from(g, "A")
    n = A.shape[0]
    names = [f"c{i}" for i in range(n)]
    rows = [(0, 0, names[a], names[b], *m, 1) for (a, b), m in zip(g["allele_pairs"], g["allele_meta"])]
    df = pd.DataFrame(rows, columns=[-1, 0, 1, 2, 3, 4, 5, 6, 7]).set_index(-1)

    class AT:
        data = df
    kp = kprune.KPruneHyperGraph(AT(), A, {nm: i for i, nm in enumerate(names)})
    kp.run()
    assert kp.allelic_prune_list == [tuple(x) for x in g["allelic"]]
    assert kp.weak_prune_list == [tuple(x) for x in g["weak"]]
    pt = kp.get_prune_table()
    assert len(pt) == len(g["allelic"]) + len(g["weak"])
    # second case: the contig index covers only part of the allele table (kprune.py:437-447)
    keep = g["sub_keep"]
    sub_names = {names[int(o)]: j for j, o in enumerate(keep)}
    kp2 = kprune.KPruneHyperGraph(AT(), A[keep][:, keep].tocsr(), sub_names)
    kp2.run()
    assert sorted(kp2.allele_group_db.items()) == [tuple(x) for x in g["sub_allele_group"]]
    assert kp2.allelic_prune_list == [tuple(x) for x in g["sub_allelic"]]
    assert kp2.weak_prune_list == [tuple(x) for x in g["sub_weak"]]
