"""The oracle against the reference's own outputs (tests/golden/ref_*.npz, made by
tests/golden/make_golden.py from /root/reference) -- pins the restatement."""
import numpy as np
import pytest
from scipy.sparse import csr_matrix

from oracle import hypergraph_oracle as orc
from helpers import GOLDEN_CASES, load_golden, csr_from, canon, canon_dup


def _prep(g):
    mq, mc, mw = g["params"]
    row, col, count, ncol = orc.hypergraph_get_data(g["in_row"], g["in_col"], g["in_count"], g["in_mapq"],
                                                    min_quality=int(mq))
    n = len(g["lengths"])
    if "remove" in g:
        k = ~np.isin(row, g["remove"])
        row, col, count = row[k], col[k], count[k]
    return row, col, count, n, ncol, int(mc), float(mw)


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_get_data_and_incidence(name):
    g = load_golden(name)
    mq = int(g["params"][0])
    row, col, count, ncol = orc.hypergraph_get_data(g["in_row"], g["in_col"], g["in_count"], g["in_mapq"], mq)
    assert np.array_equal(row, g["gd_row"]) and np.array_equal(col, g["gd_col"])
    assert np.array_equal(count, g["gd_count"]) and ncol == g["gd_shape"][1]
    row, col, count, n, ncol, mc, mw = _prep(g)
    H, rk, ck = orc.incidence_matrix(row, col, count, n, ncol, mc)
    Href = csr_from(g, "H")
    assert H.shape == Href.shape
    a = canon_dup(H); b = canon_dup(Href)
    for x, y in zip(a, b):
        assert np.array_equal(x, y)
    assert np.array_equal(np.where(rk)[0], g["H_nodes_idx"])
    if mc:
        assert np.array_equal(np.where(~rk)[0], g["H_remove_contig_idx"])


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_clique_expansion(name):
    g = load_golden(name)
    mw = float(g["params"][2])
    H = csr_from(g, "H")
    # literal float32 replay: identical pattern; values equal up to f32 summation order
    for key, kw in (("A", {}), ("A0", {"min_weight": 0}), ("Au", {"use_high_order": False}), ("Aw", {"W": g["W"]})):
        kw = dict(min_weight=mw, **kw) if "min_weight" not in kw else kw
        ref = canon(csr_from(g, key))
        A32 = canon(orc.clique_expansion(H, dtype=np.float32, **kw))
        assert np.array_equal(A32.indptr, ref.indptr) and np.array_equal(A32.indices, ref.indices), key
        np.testing.assert_allclose(A32.data, ref.data, rtol=2e-6)
        # fp64 gate vs the reference's float32 result: 1e-6 relative except entries at the threshold
        A64 = canon(orc.clique_expansion(H, dtype=np.float64, **kw))
        Aex = canon(orc.clique_expansion_exact(H, kw.get("use_high_order", True), kw.get("W")))
        if kw.get("min_weight", 0) == 0:
            assert np.array_equal(A64.indices, ref.indices)
            np.testing.assert_allclose(A64.data, ref.data, rtol=3e-6)
            assert np.array_equal(Aex.indices, ref.indices)
            np.testing.assert_allclose(Aex.data, A64.data, rtol=1e-12)


def test_toy_values():
    g = load_golden("toy_hic")
    A = csr_from(g, "A0").toarray()
    assert np.array_equal(A, np.array([[6, 2, 0], [2, 3, 1], [0, 1, 1]], dtype=np.float32))
    H = csr_from(g, "H")
    assert np.array_equal(orc.contact_counts(H).toarray(), A.astype(np.int64))


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_normalise_filter_extract(name):
    g = load_golden(name)
    H = csr_from(g, "H"); A = csr_from(g, "A")
    lengths = g["lengths"][g["H_nodes_idx"]]
    w = orc.normalize_weight_svc(H, lengths)
    np.testing.assert_allclose(w, g["NW_diag_sqrt"], rtol=1e-6)
    An = canon(orc.normalize_adjacency(A, w)); ref = canon(csr_from(g, "An"))
    assert np.array_equal(An.indices, ref.indices)
    np.testing.assert_allclose(An.data, ref.data, rtol=1e-6)
    assert np.array_equal(orc.filter_adjacency(A, lengths), g["filter_keep"])
    sub, rem, kept = orc.extract_incidence_matrix2(H, g["sub_idx"])
    ref = csr_from(g, "subH")
    assert sub.shape == ref.shape
    for x, y in zip(canon_dup(sub), canon_dup(ref)):
        assert np.array_equal(x, y)
    assert np.array_equal(np.asarray(kept, dtype=np.int64), g["sub_kept"])
    off = np.sort(orc.find_csr_data_offsets(A.tocsr(), g["pair_r"], g["pair_c"]))
    assert np.array_equal(off, g["pair_offsets"])


# ---- partition post-processing: oracle restatements vs outputs of the reference's own _merge / recluster_by_linkage
def _groups(labels):
    labels = np.asarray(labels)
    return [sorted(np.where(labels == g)[0].tolist()) for g in range(int(labels.max()) + 1)]


def _same_partition(K, labels):
    want = sorted(map(tuple, _groups(labels)))
    got = sorted(tuple(sorted(int(x) for x in g)) for g in K)
    return want == got


def test_oracle_merge_and_recluster_match_reference():
    g = load_golden("partition_post")
    A = csr_from(g, "A")
    n = A.shape[0]
    sizes = {i: float(g["lengths"][i]) for i in range(n)}
    K0 = _groups(g["K0_labels"])
    G = len(set(g["group"].tolist()))
    for method in ("mean", "sum"):
        K1 = orc.merge_groups(A, K0, sizes, G, method=method)
        assert _same_partition(K1, g[f"merge_{method}_labels"])
        assert [x[0] for x in K1] == g[f"merge_{method}_order"].tolist()
    aset = set(map(tuple, g["allelic_pairs"].tolist()))
    K2 = orc.merge_groups(A, K0, sizes, G, allelic_idx_set=aset, min_allelic_overlap=0.1)
    assert _same_partition(K2, g["merge_conflict_labels"])
    Ar = csr_from(g, "Ar")
    lab3 = g["K3_labels"]
    K3 = _groups(lab3)
    K4, moved = orc.recluster_by_linkage(Ar, K3, g["lengths"].astype(np.float32), 1.0)
    assert moved == int(g["recluster_count"][0]) and moved > 0
    assert _same_partition(K4, g["recluster_labels"])


def test_kprune_oracle_matches_reference_run():
    """oracle kprune_hypergraph against the lists of the REFERENCE's own KPruneHyperGraph.run() (tests/golden/
    make_golden_kprune.py: the reference class with igraph's 2x2 matching answered by a stand-in)."""
    g = load_golden("kprune")
    A = csr_from(g, "A")
    al, wk = orc.kprune_hypergraph(A, [tuple(x) for x in g["allele_pairs"]])
    assert al == [tuple(x) for x in g["allelic"]]
    assert wk == [tuple(x) for x in g["weak"]]
    assert sum(1 for a, b in wk if a == b) > 100          # the (c, c) entries of the reference's weak list
    grp = orc.allele_group_first([tuple(x) for x in g["allele_pairs"]])
    assert sorted(grp.items()) == [tuple(x) for x in g["allele_group"]]


def test_kprune_oracle_subset_index_matches_reference_run():
    """the contig index covers only part of the allele table: groups come from the FIRST listed partner of the unfiltered
    table and vanish when that partner is missing (kprune.py:437-447) -- second case of ref_kprune.npz"""
    g = load_golden("kprune")
    keep = g["sub_keep"]
    A = csr_from(g, "A")[keep][:, keep].tocsr()
    new = {int(o): j for j, o in enumerate(keep)}
    first = {}
    for a, b in g["allele_pairs"]:
        first.setdefault(int(a), int(b))
    first_group = {new[a]: new[b] for a, b in first.items() if a in new and b in new}
    assert sorted(first_group.items()) == [tuple(x) for x in g["sub_allele_group"]]
    pairs = [(new[int(a)], new[int(b)]) for a, b in g["allele_pairs"] if int(a) in new and int(b) in new]
    al, wk = orc.kprune_hypergraph(A, pairs, first_group=first_group)
    assert al == [tuple(x) for x in g["sub_allelic"]]
    assert wk == [tuple(x) for x in g["sub_weak"]]
    # and the shortcut the drop-in used to take (first SURVIVING partner) is NOT what the reference does
    assert sorted(orc.allele_group_first(pairs).items()) != sorted(first_group.items())
