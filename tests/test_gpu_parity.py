"""GPU parity: the CUDA path (through the libcph C-ABI and the reference-facing classes) against the
reference's golden outputs and the CPU oracle.  Integer / index work is bit-exact; weighted A is held
to 1e-6 relative against the fp64 oracle (north_star) and 3e-6 against the reference's own float32."""
import numpy as np
import pytest
from scipy.sparse import csr_matrix

from helpers import GOLDEN_CASES, load_golden, csr_from, canon, canon_dup, assert_same_structure_close
from oracle import hypergraph_oracle as orc

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def api():
    import cphasing_b200.algorithms.hypergraph as ah
    import cphasing_b200.hypergraph as hg
    from cphasing_b200 import device as D, synth
    return ah, hg, D, synth


def _edges(ah, g):
    n = len(g["lengths"])
    names = [f"c{i}" for i in range(n)]
    return ah.HyperEdges(idx={c: i for i, c in enumerate(names)}, row=g["in_row"].copy(), col=g["in_col"].copy(),
                         count=g["in_count"].copy(), contigsizes={c: int(l) for c, l in zip(names, g["lengths"])},
                         mapq=g["in_mapq"].copy()), names


def _build_H(ah, g):
    mq, mc, mw = g["params"]
    edges, names = _edges(ah, g)
    HG = ah.HyperGraph(edges, min_quality=int(mq), mapq_filter=True)
    if "remove" in g:
        HG.remove_rows([names[i] for i in g["remove"]])
    H = HG.incidence_matrix(min_contacts=int(mc))
    return HG, H, names


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_incidence_matches_reference(api, name):
    ah = api[0]
    g = load_golden(name)
    HG, H, names = _build_H(ah, g)
    ref = csr_from(g, "H")
    assert H.shape == ref.shape
    assert HG.shape == ref.shape
    for x, y in zip(canon_dup(H), canon_dup(ref)):
        assert np.array_equal(x, y)
    assert H.dtype == np.float32
    idx = {c: i for i, c in enumerate(names)}
    assert np.array_equal(np.array([idx[c] for c in HG.nodes]), g["H_nodes_idx"])
    if int(g["params"][1]):
        assert np.array_equal(np.asarray(HG.remove_contig_idx), g["H_remove_contig_idx"])
    # the triplet attributes of the reference class
    assert np.array_equal(np.sort(HG.row), np.sort(g["gd_row"])) or "remove" in g


@pytest.mark.parametrize("name", GOLDEN_CASES)
@pytest.mark.parametrize("from_host", [False, True])
def test_clique_expansion_matches_reference(api, name, from_host):
    ah = api[0]
    g = load_golden(name)
    mw = float(g["params"][2])
    if from_host:
        H = csr_from(g, "H")                         # plain scipy input: uploaded by clique_expansion_init
    else:
        _, H, _ = _build_H(ah, g)                    # device twin travels with the matrix
    Hs = csr_from(g, "H")
    cases = [("A", dict(min_weight=mw)), ("A0", dict(min_weight=0)), ("Au", dict(min_weight=mw, use_high_order=False))]
    if from_host or int(g["params"][1]):
        cases.append(("Aw", dict(min_weight=mw, W=g["W"])))
    for key, kw in cases:
        A = ah.HyperGraph.clique_expansion_init(H, **kw)
        assert A.dtype == np.float32 and A.shape == (Hs.shape[0],) * 2
        ref = csr_from(g, key)
        thr = kw["min_weight"] if kw["min_weight"] > 0 else None
        assert_same_structure_close(A, ref, rtol=3e-6, thr=thr)
        # fp64 gate: 1e-6 relative against the exact fp64 oracle
        ex = orc.clique_expansion_exact(Hs, kw.get("use_high_order", True), kw.get("W"))
        h = ah.device_handle(A)
        A64 = csr_matrix((h.data64.cpu().numpy(), h.indices.cpu().numpy(), h.indptr.cpu().numpy()), shape=A.shape)
        if thr is None:
            exc = canon(ex); exc.eliminate_zeros()
            assert np.array_equal(canon(A64).indices, exc.indices)
            np.testing.assert_allclose(canon(A64).data, exc.data, rtol=1e-9)
        else:
            E = ex.tocsr().copy(); E.data[E.data.astype(np.float32) < np.float32(thr)] = 0; E.eliminate_zeros()
            assert_same_structure_close(A64, E, rtol=1e-6, thr=thr)


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_contact_counts_bit_exact(api, name):
    ah, _, D, _ = api
    g = load_golden(name)
    Hs = csr_from(g, "H")
    dH = ah.upload_incidence(Hs)
    A, acc = D.clique_expansion(dH, use_high_order=False, min_weight=0.0)
    assert acc.frac_bits == 0
    C = canon(orc.contact_counts(Hs)); C.eliminate_zeros()
    got = csr_matrix((A.data64.cpu().numpy(), A.indices.cpu().numpy(), A.indptr.cpu().numpy()), shape=C.shape)
    assert np.array_equal(got.indptr, C.indptr) and np.array_equal(got.indices, C.indices)
    assert np.array_equal(got.data.astype(np.int64), C.data)      # bit-exact integers


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_normalise_reweight_extract(api, name):
    ah = api[0]
    g = load_golden(name)
    A = csr_from(g, "A")
    w = g["NW_diag_sqrt"]
    An = ah.HyperGraph.normalize_adjacency_matrix(A, w)
    ref = canon(csr_from(g, "An"))
    assert np.array_equal(canon(An).indices, ref.indices)
    np.testing.assert_allclose(canon(An).data, ref.data, rtol=1e-6)
    # dense NW accepted too (the reference's calling convention)
    if A.shape[0] <= 200:
        An2 = ah.HyperGraph.normalize_adjacency_matrix(A, np.outer(w, w).astype(np.float32))
        np.testing.assert_allclose(canon(An2).data, ref.data, rtol=2e-6)
    # fused normalisation inside the expansion epilogue
    H = csr_from(g, "H")
    Af = ah.HyperGraph.clique_expansion_init(H, NW=w, min_weight=float(g["params"][2]))
    # entries that sit exactly on the cut (e.g. 6 * 1/3 = 2.0) are kept or dropped by the reference depending on its
    # float32 rounding of sqrt(d)^2, so the threshold-aware comparison is used, scaled by the entry's node weights
    mw = float(g["params"][2])
    if mw == 0:
        assert_same_structure_close(Af, ref, rtol=3e-6)
    else:
        U = ah.HyperGraph.clique_expansion_init(H, min_weight=mw)        # same pattern as Af, un-normalised
        assert np.array_equal(canon(U).indices, canon(Af).indices)
        assert_same_structure_close(U, csr_from(g, "A"), rtol=3e-6, thr=mw)
        common = canon(Af).multiply(canon(ref) != 0).tocsr(); common.eliminate_zeros()
        rc = canon(ref).multiply(canon(Af) != 0).tocsr(); rc.eliminate_zeros()
        np.testing.assert_allclose(canon(common).data, canon(rc).data, rtol=3e-6)
    # extract_incidence_matrix2
    sub, rem, kept = ah.extract_incidence_matrix2(H, g["sub_idx"])
    rs = csr_from(g, "subH")
    assert sub.shape == rs.shape
    for x, y in zip(canon_dup(sub), canon_dup(rs)):
        assert np.array_equal(x, y)
    assert np.array_equal(np.asarray(kept, dtype=np.int64), g["sub_kept"])
    # find_csr_data_offsets + IRMM-style re-weighting
    from cphasing_b200 import graphops
    a = graphops.upload_adjacency(A)
    off = graphops.find_pairs(a, g["pair_r"], g["pair_c"]).cpu().numpy()
    assert np.array_equal(np.sort(off[off >= 0]), g["pair_offsets"])
    P = (g["pair_r"], g["pair_c"])
    Ar = ah.HyperGraph.reweight_pairs(A, P, None, -1.0)
    exp = canon(A).copy(); exp.data[g["pair_offsets"]] *= -1.0
    assert np.array_equal(canon(Ar).data, exp.data)


def _oracle_vs_device_porec(D, t, n, **kw):
    exp = orc.porec_edges(t["read_idx"], t["contig_idx"], t["mapq"], t.get("start"), t.get("end"), kw.pop("contig_len", None),
                          n_contigs=n, **kw)
    return exp


@pytest.mark.parametrize("law,edge_length,min_mapq,shuffle", [("poisson", 0, 1, False), ("exp", 0, 0, False),
                                                             ("poisson", 300000, 1, False), ("poisson", 0, 10, True),
                                                             ("exp", 500000, 30, True)])
def test_porec_edges_bit_exact(api, law, edge_length, min_mapq, shuffle):
    _, hg, D, synth = api
    g = synth.make_genome(150, 5, 2, seed=5)
    t = synth.porec_table(g, 20000, seed=6, law=law)
    if shuffle:
        p = np.random.default_rng(1).permutation(len(t["read_idx"]))
        t = {k: v[p] for k, v in t.items()}
    exp = orc.porec_edges(t["read_idx"], t["contig_idx"], t["mapq"], t["start"], t["end"], g.length, min_mapq=min_mapq,
                          min_order=2, max_order=50, edge_length=edge_length, n_contigs=g.n)
    raw = D.porec_to_edges(t["read_idx"], t["contig_idx"], t["mapq"], g.n, start=t["start"], end=t["end"],
                           contig_len=g.length, min_mapq=min_mapq, min_order=2, max_order=50, edge_length=edge_length)
    row, col, cnt, mq = D.edges_to_triplets(raw)
    got = (row.cpu().numpy(), col.cpu().numpy(), cnt.view(__import__("torch").int32).cpu().numpy().view(np.uint32),
           mq.cpu().numpy())
    for a, b in zip(got, exp):
        assert a.dtype == b.dtype and np.array_equal(a, b)
    assert raw.n_cols == (int(exp[1].max()) + 1 if len(exp[1]) else 0)
    # and through the reference-facing class
    ex = hg.HyperExtractor.from_arrays(t["read_idx"], t["contig_idx"], t["mapq"], g.contig_idx, g.contigsizes,
                                       start=t["start"], end=t["end"], min_quality=min_mapq, edge_length=edge_length)
    assert np.array_equal(np.asarray(ex.edges.row), exp[0]) and np.array_equal(np.asarray(ex.edges.count), exp[2])


def test_porec_order_bounds_and_empty(api):
    _, _, D, _ = api
    # read 0: one contig twice (order 1 -> dropped); read 1: contigs 0,1 ; read 2: 3 contigs (max_order=2 -> dropped)
    r = np.array([0, 0, 1, 1, 1, 2, 2, 2], np.uint32)
    c = np.array([4, 4, 0, 1, 0, 1, 2, 3], np.uint32)
    q = np.array([60, 60, 5, 60, 50, 60, 60, 60], np.int8)
    raw = D.porec_to_edges(r, c, q, 5, min_mapq=1, min_order=2, max_order=2)
    row, col, cnt, mq = [x.cpu().numpy() for x in D.edges_to_triplets(raw)]
    assert row.tolist() == [0, 1] and col.tolist() == [1, 1] and cnt.view(np.uint32).tolist() == [2, 1]
    assert mq.tolist() == [50, 60]
    raw = D.porec_to_edges(r, c, q, 5, min_mapq=61)
    assert raw.nnz == 0 and raw.n_cols == 0
    raw = D.porec_to_edges(np.zeros(0, np.uint32), np.zeros(0, np.uint32), np.zeros(0, np.int8), 5)
    assert raw.nnz == 0


@pytest.mark.parametrize("edge_length,min_mapq", [(0, 1), (400000, 1), (0, 0)])
def test_pairs_edges_bit_exact(api, edge_length, min_mapq):
    _, hg, D, synth = api
    g = synth.make_genome(80, 4, 2, seed=8)
    t = synth.pairs_table(g, 50000, seed=9)
    exp = orc.pairs_edges(t["chrom1"], t["chrom2"], t["mapq"], t["pos1"], t["pos2"], g.length, min_mapq=min_mapq,
                          edge_length=edge_length, n_contigs=g.n)
    ex = hg.Extractor.from_arrays(t["chrom1"], t["chrom2"], t["mapq"], g.contig_idx, g.contigsizes, pos1=t["pos1"],
                                  pos2=t["pos2"], min_quality=min_mapq, edge_length=edge_length)
    e = ex.edges
    # duplicated (row, col) of intra-contig pairs: compare as multisets per (row, col, mapq)
    got = np.lexsort((np.asarray(e.mapq), np.asarray(e.col), np.asarray(e.row)))
    want = np.lexsort((exp[3], exp[1], exp[0]))
    for a, b in zip((e.row, e.col, e.count, e.mapq), exp):
        assert np.array_equal(np.asarray(a)[got], b[want])


def test_hyperedges_msgpack_roundtrip(api, tmp_path):
    ah, hg, D, synth = api
    import msgspec
    g = synth.make_genome(30, 3, 2, seed=3)
    t = synth.porec_table(g, 500, seed=4)
    ex = hg.HyperExtractor.from_arrays(t["read_idx"], t["contig_idx"], t["mapq"], g.contig_idx, g.contigsizes, edge_length=0)
    p = tmp_path / "x.hg"
    ex.save(str(p))
    e2 = msgspec.msgpack.decode(open(p, "rb").read(), type=ah.HyperEdges, dec_hook=ah.numpy_dec_hook)
    e2.to_numpy()
    assert np.array_equal(e2.row, np.asarray(ex.edges.row)) and e2.idx == g.contig_idx
    H1 = ah.HyperGraph(e2, min_quality=1).incidence_matrix(5)
    H2 = ah.HyperGraph(ex.edges, min_quality=1).incidence_matrix(5)
    assert (H1 != H2).nnz == 0


def test_empty_edges_raise(api):
    ah = api[0]
    e = ah.HyperEdges(idx={"a": 0}, row=np.zeros(0, np.int32), col=np.zeros(0, np.int32), count=np.zeros(0, np.uint32),
                      contigsizes={"a": 10}, mapq=np.zeros(0, np.int8))
    with pytest.raises(ValueError, match="No hyperedges found"):
        ah.HyperGraph(e)


def test_expansion_properties_medium(api):
    """size-independent properties at a size the oracle does not need: symmetry, the sum identity
    sum_ij C_ij = sum_e (sum_i c_ie)^2, and idempotence of the build."""
    ah, hg, D, synth = api
    import torch
    g = synth.make_genome(1500, 6, 4, seed=41)
    t = synth.porec_table(g, 300000, seed=42, law="exp")
    raw = D.porec_to_edges(t["read_idx"], t["contig_idx"], t["mapq"], g.n, min_mapq=1)
    H = D.incidence(raw, min_contacts=25)
    A, acc = D.clique_expansion(H, use_high_order=False, min_weight=0.0)
    M = csr_matrix((A.data64.cpu().numpy(), A.indices.cpu().numpy(), A.indptr.cpu().numpy()), shape=(A.n, A.n))
    assert (abs(M - M.T)).nnz == 0
    cnt = (H.pk.view(torch.int32) & 0xFFFFFF).to(torch.int64)
    colsum = torch.zeros(H.n_cols, dtype=torch.int64, device=cnt.device).index_add_(0, H.indices.long(), cnt)
    assert int(M.sum()) == int((colsum * colsum).sum().item())
    A2, _ = D.clique_expansion(H, use_high_order=False, min_weight=0.0)
    assert torch.equal(A2.data64, A.data64) and torch.equal(A2.indices, A.indices)
    # weighted: deterministic run to run (fixed-point accumulation)
    B1, _ = D.clique_expansion(H, use_high_order=True, min_weight=0.1)
    B2, _ = D.clique_expansion(H, use_high_order=True, min_weight=0.1)
    assert torch.equal(B1.data64, B2.data64)
    Hs = ah.wrap_incidence(H)
    ex = orc.clique_expansion_exact(csr_matrix(Hs), True)
    E = ex.tocsr(); E.data[E.data.astype(np.float32) < np.float32(0.1)] = 0; E.eliminate_zeros()
    B = csr_matrix((B1.data64.cpu().numpy(), B1.indices.cpu().numpy(), B1.indptr.cpu().numpy()), shape=E.shape)
    assert_same_structure_close(B, E, rtol=1e-6, thr=0.1)


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_porec_long_reads_and_tile_boundaries(api, seed):
    """reads far longer than the shared-memory window (slow cooperative path), reads straddling tile
    boundaries, invalid records inside runs -- against the oracle, bit-exact."""
    _, _, D, _ = api
    import torch
    rng = np.random.default_rng(seed)
    n = 300
    lens = []
    while sum(lens) < 60000:
        u = rng.random()
        if u < 0.002:
            lens.append(int(rng.integers(1100, 9000)))      # longer than PT_HALO / whole tiles
        elif u < 0.02:
            lens.append(int(rng.integers(100, 1000)))
        else:
            lens.append(int(rng.integers(1, 12)))
    lens = np.array(lens)
    read = np.repeat(np.arange(len(lens), dtype=np.uint32) * 3 + 5, lens)
    ncont = np.where(lens > 60, rng.integers(2, 40, size=len(lens)), 1000)
    contig = np.concatenate([rng.integers(0, min(n, k), size=l) for l, k in zip(lens, ncont)]).astype(np.uint32)
    contig[rng.random(contig.size) < 0.02] = n + 7          # outside the contig index -> dropped
    mapq = rng.choice(np.array([0, 3, 20, 60], dtype=np.int8), size=contig.size)
    for min_mapq, max_order in ((1, 50), (0, 8), (21, 50)):
        exp = orc.porec_edges(read, contig, mapq, min_mapq=min_mapq, min_order=2, max_order=max_order, n_contigs=n)
        raw = D.porec_to_edges(read, contig, mapq, n, min_mapq=min_mapq, min_order=2, max_order=max_order, read_offset=11)
        row, col, cnt, mq = D.edges_to_triplets(raw)
        got = (row.cpu().numpy(), col.cpu().numpy() - 11, cnt.view(torch.int32).cpu().numpy().view(np.uint32), mq.cpu().numpy())
        for a, b in zip(got, exp):
            assert np.array_equal(a, b)
        assert raw.n_cols == (int(exp[1].max()) + 12 if len(exp[1]) else 0)
        # head bits: exactly one per surviving read, on its first entry
        heads = (raw.row.view(torch.int32) < 0).cpu().numpy()
        cols = raw.col.view(torch.int32).cpu().numpy()
        assert heads.sum() == len(np.unique(cols))
        assert np.all(heads[np.concatenate(([True], cols[1:] != cols[:-1]))])


def test_incidence_long_hyperedges(api):
    """hyperedges longer than the tile window (slow cooperative path) and the int8 wrap of D_e
    (algorithms/hypergraph.py:385) -- against the oracle."""
    ah, _, D, _ = api
    from scipy.sparse import random as sprandom, hstack, csr_matrix as csr
    rng = np.random.default_rng(3)
    n = 3500
    base = sprandom(n, 4000, density=0.001, random_state=5, format="csc")
    base.data = rng.integers(1, 4, size=base.nnz).astype(np.float32)
    dense_cols = []
    for k in (1500, 3300, 130, 257):
        col = np.zeros((n, 1), np.float32)
        col[rng.choice(n, size=k, replace=False), 0] = rng.integers(1, 3, size=k)
        dense_cols.append(csr(col))
    H = hstack([base[:, :2000], dense_cols[0], dense_cols[1], base[:, 2000:], dense_cols[2], dense_cols[3]]).tocsr().astype(np.float32)
    A = ah.HyperGraph.clique_expansion_init(H, min_weight=0)
    ex = canon(orc.clique_expansion_exact(H, True)); ex.eliminate_zeros()
    h = ah.device_handle(A)
    A64 = csr((h.data64.cpu().numpy(), h.indices.cpu().numpy(), h.indptr.cpu().numpy()), shape=A.shape)
    assert np.array_equal(canon(A64).indptr, ex.indptr) and np.array_equal(canon(A64).indices, ex.indices)
    np.testing.assert_allclose(canon(A64).data, ex.data, rtol=1e-9)
    # sub-incidence of the same matrix (packed input) keeps / drops hyperedges like the reference
    idx = np.arange(0, n, 2)
    sub, rem, kept = ah.extract_incidence_matrix2(H, idx)
    so, ro, ko = orc.extract_incidence_matrix2(H, idx)
    assert sub.shape == so.shape
    for x, y in zip(canon_dup(sub), canon_dup(so)):
        assert np.array_equal(x, y)
    assert np.array_equal(kept, ko)


def test_porec_edges_unaligned_device_columns(api):
    """the bulk (TMA) window loads need 16-byte aligned columns; views that start 4 bytes off take the plain-load path
    and must give the same triplets."""
    import torch
    _, hg, D, synth = api
    g = synth.make_genome(150, 5, 2, seed=5)
    t = synth.porec_table(g, 20000, seed=6)
    exp = orc.porec_edges(t["read_idx"], t["contig_idx"], t["mapq"], min_mapq=1, min_order=2, max_order=50, n_contigs=g.n)
    dev = torch.device("cuda")
    pad = lambda a, dt: torch.cat([torch.zeros(1, dtype=dt, device=dev), torch.from_numpy(a.view(
        {torch.int32: np.int32, torch.int8: np.int8}[dt])).to(dev)])[1:]
    r = pad(t["read_idx"], torch.int32).view(torch.uint32)
    c = pad(t["contig_idx"], torch.int32).view(torch.uint32)
    q = pad(t["mapq"], torch.int8)
    assert r.data_ptr() % 16 != 0
    for cols in ((r, c, q), (r.clone(), c.clone(), q.clone())):          # unaligned views, then aligned copies
        raw = D.porec_to_edges(cols[0], cols[1], cols[2], g.n, min_mapq=1, min_order=2, max_order=50)
        row, col, cnt, mq = D.edges_to_triplets(raw)
        got = (row.cpu().numpy(), col.cpu().numpy(), cnt.view(torch.int32).cpu().numpy().view(np.uint32), mq.cpu().numpy())
        for a, b in zip(got, exp):
            assert np.array_equal(a, b)


@pytest.mark.parametrize("chunk", [3584, 3 * 3584, 1 << 20])
@pytest.mark.parametrize("pinned", [False, True])
@pytest.mark.parametrize("pack", [0, 3])
def test_porec_host_streamed_chunks(api, chunk, pinned, pack, monkeypatch):
    """cph_porec_to_edges_host: host columns copied in chunks with the tile kernel launched per landed chunk must give
    exactly what the one-launch device path and the oracle give -- long reads crossing chunk borders included.
    pack = 3: three host threads re-encode the chunks (read-id deltas, 16-bit contigs) and the device restores them."""
    _, _, D, _ = api
    import torch
    rng = np.random.default_rng(chunk % 97 + pinned)
    n = 200
    lens = []
    while sum(lens) < 50000:
        u = rng.random()
        lens.append(int(rng.integers(3000, 9000)) if u < 0.002 else int(rng.integers(400, 900)) if u < 0.01
                    else int(rng.integers(1, 12)))
    lens = np.array(lens)
    read = np.repeat(np.arange(len(lens), dtype=np.uint32) * 2 + 1, lens)
    ncont = np.where(lens > 60, rng.integers(2, 45, size=len(lens)), 1000)
    contig = np.concatenate([rng.integers(0, min(n, k), size=l) for l, k in zip(lens, ncont)]).astype(np.uint32)
    mapq = rng.choice(np.array([0, 3, 20, 60], dtype=np.int8), size=contig.size)
    monkeypatch.setattr(D, "STREAM_MIN_RECORDS", 1)
    monkeypatch.setattr(D, "STREAM_CHUNK_RECORDS", chunk)
    monkeypatch.setattr(D, "STREAM_PACK_CHUNK_RECORDS", chunk)
    monkeypatch.setattr(D, "STREAM_PACK_THREADS", pack)
    cols = (read, contig, mapq)
    if pinned:
        cols = tuple(torch.from_numpy(x.view(np.int32) if x.dtype == np.uint32 else x).pin_memory() for x in cols)
    exp = orc.porec_edges(read, contig, mapq, min_mapq=1, min_order=2, max_order=50, n_contigs=n)
    from cphasing_b200 import _lib
    l0 = _lib.launches()
    raw = D.porec_to_edges(*cols, n, min_mapq=1, min_order=2, max_order=50, read_offset=3)
    nl = _lib.launches() - l0
    if chunk < 50000:
        assert nl >= 2, "expected one kernel launch per landed chunk"
    row, col, cnt, mq = D.edges_to_triplets(raw)
    got = (row.cpu().numpy(), col.cpu().numpy() - 3, cnt.view(torch.int32).cpu().numpy().view(np.uint32), mq.cpu().numpy())
    for a, b in zip(got, exp):
        assert np.array_equal(a, b)
    # a table that is NOT grouped by read falls through to the general (sorting) path on the staged columns
    p = rng.permutation(read.size)
    raw2 = D.porec_to_edges(read[p], contig[p], mapq[p], n, min_mapq=1, min_order=2, max_order=50, read_offset=3)
    row2, col2, cnt2, mq2 = D.edges_to_triplets(raw2)
    assert np.array_equal(row2.cpu().numpy(), exp[0]) and np.array_equal(col2.cpu().numpy() - 3, exp[1])
    assert np.array_equal(cnt2.view(torch.int32).cpu().numpy().view(np.uint32), exp[2])


@pytest.mark.parametrize("n", [300, 70000])
def test_porec_host_packed_fallbacks(api, n, monkeypatch):
    """Packed host transfer: chunks whose read ids jump by more than 255 travel unpacked next to packed ones, and more
    than 65536 contigs switch the packed contig column to 32 bits -- same edges as the oracle either way."""
    _, _, D, _ = api
    import torch
    rng = np.random.default_rng(n)
    lens = rng.integers(1, 9, size=40000)
    ids = np.cumsum(rng.integers(1, 4, size=lens.size)).astype(np.int64)
    ids[lens.size // 3:] += 1000            # one big step: that chunk cannot be delta-encoded
    ids[2 * lens.size // 3:] += 255         # a step of 255+ at another chunk
    read = np.repeat(ids.astype(np.uint32), lens)
    contig = rng.integers(0, n, size=read.size).astype(np.uint32)
    k = contig[1::5].size
    contig[:5 * k:5] = contig[1::5]                         # duplicates inside reads
    mapq = rng.choice(np.array([0, 3, 20, 60], dtype=np.int8), size=contig.size)
    monkeypatch.setattr(D, "STREAM_MIN_RECORDS", 1)
    monkeypatch.setattr(D, "STREAM_PACK_CHUNK_RECORDS", 4096)
    monkeypatch.setattr(D, "STREAM_PACK_THREADS", 4)
    exp = orc.porec_edges(read, contig, mapq, min_mapq=1, min_order=2, max_order=50, n_contigs=n)
    raw = D.porec_to_edges(read, contig, mapq, n, min_mapq=1, min_order=2, max_order=50)
    row, col, cnt, mq = D.edges_to_triplets(raw)
    got = (row.cpu().numpy(), col.cpu().numpy(), cnt.view(torch.int32).cpu().numpy().view(np.uint32), mq.cpu().numpy())
    for a, b in zip(got, exp):
        assert np.array_equal(a, b)


def test_extract_rows_with_explicit_zero_counts(api):
    """extract_incidence_matrix2 (algorithms/hypergraph.py:938-954) on a host matrix that stores explicit zeros: the
    column test is on the SUMMED count, so a hyperedge with two surviving entries of which one is a stored 0 must go
    (the kernel's `k >= col_min_sum` shortcut has to switch itself off for such tiles)."""
    ah = api[0]
    rng = np.random.default_rng(11)
    n, m = 60, 9000
    k = rng.integers(2, 7, size=m)
    cols = np.repeat(np.arange(m), k)
    rows = np.concatenate([rng.choice(n, size=x, replace=False) for x in k])
    data = rng.integers(0, 3, size=rows.size).astype(np.float32)          # about a third explicit zeros
    H = csr_matrix((data, (rows, cols)), shape=(n, m))
    assert (H.data == 0).sum() > 1000                                      # zeros really are stored
    idx = np.sort(rng.choice(n, size=25, replace=False))
    sub, removed, kept = ah.extract_incidence_matrix2(H, idx)
    eo, ro, ko = orc.extract_incidence_matrix2(H, idx)
    assert np.array_equal(kept, ko) and np.array_equal(np.asarray(removed), ro)
    assert sub.shape == eo.shape
    for x, y in zip(canon_dup(sub), canon_dup(eo)):
        assert np.array_equal(x, y)


@pytest.mark.parametrize("form", ["swept", "swept-fallback", "binned"])
@pytest.mark.parametrize("use_high_order", [True, False])
def test_binned_expansion_is_bit_identical(api, use_high_order, form, monkeypatch):
    """Accumulators beyond L2 (n > 5.6k) take the swept kernels (runs sorted by contig and packed to 4 bytes, then one
    streaming pass per L2-sized row band of the triangle) or the binned ones (pair records grouped by 2-D block).  Integer
    fixed-point sums do not depend on the order of addition, so triangle and diagonal must equal the direct kernel's bit
    for bit -- including large counts (swept: up to its count field; beyond it the device flag sends the call to the
    direct kernel; binned: pairs whose count product does not fit the 34-bit record field)."""
    _, _, D, synth = api
    import torch
    g = synth.make_genome(6000, 10, 4, seed=91)
    t = synth.porec_table_device(g, 1_500_000, 92, law="poisson", first=0, device="cuda")
    raw = D.porec_to_edges(t["read_idx"], t["contig_idx"], t["mapq"], g.n, min_mapq=1)
    H = D.incidence(raw, min_contacts=5, min_mapq=0)
    assert H.n_rows > D.BINNED_MIN_N and H.nnz >= D.BINNED_MIN_NNZ
    pk = H.pem_pk.view(torch.int32)
    big = torch.arange(0, H.nnz, 100_003, device=pk.device)
    hi, lo = (2_000, 1_500) if form == "swept" else (200_000, 150_000)   # n ~ 6k: 13 row bits, counts < 2^11 fit the field
    pk[big] = (pk[big] & ~0x00FFFFFF) | hi               # products above 2^34 with one another are unlikely; with
    pk[big + 1] = (pk[big + 1] & ~0x00FFFFFF) | lo       # the neighbour entry (same hyperedge most of the time) they occur
    H._bound = None
    monkeypatch.setenv("CPH_EXPAND_FORM", "binned" if form == "binned" else "swept")
    monkeypatch.setenv("CPH_BAND_MB", "3")              # ~48 bands at this n
    l0 = __import__("cphasing_b200")._lib.launches()
    a1 = D.expand_accumulate(H, use_high_order=use_high_order)
    binned_launches = __import__("cphasing_b200")._lib.launches() - l0
    monkeypatch.setattr(D, "BINNED_MIN_N", 1 << 30)
    a2 = D.expand_accumulate(H, use_high_order=use_high_order)
    assert binned_launches >= 3                          # binned: count, (scan..,) fill, apply; swept: prep + one per band
    if form == "swept":
        assert binned_launches >= 40                     # it really took the band passes (no fallback)
    assert a1.frac_bits == a2.frac_bits
    assert torch.equal(a1.diag, a2.diag)
    assert torch.equal(a1.acc, a2.acc)
    assert int((a1.acc != 0).sum().item()) > 1_000_000


@pytest.mark.parametrize("longest", [64, 65, 90])
def test_swept_expansion_run_length_boundary(api, longest, monkeypatch):
    """the swept form holds hyperedge runs of up to 64 entries in its halo; a longer run raises the device flag and the call
    falls back to the direct kernel -- either way triangle and diagonal equal the direct kernel's"""
    ah, _, D, _ = api
    import torch
    from scipy.sparse import csr_matrix
    rng = np.random.default_rng(longest)
    n, m = 300, 4000
    k = rng.integers(2, 9, size=m)
    k[17] = longest                                      # one long hyperedge in the middle of a tile
    k[m - 1] = min(longest, 64)                          # and one at the very end
    cols = np.repeat(np.arange(m), k)
    rows = np.concatenate([rng.choice(n, size=x, replace=False) for x in k])
    data = rng.integers(1, 4, size=rows.size).astype(np.float32)
    H = ah.upload_incidence(csr_matrix((data, (rows, cols)), shape=(n, m)))
    monkeypatch.setenv("CPH_EXPAND_FORM", "direct")
    ref = D.expand_accumulate(H, use_high_order=True)
    monkeypatch.setenv("CPH_EXPAND_FORM", "swept")
    monkeypatch.setenv("CPH_BAND_MB", "0.05")
    l0 = __import__("cphasing_b200")._lib.launches()
    got = D.expand_accumulate(H, use_high_order=True)
    launches = __import__("cphasing_b200")._lib.launches() - l0
    assert torch.equal(got.acc, ref.acc) and torch.equal(got.diag, ref.diag)
    if longest <= 64:
        assert launches >= 5                              # the band passes did run
    else:
        assert launches <= 3                              # flag read before any band: prep, then the direct kernel only
