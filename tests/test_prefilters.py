"""Extractor pre-filters the reference's default CLI can reach: the HCR-BED midpoint filter (hypergraph.py:1721-1752 for
Pore-C tables, :786-893 for pairs) and the split-contig row index (:1406-1424, :994-1023).

CPU: the oracle's per-record restatement of polars' join_asof(backward) against pandas.merge_asof (an independent
implementation of the same join).  GPU: the kernels (cph_porec_to_edges_pf / cph_pairs_to_edges_pf) bit-exact against the
oracle, through the device module and through the reference-facing classes with a BED file on disk."""
import numpy as np
import pandas as pd
import pytest

from oracle import hypergraph_oracle as orc


def make_bed(g, rng, per_contig=3, nested=True):
    names, st, en = [], [], []
    for c in range(g.n):
        L = int(g.length[c])
        k = int(rng.integers(0, per_contig + 1))
        for _ in range(k):
            a = int(rng.integers(0, max(L - 10, 1)))
            b = int(min(L, a + rng.integers(1, max(L // 3, 2))))
            names.append(c); st.append(a); en.append(b)
        if nested and k and rng.random() < 0.3:          # an interval nested in the previous one: join_asof sees only the
            names.append(c); st.append(st[-1] + 1); en.append(st[-1] + 2)   # LAST start <= pos
    o = rng.permutation(len(names))                       # file order is arbitrary
    return np.array(names)[o], np.array(st, np.uint32)[o], np.array(en, np.uint32)[o]


def split_maps(g, split_length):
    """split bins as cli.py:4776-4790 builds them: init index per original contig, names `ctg|start_end`, bin lengths"""
    init, names, lens = [], [], []
    for c in range(g.n):
        L = int(g.length[c])
        init.append(len(names))
        for a in range(0, L, split_length):
            b = min(L, a + split_length)
            names.append(f"{g.names[c]}|{a}_{b}")
            lens.append(b - a)
    return np.array(init, np.uint32), names, np.array(lens, np.int64)


def test_hcr_restatement_matches_merge_asof():
    from cphasing_b200 import synth
    rng = np.random.default_rng(3)
    g = synth.make_genome(60, 3, 2, seed=4)
    bn, bs, be = make_bed(g, rng)
    n = 20000
    names = rng.integers(0, g.n, n)
    pos = (rng.random(n) * g.length[names]).astype(np.int64)
    got = orc.hcr_hits(names, pos, bn, bs, be)
    left = pd.DataFrame({"chrom": names, "pos": pos, "i": np.arange(n)}).sort_values("pos", kind="stable")
    right = pd.DataFrame({"chrom": bn, "pos": bs.astype(np.int64), "end_bed": be.astype(np.int64)}).sort_values("pos", kind="stable")
    j = pd.merge_asof(left, right, on="pos", by="chrom", direction="backward").sort_values("i")
    want = (j["pos"] < j["end_bed"]).fillna(False).to_numpy() & j["end_bed"].notna().to_numpy()
    assert np.array_equal(got, want)
    assert 0.05 < got.mean() < 0.95


def test_split_index_toy():
    # two contigs of 250 and 90 bp, 100-bp bins: rows 0,1,2 | 3 ; name code 2 is not in the map
    row, ok = orc.split_index(np.array([0, 0, 0, 1, 2]), np.array([0, 100, 249, 89, 5]), np.array([0, 3, -1]), 100)
    assert ok.tolist() == [True, True, True, True, False] and row[:4].tolist() == [0, 1, 2, 3]


def _triplets(D, raw):
    row, col, cnt, mq = D.edges_to_triplets(raw)
    return D.to_host(row), D.to_host(col), D.to_host(cnt).view(np.uint32), D.to_host(mq)


@pytest.fixture(scope="module")
def api():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from cphasing_b200 import device as D, hypergraph as hg, synth
    return D, hg, synth


@pytest.mark.gpu
@pytest.mark.parametrize("hcr,invert,split_length,shuffle", [(True, False, 0, False), (True, True, 0, False),
                                                             (False, False, 700_000, False), (True, False, 500_000, False),
                                                             (True, True, 900_000, True)])
def test_porec_prefilters_bit_exact(api, tmp_path, hcr, invert, split_length, shuffle):
    D, hg, synth = api
    rng = np.random.default_rng(11)
    g = synth.make_genome(120, 4, 2, seed=12)
    t = synth.porec_table(g, 30000, seed=13, law="poisson")
    if shuffle:
        p = rng.permutation(len(t["read_idx"]))
        t = {k: v[p] for k, v in t.items()}
    bn, bs, be = make_bed(g, rng)
    n_rows, init, names, lens = g.n, None, None, None
    if split_length:
        init, names, lens = split_maps(g, split_length)
        n_rows = len(names)
    exp = orc.porec_edges(t["read_idx"], t["contig_idx"], t["mapq"], t["start"], t["end"], g.length, min_mapq=1, min_order=2,
                          max_order=50, edge_length=300000, n_contigs=n_rows,
                          hcr=(bn, bs, be, invert) if hcr else None,
                          split=(init.astype(np.int64), split_length) if split_length else None)
    assert len(exp[0]) > 100
    pf = D.PreFilter(n_names=g.n, hcr_invert=invert)
    if hcr:
        o = np.lexsort((bs, bn))
        pf.bed_ptr = np.concatenate(([0], np.cumsum(np.bincount(bn, minlength=g.n)))).astype(np.int64)
        pf.bed_start, pf.bed_end = bs[o], be[o]
    if split_length:
        pf.split_length, pf.split_init = split_length, init
    raw = D.porec_to_edges(t["read_idx"], t["contig_idx"], t["mapq"], n_rows, start=t["start"], end=t["end"],
                           contig_len=g.length, min_mapq=1, min_order=2, max_order=50, edge_length=300000, prefilter=pf)
    for a, b in zip(_triplets(D, raw), exp):
        assert a.dtype == b.dtype and np.array_equal(a, b)
    # through the reference-facing class, BED read from a file, names as the CLI builds them
    bed = tmp_path / "hcr.bed"
    pd.DataFrame({"c": [g.names[i] for i in bn], "s": bs, "e": be}).to_csv(bed, sep="\t", header=False, index=False)
    if split_length:
        contig_idx = {nm: i for i, nm in enumerate(names)}
        sizes = dict(zip(names, lens.tolist()))
        boundarys = {g.names[c]: int(init[c]) for c in range(g.n)}
    else:
        contig_idx, sizes, boundarys = g.contig_idx, g.contigsizes, None
    ex = hg.HyperExtractor.from_arrays(t["read_idx"], t["contig_idx"], t["mapq"], contig_idx, sizes, start=t["start"],
                                       end=t["end"], min_quality=1, edge_length=300000, hcr_bed=str(bed) if hcr else None,
                                       hcr_invert=invert, split_length=split_length or None, split_contig_boundarys=boundarys)
    e = ex.edges
    for a, b in zip((e.row, e.col, e.count, e.mapq), exp):
        assert np.array_equal(np.asarray(a), b)


@pytest.mark.gpu
@pytest.mark.parametrize("hcr,invert,split_length", [(True, False, 0), (True, True, 0), (False, False, 600_000)])
def test_pairs_prefilters_bit_exact(api, hcr, invert, split_length):
    D, hg, synth = api
    rng = np.random.default_rng(21)
    g = synth.make_genome(70, 4, 2, seed=22)
    t = synth.pairs_table(g, 60000, seed=23)
    bn, bs, be = make_bed(g, rng)
    n_rows, init, names, lens = g.n, None, None, None
    if split_length:
        init, names, lens = split_maps(g, split_length)
        n_rows = len(names)
    exp = orc.pairs_edges(t["chrom1"], t["chrom2"], t["mapq"], t["pos1"], t["pos2"], g.length, min_mapq=1, edge_length=0,
                          n_contigs=n_rows, hcr=(bn, bs, be, invert) if hcr else None,
                          split=(init.astype(np.int64), split_length) if split_length else None)
    assert len(exp[0]) > 100
    if split_length:
        contig_idx = {nm: i for i, nm in enumerate(names)}
        sizes = dict(zip(names, lens.tolist()))
        boundarys = {g.names[c]: int(init[c]) for c in range(g.n)}
    else:
        contig_idx, sizes, boundarys = g.contig_idx, g.contigsizes, None
    ex = hg.Extractor.from_arrays(t["chrom1"], t["chrom2"], t["mapq"], contig_idx, sizes, pos1=t["pos1"], pos2=t["pos2"],
                                  min_quality=1, edge_length=0,
                                  hcr_bed=([g.names[i] for i in bn], bs, be) if hcr else None, hcr_invert=invert,
                                  split_length=split_length or None, split_contig_boundarys=boundarys)
    e = ex.edges
    got = np.lexsort((np.asarray(e.mapq), np.asarray(e.col), np.asarray(e.row)))
    want = np.lexsort((exp[3], exp[1], exp[0]))
    for a, b in zip((e.row, e.col, e.count, e.mapq), exp):
        assert np.array_equal(np.asarray(a)[got], b[want])
