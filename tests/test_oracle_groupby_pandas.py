"""The table -> triplets step has TWO independent CPU restatements (numpy in oracle/hypergraph_oracle.py, pandas in
oracle/groupby_pandas.py, each citing the reference's polars lines): they must agree on generated tables that hit the
cases the reference's filters distinguish -- duplicated (read, contig) records, contigs missing from the index (inner
join), reads of order 1 and beyond max_order, MAPQ ties and sub-threshold MAPQ, the edge filter (the goldens start from
triplets, so they hold no raw table to add here).  polars itself is absent, so neither is pinned to it; this removes the single-restatement risk."""
import numpy as np
import pandas as pd
import pytest
from hypothesis import given, settings, strategies as st

from oracle import groupby_pandas as gp
from oracle import hypergraph_oracle as orc


def make_index(n, n_unknown, rng):
    names = [f"ctg{i:04d}" for i in range(n + n_unknown)]
    known = names[:n]
    lens = {c: int(rng.integers(20_000, 3_000_000)) for c in names}
    return names, {c: i for i, c in enumerate(known)}, lens


def porec_table(rng, n_reads, names, lens, max_len=70):
    k = np.minimum(rng.geometric(0.25, n_reads), max_len)               # alignments per read: 1 .. 70
    read = np.repeat(np.arange(n_reads), k) + 5                          # read ids need not start at 0
    pool = rng.choice(len(names), size=(n_reads, 6))                     # few contigs per read -> duplicates are common
    ci = pool[np.repeat(np.arange(n_reads), k), rng.integers(0, 6, read.size)]
    wide = rng.random(n_reads) < 0.05                                    # some reads touch many contigs (order > max_order)
    w = np.repeat(wide, k)
    ci[w] = rng.integers(0, len(names), int(w.sum()))
    chrom = np.array(names, dtype=object)[ci]
    L = np.array([lens[c] for c in chrom])
    start = (rng.random(read.size) * L * 0.98).astype(np.int64)
    end = np.minimum(start + rng.integers(50, 5000, read.size), L)
    mapq = rng.choice([0, 1, 1, 5, 30, 60, 60], read.size).astype(np.int8)
    return pd.DataFrame({"read_idx": read.astype(np.uint32), "chrom": chrom, "start": start.astype(np.uint32),
                         "end": end.astype(np.uint32), "mapping_quality": mapq})


def codes(series, contig_idx):
    """what the product's ingest hands the kernels and the numpy oracle: index of the name, n_contigs for unknown names"""
    n = len(contig_idx)
    return series.map(lambda c: contig_idx.get(c, n)).to_numpy().astype(np.uint32)


@pytest.mark.parametrize("seed", range(6))
@pytest.mark.parametrize("edge_length", [0, 100_000])
@pytest.mark.parametrize("min_q,min_o,max_o", [(1, 2, 50), (0, 2, 5), (30, 3, 8), (1, 1, 50)])
def test_porec_group_by_two_engines_agree(seed, edge_length, min_q, min_o, max_o):
    rng = np.random.default_rng(seed)
    names, idx, lens = make_index(40, 4, rng)
    df = porec_table(rng, 3000, names, lens)
    a = gp.porec_canonical(gp.porec_triplets(df, idx, lens, min_quality=min_q, min_order=min_o, max_order=max_o,
                                             edge_length=edge_length))
    clen = np.array([lens[c] for c in idx], dtype=np.int64)
    b = orc.porec_edges(df["read_idx"].to_numpy(), codes(df["chrom"], idx), df["mapping_quality"].to_numpy(),
                        start=df["start"].to_numpy(), end=df["end"].to_numpy(), contig_len=clen, min_mapq=min_q,
                        min_order=min_o, max_order=max_o, edge_length=edge_length, n_contigs=len(idx))
    assert len(a[0]) > 20
    for x, y, what in zip(a, b, ("row", "col", "count", "mapq")):
        assert x.dtype == y.dtype and np.array_equal(x, y), what
    assert edge_length or a[2].max() > 1                   # duplicated (read, contig) records were merged into counts


@given(st.lists(st.tuples(st.integers(0, 6), st.integers(0, 5), st.sampled_from([0, 1, 2, 60])), min_size=0, max_size=40),
       st.integers(0, 2), st.integers(1, 3), st.integers(2, 4))
@settings(max_examples=200, deadline=None)
def test_porec_group_by_small_tables(records, min_q, min_o, max_o):
    """tiny tables, every combination: empty input, a single read, all records filtered, contig 5 unknown"""
    idx = {f"c{i}": i for i in range(5)}
    df = pd.DataFrame({"read_idx": np.array([r for r, _, _ in records], dtype=np.uint32),
                       "chrom": np.array([f"c{c}" for _, c, _ in records], dtype=object),
                       "start": np.zeros(len(records), np.uint32), "end": np.ones(len(records), np.uint32),
                       "mapping_quality": np.array([q for _, _, q in records], dtype=np.int8)})
    a = gp.porec_canonical(gp.porec_triplets(df, idx, None, min_quality=min_q, min_order=min_o, max_order=max_o))
    b = orc.porec_edges(df["read_idx"].to_numpy(), codes(df["chrom"], idx), df["mapping_quality"].to_numpy(), min_mapq=min_q,
                        min_order=min_o, max_order=max_o, n_contigs=5)
    for x, y in zip(a, b):
        assert np.array_equal(x, y)


@pytest.mark.parametrize("seed", range(6))
@pytest.mark.parametrize("edge_length", [0, 150_000])
@pytest.mark.parametrize("min_q", [0, 1, 30])
def test_pairs_two_engines_agree(seed, edge_length, min_q):
    rng = np.random.default_rng(100 + seed)
    names, idx, lens = make_index(30, 3, rng)
    m = 5000
    c1 = rng.integers(0, len(names), m)
    c2 = np.where(rng.random(m) < 0.3, c1, rng.integers(0, len(names), m))       # intra-contig pairs: duplicated (row, col)
    L1 = np.array([lens[names[c]] for c in c1]); L2 = np.array([lens[names[c]] for c in c2])
    p = pd.DataFrame({"chrom1": np.array(names, dtype=object)[c1], "pos1": (rng.random(m) * L1).astype(np.int64),
                      "chrom2": np.array(names, dtype=object)[c2], "pos2": (rng.random(m) * L2).astype(np.int64),
                      "mapq": rng.choice([0, 1, 20, 60], m).astype(np.int8)})
    a = gp.pairs_canonical(gp.pairs_triplets(p, idx, lens, min_mapq=min_q, edge_length=edge_length))
    clen = np.array([lens[c] for c in idx], dtype=np.int64)
    b = orc.pairs_edges(codes(p["chrom1"], idx), codes(p["chrom2"], idx), p["mapq"].to_numpy(), pos1=p["pos1"].to_numpy(),
                        pos2=p["pos2"].to_numpy(), contig_len=clen, min_mapq=min_q, edge_length=edge_length, n_contigs=len(idx))
    assert len(a[0]) > 50
    for x, y, what in zip(a, b, ("row", "col", "count", "mapq")):
        assert x.dtype == y.dtype and np.array_equal(x, y), what
