"""Files in -> hyperedges out through the streaming decode (cphasing_b200/ingest.py feeding the extractor kernels chunk by
chunk): must equal the in-memory path bit for bit, whatever the block size; tables that are not ordered by read fall back
to the whole-table decode; pairs files and PQS directories feed `Extractor`."""
import gzip

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def write_porec(path, t, names, order=None):
    idx = np.arange(len(t["read_idx"])) if order is None else order
    op = gzip.open if str(path).endswith(".gz") else open
    with op(path, "wt") as fh:
        fh.write("#read_idx\tread_length\tread_start\tread_end\tstrand\tchrom\tstart\tend\tmapping_quality\tidentity\tname\n")
        for i in idx:
            fh.write(f"{t['read_idx'][i]}\t100\t0\t90\t+\t{names[t['contig_idx'][i]]}\t{t['start'][i]}\t{t['end'][i]}\t"
                     f"{t['mapq'][i]}\t0.9\tr\n")


def edges_tuple(e):
    return tuple(np.asarray(getattr(e, f)) for f in ("row", "col", "count", "mapq"))


@pytest.mark.parametrize("block", [40_000, 1 << 20])
def test_streamed_table_equals_in_memory_path(tmp_path, monkeypatch, block):
    from cphasing_b200 import hypergraph as hg, ingest, synth
    g = synth.make_genome(300, 5, 2, seed=8)
    t = synth.porec_table(g, 60_000, seed=9)
    p = tmp_path / "reads.porec.gz"
    write_porec(p, t, g.names)
    kw = dict(min_quality=1, edge_length=200_000)
    ref = hg.HyperExtractor.from_arrays(t["read_idx"], t["contig_idx"], t["mapq"], g.contig_idx, g.contigsizes,
                                        start=t["start"], end=t["end"], **kw)
    monkeypatch.setattr(ingest, "BLOCK_BYTES", block)
    calls = []
    real = hg.D.porec_to_edges
    monkeypatch.setattr(hg.D, "porec_to_edges", lambda *a, **k: (calls.append(1), real(*a, **k))[1])
    ex = hg.HyperExtractor([str(p)], g.contig_idx, g.contigsizes, threads=3, **kw)
    assert len(calls) > (20 if block < 100_000 else 0)          # small blocks: really chunk by chunk
    for a, b in zip(edges_tuple(ex.edges), edges_tuple(ref.edges)):
        assert np.array_equal(a, b)


def test_table_not_ordered_by_read_falls_back(tmp_path, monkeypatch):
    from cphasing_b200 import hypergraph as hg, ingest, synth
    g = synth.make_genome(200, 4, 2, seed=18)
    t = synth.porec_table(g, 20_000, seed=19)
    order = np.random.default_rng(1).permutation(len(t["read_idx"]))
    p = tmp_path / "shuffled.porec"
    write_porec(p, t, g.names, order)
    monkeypatch.setattr(ingest, "BLOCK_BYTES", 30_000)
    ex = hg.HyperExtractor([str(p)], g.contig_idx, g.contigsizes, min_quality=1, edge_length=0)
    ref = hg.HyperExtractor.from_arrays(t["read_idx"], t["contig_idx"], t["mapq"], g.contig_idx, g.contigsizes,
                                        min_quality=1, edge_length=0)
    # the group-by of an unordered table numbers nothing differently: same multiset of (row, col, count, mapq)
    key = lambda e: np.lexsort(edges_tuple(e)[::-1])
    for a, b in zip(edges_tuple(ex.edges), edges_tuple(ref.edges)):
        assert np.array_equal(a[key(ex.edges)], b[key(ref.edges)])


def test_pairs_file_and_pqs_directory(tmp_path):
    import pyarrow as pa
    import pyarrow.parquet as pq
    from cphasing_b200 import hypergraph as hg, synth
    g = synth.make_genome(150, 3, 2, seed=28)
    t = synth.pairs_table(g, 40_000, seed=29) if hasattr(synth, "pairs_table") else None
    if t is None:
        rng = np.random.default_rng(29)
        t = {"chrom1": rng.integers(0, g.n, 40_000).astype(np.uint32), "chrom2": rng.integers(0, g.n, 40_000).astype(np.uint32),
             "mapq": rng.integers(0, 61, 40_000).astype(np.int8)}
    m = len(t["mapq"])
    pos1 = np.arange(m, dtype=np.uint32) % 1000
    pos2 = (np.arange(m, dtype=np.uint32) * 7) % 1000
    names = np.array(g.names)
    pf = tmp_path / "x.pairs.gz"
    with gzip.open(pf, "wt") as fh:
        fh.write("## pairs format v1.0\n#columns: readID chrom1 pos1 chrom2 pos2 strand1 strand2 mapq\n")
        for i in range(m):
            fh.write(f"r{i}\t{names[t['chrom1'][i]]}\t{pos1[i]}\t{names[t['chrom2'][i]]}\t{pos2[i]}\t+\t-\t{t['mapq'][i]}\n")
    ref = hg.Extractor.from_arrays(t["chrom1"], t["chrom2"], t["mapq"], g.contig_idx, g.contigsizes, min_quality=1, edge_length=0)
    ex = hg.Extractor([str(pf)], g.contig_idx, g.contigsizes, min_quality=1, edge_length=0)
    for a, b in zip(edges_tuple(ex.edges), edges_tuple(ref.edges)):
        assert np.array_equal(a, b)
    # the same pairs as a PQS store: q1 holds the MAPQ >= 1 pairs in two chunks
    d = tmp_path / "x.pqs"
    (d / "q0").mkdir(parents=True); (d / "q1").mkdir()
    (d / "_metadata").write_text(str({"is_pqs": True, "format": "pairs", "is_with_mapq": True}))
    (d / "_contigsizes").write_text("".join(f"{n}\t{l}\n" for n, l in zip(g.names, g.length)))
    keep = np.flatnonzero(t["mapq"] >= 1)
    for part, sel in enumerate(np.array_split(keep, 2)):
        pq.write_table(pa.table({"chrom1": names[t["chrom1"][sel]], "pos1": pos1[sel], "chrom2": names[t["chrom2"][sel]],
                                 "pos2": pos2[sel], "mapq": t["mapq"][sel]}), d / "q1" / f"{part}.parquet")
    pq.write_table(pa.table({"chrom1": names[t["chrom1"]], "pos1": pos1, "chrom2": names[t["chrom2"]], "pos2": pos2,
                             "mapq": t["mapq"]}), d / "q0" / "0.parquet")
    exq = hg.Extractor([str(d)], g.contig_idx, g.contigsizes, min_quality=1, edge_length=0)
    refq = hg.Extractor.from_arrays(t["chrom1"][keep], t["chrom2"][keep], t["mapq"][keep], g.contig_idx, g.contigsizes,
                                    min_quality=1, edge_length=0)
    for a, b in zip(edges_tuple(exq.edges), edges_tuple(refq.edges)):
        assert np.array_equal(a, b)
    with pytest.raises(NotImplementedError):
        hg.Extractor([str(d)], g.contig_idx, g.contigsizes, min_quality=0, edge_length=0)
