"""Host-side table decode (cphasing_b200/ingest.py + csrc/ingest.cu): the multi-threaded TSV parser of the
C-ABI library against pandas on the same text, block / thread boundaries, gzip, malformed input, PQS directories.
No GPU needed: the parser is host code inside libcph.so."""
import gzip

import numpy as np
import pandas as pd
import pytest

from cphasing_b200 import ingest


def porec_text(n_reads=3000, n_contigs=40, seed=3, crlf=False):
    rng = np.random.default_rng(seed)
    names = [f"ctg{i:04d}" if i % 3 else f"scaffold_{i}|arrow" for i in range(n_contigs)]
    lines = ["#read_idx\tread_length\tread_start\tread_end\tstrand\tchrom\tstart\tend\tmapping_quality\tidentity\tfilter"]
    rows = []
    for r in range(n_reads):
        for _ in range(int(rng.integers(1, 7))):
            c = names[int(rng.integers(0, n_contigs))] if rng.random() > 0.03 else "not_in_index"
            st = int(rng.integers(0, 4_000_000_000))
            en = min(st + int(rng.integers(1, 5000)), 4_294_967_295)
            mq = int(rng.integers(0, 61))
            rows.append((r * 3 + 7, c, st, en, mq))
            lines.append(f"{r * 3 + 7}\t12000\t10\t900\t+\t{c}\t{st}\t{en}\t{mq}\t0.97\tpass")
        if r % 500 == 0:
            lines.append("# a comment in the middle")
            lines.append("")
    text = ("\r\n" if crlf else "\n").join(lines) + ("\r\n" if crlf else "\n")
    return text, names, rows


def expect(rows, names):
    idx = {c: i for i, c in enumerate(names)}
    return {"read_idx": np.array([r[0] for r in rows], dtype=np.uint32),
            "contig": np.array([idx.get(r[1], len(names)) for r in rows], dtype=np.uint32),
            "start": np.array([r[2] for r in rows], dtype=np.uint32), "end": np.array([r[3] for r in rows], dtype=np.uint32),
            "mapq": np.array([r[4] for r in rows], dtype=np.int8)}


@pytest.mark.parametrize("crlf", [False, True])
@pytest.mark.parametrize("threads", [1, 7])
def test_parse_block_matches_the_rows(threads, crlf):
    text, names, rows = porec_text(crlf=crlf)
    nt = ingest.NameTable({c: i for i, c in enumerate(names)})
    cols, bad = ingest.parse_block(text.encode(), ingest.POREC_FIELDS, nt, threads=threads)
    want = expect(rows, names)
    assert bad == 0
    for k, v in want.items():
        assert cols[k].dtype == v.dtype and np.array_equal(cols[k], v), k


def test_agrees_with_pandas_on_a_gzip_file_in_many_blocks(tmp_path):
    text, names, rows = porec_text(n_reads=20000, seed=5)
    p = tmp_path / "t.porec.gz"
    with gzip.open(p, "wt") as fh:
        fh.write(text)
    df = pd.read_csv(p, sep="\t", header=None, comment="#", usecols=[0, 5, 6, 7, 8],
                     names=["read_idx", "chrom", "start", "end", "mapq"])
    idx = {c: i for i, c in enumerate(names)}
    parts = list(ingest.read_table_chunks(p, ingest.POREC_FIELDS, idx, threads=3, block_bytes=50_000))
    assert len(parts) > 10                                     # really went block by block
    cols = {k: np.concatenate([c[k] for c, _ in parts]) for k in ("read_idx", "contig", "start", "end", "mapq")}
    assert sum(b for _, b in parts) == 0
    assert np.array_equal(cols["read_idx"], df["read_idx"].to_numpy().astype(np.uint32))
    assert np.array_equal(cols["contig"], df["chrom"].map(lambda c: idx.get(c, len(names))).to_numpy().astype(np.uint32))
    assert np.array_equal(cols["start"], df["start"].to_numpy().astype(np.uint32))
    assert np.array_equal(cols["end"], df["end"].to_numpy().astype(np.uint32))
    assert np.array_equal(cols["mapq"], df["mapq"].to_numpy().astype(np.int8))
    whole, bad = ingest.read_table(p, ingest.POREC_FIELDS, idx, threads=2)
    assert bad == 0 and all(np.array_equal(whole[k], cols[k]) for k in cols)


def test_malformed_lines_are_counted_and_skipped_in_order():
    good = "5\tx\tx\tx\tx\tA\t1\t2\t60\n"
    text = good + "6\tx\tx\n" + good.replace("5\t", "7\t") + "8\tx\tx\tx\tx\tA\tone\t2\t60\n" + good.replace("5\t", "9\t") + \
        "10\tx\tx\tx\tx\tB\t3\t4\t200\n"                       # short line, non-numeric field; MAPQ 200 saturates
    big = (text * 3000).encode()                               # several thread ranges, gaps closed in file order
    nt = ingest.NameTable({"A": 0, "B": 1})
    cols, bad = ingest.parse_block(big, ingest.POREC_FIELDS, nt, threads=5)
    assert bad == 2 * 3000
    assert np.array_equal(cols["read_idx"], np.tile(np.array([5, 7, 9, 10], dtype=np.uint32), 3000))
    assert np.array_equal(cols["mapq"][:4], np.array([60, 60, 60, 127], dtype=np.int8))
    assert np.array_equal(cols["contig"][:4], np.array([0, 0, 0, 1], dtype=np.uint32))


def test_name_table_with_many_names_and_pairs_fields():
    names = {f"c{i}_{i * 7919 % 1013}": i for i in range(5000)}
    lines = [f"r{i}\t{nm}\t{i}\tc{(i * 31) % 5000}_{(i * 31) % 5000 * 7919 % 1013}\t{i + 1}\t+\t-\t{i % 61}\tUU\n"
             for nm, i in names.items()]
    cols, bad = ingest.parse_block("".join(lines).encode(), ingest.PAIRS_FIELDS, ingest.NameTable(names), threads=4)
    assert bad == 0
    assert np.array_equal(cols["chrom1"], np.arange(5000, dtype=np.uint32))
    assert np.array_equal(cols["chrom2"], (np.arange(5000) * 31 % 5000).astype(np.uint32))
    assert np.array_equal(cols["pos2"], np.arange(1, 5001, dtype=np.uint32))
    assert np.array_equal(cols["mapq"], (np.arange(5000) % 61).astype(np.int8))


def test_pqs_directory(tmp_path):
    import pyarrow as pa
    import pyarrow.parquet as pq
    d = tmp_path / "x.pqs"
    (d / "q0").mkdir(parents=True)
    (d / "q1").mkdir()
    (d / "_metadata").write_text(str({"is_pqs": True, "format": "pairs", "is_with_mapq": True}))
    (d / "_contigsizes").write_text("a\t100\nb\t200\n")
    for sub, mq in (("q0", [0, 5, 0, 60]), ("q1", [5, 60])):
        for part in (0, 1, 10):
            t = pa.table({"read_idx": list(range(len(mq))), "chrom1": ["a", "b", "a", "b"][:len(mq)], "pos1": [1, 2, 3, 4][:len(mq)],
                          "chrom2": ["b", "b", "zz", "a"][:len(mq)], "pos2": [9, 8, 7, 6][:len(mq)], "mapq": mq})
            pq.write_table(t, d / sub / f"{part}.parquet")
    assert ingest.is_pqs(d) and not ingest.is_pqs(tmp_path)
    files, with_mapq = ingest.pqs_chunks(d, 1)
    assert [f.name for f in files] == ["0.parquet", "1.parquet", "10.parquet"] and files[0].parent.name == "q1" and with_mapq
    assert ingest.pqs_chunks(d, 0)[0][0].parent.name == "q0"
    idx = {"a": 0, "b": 1}
    lut = lambda ser: np.array([idx.get(c, 2) for c in ser], dtype=np.uint32)
    c = ingest.read_pqs_chunk(files[0], lut, 30, with_mapq)
    assert c["chrom1"].tolist() == [1] and c["chrom2"].tolist() == [1] and c["mapq"].tolist() == [60]
    c = ingest.read_pqs_chunk(files[0], lut, 1, with_mapq)
    assert c["mapq"].tolist() == [5, 60] and c["pos1"].dtype == np.uint32
