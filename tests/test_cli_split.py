"""`--split-length` handling of the command line (cli.py:4757-4790) against the reference's own utilities (`binnify`,
`determine_split_length`, utilities.py:1430-1480, 1709-1744): the `auto` rule, bins in name order, trimmed ends."""
import os
import sys

import numpy as np
import pandas as pd
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "baseline"))


def test_split_and_auto_rule_match_reference():
    import ref_loader
    if not ref_loader.available():
        pytest.skip("the reference install (baseline/_ref) is not present")
    ref_loader.load()
    from cphasing.utilities import binnify, determine_split_length
    from cphasing_b200 import cli
    rng = np.random.default_rng(1)
    seen_auto = set()
    for case in range(4):
        L = rng.integers(50_000, 8_000_000, 40)
        if case:
            L[3] = 120_000_000                   # one outlier contig: `auto` switches the splitting on
        if case == 2:
            L[7] = 30_000
        if case == 3:
            L = L[:8]                            # fewer than 10 contigs: never split
        sizes = pd.DataFrame({"length": L}, index=[f"ctg{int(i):03d}" for i in rng.permutation(len(L))])
        a, b = cli.auto_split_length(sizes), determine_split_length(sizes)
        assert a == b
        seen_auto.add(a)
        sl = a or 3_000_000
        for trim in (0, 25_000):
            mine, first = cli.split_contigs(sizes, sl, trim)
            ref = binnify(sizes["length"], sl, trim_length=trim)
            names = [f"{c}|{int(s)}_{int(e)}" for c, s, e in zip(ref["chrom"], ref["start"], ref["end"])]
            assert list(mine.index) == names
            assert list(mine["length"]) == list((ref["end"] - ref["start"]).astype(int))
            assert first == ref.reset_index().groupby("chrom")["index"].apply(lambda x: x.min()).to_dict()
    assert None in seen_auto and len(seen_auto) > 1


def test_whitelist_and_split_length_together(tmp_path):
    """`hypergraph -sl N -wl file`: the rows are the split bins of EVERY contig (the reference overwrites the whitelisted
    list with all bins, cli.py:3890-3906) and each contig's first-bin index points into that same numbering; without
    splitting the whitelist selects the indexed contigs."""
    from types import SimpleNamespace
    from cphasing_b200 import cli
    sizes = pd.DataFrame({"length": [2_500_000, 900_000, 4_100_000]}, index=pd.Index(["ctgB", "ctgA", "ctgC"], name="contig"))
    wl = tmp_path / "wl.txt"
    wl.write_text("ctgC\nctgA\n")
    a = SimpleNamespace(split_length="1m", extract_whitelist=str(wl), alleles_trim_length=0)
    sz, idx, sl, first = cli._index_contigs(a, sizes)
    assert sl == 1_000_000
    assert list(sz.index) == ["ctgA|0_900000", "ctgB|0_1000000", "ctgB|1000000_2000000", "ctgB|2000000_2500000",
                              "ctgC|0_1000000", "ctgC|1000000_2000000", "ctgC|2000000_3000000", "ctgC|3000000_4000000",
                              "ctgC|4000000_4100000"]
    assert first == {"ctgA": 0, "ctgB": 1, "ctgC": 4}
    for c, f in first.items():                        # row of an alignment = first[c] + mid // sl lands on a bin of c
        nb = -(-int(sizes.loc[c, "length"]) // sl)
        assert all(sz.index[f + k].split("|")[0] == c for k in range(nb))
        assert idx[sz.index[f]] == f
    b = SimpleNamespace(split_length="none", extract_whitelist=str(wl), alleles_trim_length=0)
    sz, idx, sl, first = cli._index_contigs(b, sizes)
    assert sl is None and first is None and list(sz.index) == ["ctgA", "ctgC"] and idx == {"ctgA": 0, "ctgC": 1}
