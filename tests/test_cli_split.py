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
