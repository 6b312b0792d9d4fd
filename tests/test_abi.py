"""CPU-side checks of the drop-in boundary: libcph.so loads, exports every symbol include/cph.h
declares, and the ctypes table in cphasing_b200/_lib.py covers exactly that set.  No compute calls."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "cph.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(cph_[a-z0-9_]+)\s*\(", src)))


def test_header_and_ctypes_table_agree():
    from cphasing_b200 import _lib
    assert header_symbols() == _lib.declared_symbols()


def test_library_exports_every_declared_symbol():
    from cphasing_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        import __graft_entry__
        __graft_entry__.build()
    L = ctypes.CDLL(_lib.LIB_PATH)
    for s in header_symbols():
        assert hasattr(L, s), f"{s} declared in include/cph.h but not exported by libcph.so"
    assert _lib.lib().cph_version() >= 100
    assert _lib.lib().cph_launch_count() == 0


def test_product_never_imports_oracle():
    """the oracle is test infrastructure (tier rule 3): nothing under cphasing_b200/ may import it."""
    bad = []
    for d, _, files in os.walk(os.path.join(ROOT, "cphasing_b200")):
        for f in files:
            if f.endswith(".py"):
                s = open(os.path.join(d, f)).read()
                if re.search(r"^\s*(from|import)\s+oracle\b", s, flags=re.M):
                    bad.append(f)
    assert not bad, bad


def test_no_cuda_means_loud_failure():
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    from cphasing_b200 import device as D, _lib
    import numpy as np
    with pytest.raises(_lib.CphError):
        D.porec_to_edges(np.zeros(1, np.uint32), np.zeros(1, np.uint32), np.zeros(1, np.int8), 1)


def test_kprune_cli_contract_and_unpinned_modes():
    """`cphasing kprune AlleleTable INPUT_CONTACTS_PATH -o out -n .. -m .. -fc .. -t ..` (cli.py:3525-3645) parses; options whose
    meaning only the absent cphasing-rs binary defines refuse loudly instead of being ignored."""
    from cphasing_b200 import cli
    from cphasing_b200.kprune import KPrunerRust, UNPINNED
    a = cli.build_parser().parse_args(["kprune", "allele.table", "c.contacts", "-o", "p.table", "-n", "cis", "-m", "fast",
                                       "-fc", "first.clusters.txt", "-t", "8"])
    assert (a.alleletable, a.contacts, a.output, a.norm_method, a.method, a.first_cluster, a.threads) == \
        ("allele.table", "c.contacts", "p.table", "cis", "fast", "first.clusters.txt", 8)
    d = cli.build_parser().parse_args(["kprune", "a", "c"])
    assert (d.output, d.method, d.norm_method, d.threads) == ("prune.contig.table", "precise", "none", 4)
    for kw in (dict(method="greedy"), dict(method="fast", norm_method="cis_unique"), dict(method="fast", norm_method="auto"),
               dict(method="fast", count_re="x.counts")):
        with pytest.raises(NotImplementedError, match="parity unpinned"):
            KPrunerRust("a", "c", "o", **kw)
    assert "cphasing-rs" in UNPINNED
