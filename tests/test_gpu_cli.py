"""The command-line face (cphasing_b200/cli.py; reference cli.py:3525-5360) end to end on files: Pore-C table (TSV.gz) ->
`hypergraph` -> .hg (the reference's msgpack layout) -> `hyperpartition -n 5:4 -at ...` -> clusters file, and `kprune`."""
import gzip
import os
import subprocess
import sys

import numpy as np
import pytest

from oracle import community_oracle as co

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run(*args):
    r = subprocess.run([sys.executable, "-m", "cphasing_b200", *args], cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-3000:]
    return r


def test_cli_files_end_to_end(tmp_path):
    from cphasing_b200 import synth
    g = synth.make_genome(400, 5, 4, seed=20240)
    t = synth.porec_table(g, 120_000, seed=20240)
    names = np.array(g.names)
    table = tmp_path / "reads.porec.gz"
    with gzip.open(table, "wt") as fh:      # the 11 columns of a Pore-C table; the path reads 0, 5, 6, 7, 8 (hypergraph.py:1706-1717)
        for r, c, s, e, q in zip(t["read_idx"], t["contig_idx"], t["start"], t["end"], t["mapq"]):
            fh.write(f"{r}\t100\t0\t90\t+\t{names[c]}\t{s}\t{e}\t{q}\t0.9\tread{r}\n")
    sizes = tmp_path / "contigs.sizes"
    sizes.write_text("".join(f"{n}\t{l}\n" for n, l in zip(g.names, g.length)))
    at = tmp_path / "allele.table"
    with open(at, "w") as fh:
        for i, (a, b, m1, m2, sh, sim) in enumerate(synth.allele_pairs(g, seed=3)):
            fh.write(f"{i}\t{i}\t{names[a]}\t{names[b]}\t{m1}\t{m2}\t{sh}\t{sim:.4f}\t1\n")
    hgf = tmp_path / "reads.hg"
    run("hypergraph", str(table), str(sizes), str(hgf), "-e", "0")
    assert hgf.stat().st_size > 100_000
    out = tmp_path / "out.clusters.txt"
    run("hyperpartition", str(hgf), str(sizes), str(out), "-n", "5:4", "-at", str(at), "-ir1", "0.8", "-ir2", "0.8")
    lines = out.read_text().strip().split("\n")
    assert len(lines) == 20
    assert [l.split("\t")[0] for l in lines[:5]] == ["Chr01g1", "Chr01g2", "Chr01g3", "Chr01g4", "Chr02g1"]
    truth = {n: int(x) for n, x in zip(g.names, g.group)}
    lab, tru = [], []
    for i, l in enumerate(lines):
        f = l.split("\t")
        members = f[2].split(" ")
        assert int(f[1]) == len(members)
        lab += [i] * len(members); tru += [truth[m] for m in members]
    assert len(lab) > 0.95 * g.n and co.ari(lab, tru) >= 0.99
    # one round straight from the table, no .hg in between (cli.py --porec)
    out1 = tmp_path / "basal.clusters.txt"
    run("hyperpartition", str(table), str(sizes), str(out1), "--porec", "-e", "0", "-n", "20")
    l1 = out1.read_text().strip().split("\n")
    assert len(l1) == 20 and l1[0].startswith("Chr01\t")
    # kprune, the reference's file contract (cli.py:3525-3669): AlleleTable + 3-column contacts -> 7-column prune table
    # (core.py:711-723).  The contacts come from `utils hg2contacts` (cli.py:9290-9401), as in hyperpartition.py:1566-1572.
    contacts = tmp_path / "hypergraph.expansion.contacts"
    run("utils", "hg2contacts", str(hgf), str(sizes), "-o", str(contacts), "-mc", "5")
    first = contacts.read_text().split("\n")[0].split("\t")
    assert len(first) == 3 and first[0] in truth and float(first[2]) >= 1
    pt = tmp_path / "prune.table"
    run("kprune", str(at), str(contacts), "-o", str(pt), "-m", "fast", "-n", "none", "-t", "2")
    rows = [l.split("\t") for l in pt.read_text().strip().split("\n")]
    assert all(len(r) == 7 for r in rows) and {r[6] for r in rows} == {"0", "1"}
    # the same decision as the in-tree Python scorer restated by the oracle (KPruneHyperGraph, kprune.py:385-552) on the A
    # the contacts file describes
    from cphasing_b200.kprune import read_contacts, allele_frame
    from oracle import hypergraph_oracle as orc
    cn, A = read_contacts(str(contacts))
    ci = {c: i for i, c in enumerate(cn)}
    af = allele_frame(str(at))
    fg = af.groupby(1, sort=False)[2].first()
    first_group = {ci[k]: ci[v] for k, v in fg.items() if k in ci and v in ci}
    pairs = [(ci[a], ci[b]) for a, b in zip(af[1], af[2]) if a in ci and b in ci]
    al, wk = orc.kprune_hypergraph(A, pairs, first_group=first_group)
    got_al = sorted((ci[r[0]], ci[r[1]]) for r in rows if r[6] == "0")
    got_wk = sorted((ci[r[0]], ci[r[1]]) for r in rows if r[6] == "1")
    assert got_al == al and got_wk == wk
    # -fc: only contacts inside one first-round group are candidates -> a subset of the pairs above, none across groups
    pt2 = tmp_path / "prune.fc.table"
    chrom = tmp_path / "first.clusters.txt"                 # chromosome-level groups: the 4 haplotypes of a chromosome together
    run("hyperpartition", str(hgf), str(sizes), str(chrom), "-n", "5", "-ir1", "0.3", "-r1", "0.3")
    lc = chrom.read_text().strip().split("\n")
    run("kprune", str(at), str(contacts), "-o", str(pt2), "-m", "fast", "-fc", str(chrom))
    grp_of = {}
    for gi, l in enumerate(lc):
        for m in l.split("\t")[2].split(" "):
            grp_of[m] = gi
    rows2 = [l.split("\t") for l in pt2.read_text().strip().split("\n")]
    assert 0 < len(rows2) <= len(rows)
    assert all(grp_of[r[0]] == grp_of[r[1]] for r in rows2)
    # -m precise scores every listed allele (KPruner, kprune.py:269-335): runs and still writes both types
    pt3 = tmp_path / "prune.precise.table"
    run("kprune", str(at), str(contacts), "-o", str(pt3))
    assert {l.split("\t")[6] for l in pt3.read_text().strip().split("\n")} == {"0", "1"}
    # a command line written for the reference parses (cli.py:3945-4597 options that steer other stages are accepted, the ones
    # that would change the result here refuse loudly): --merge-cluster merges the 20 groups of `out` down to 5 (cli.py:5013)
    merged = tmp_path / "merged.clusters.txt"
    run("hyperpartition", str(hgf), str(sizes), str(merged), "-n", "5", "--merge-cluster", str(out), "-knm", "cis", "-apf", "3.0",
        "--disable-merge-use-allele", "-c", str(contacts), "--enable-misassembly-remove", "-alleles-k", "19")
    lm = merged.read_text().strip().split("\n")
    assert len(lm) == 5
    before = sorted(m for l in lines for m in l.split("\t")[2].split(" "))
    after = sorted(m for l in lm for m in l.split("\t")[2].split(" "))
    assert after == before                                                  # the same contigs, now in 5 groups (_merge, :2711)
    out_hg = tmp_path / "reads.hg2"
    r = subprocess.run([sys.executable, "-m", "cphasing_b200", "hyperpartition", str(hgf), str(sizes), str(tmp_path / "x.txt"),
                        "-n", "5:4", "--refine"], cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert r.returncode != 0 and "not implemented" in r.stderr
