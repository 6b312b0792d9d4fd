"""The drop-in boundary, shown rather than asserted.

* ``test_dropin_matches_reference_driver_golden``: cphasing_b200.HyperPartition against groups produced by the REFERENCE's
  own driver (single_partition, incremental_partition / _incremental_partition) on its own HyperGraph / IRMM
  (tests/golden/make_golden_driver.py -> ref_driver.npz).
* ``test_reference_driver_on_dropin_classes``: the reference's ``cphasing.hyperpartition`` module (the unmodified install under
  baseline/_ref, which travels to the GPU box) with ``HyperGraph, IRMM, extract_incidence_matrix2`` rebound to
  cphasing_b200's -- the import switch of INTEGRATION.md -- runs ITS single_partition and incremental_partition; the groups
  must equal those of cphasing_b200.HyperPartition.
* ``test_merge_split_matches_reference`` (CPU): the split-contig merge against the reference's own method.
"""
import contextlib
import io
import os
import sys

import numpy as np
import pandas as pd
import pytest

from helpers import GOLDEN

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "baseline"))
sys.path.insert(0, GOLDEN)


def partition_of(hp, names):
    idx = {nm: i for i, nm in enumerate(names)}
    return {frozenset(idx[hp.vertices[v]] for v in grp) for grp in hp.K if len(grp)}


def partition_of_labels(lab):
    return {frozenset(np.where(lab == g)[0].tolist()) for g in np.unique(lab[lab >= 0])}


def reference_module():
    import ref_loader
    if not ref_loader.available():
        pytest.skip("the reference install (baseline/_ref) is not present")
    import make_golden_driver as mgd
    mgd.pandas_writable_shim()
    ref_loader.load()
    import cphasing.hyperpartition as rhp
    return rhp, mgd


@pytest.fixture(scope="module")
def setup(tmp_path_factory):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    import make_golden_driver as mgd
    from cphasing_b200 import hypergraph as hg, synth
    g = synth.make_genome(mgd.N_CONTIGS, mgd.CHROMS, mgd.PLOIDY, seed=mgd.SEED)
    t = synth.porec_table(g, mgd.M, seed=mgd.SEED + 1)
    ex = hg.HyperExtractor.from_arrays(t["read_idx"], t["contig_idx"], t["mapq"], g.contig_idx, g.contigsizes,
                                       min_quality=1, edge_length=0)
    g2, trip = mgd.inputs()
    pt = mgd.prune_table(g2, trip)
    wd = tmp_path_factory.mktemp("driver")
    ptf = str(wd / "prune.table")
    pt.to_csv(ptf, sep="\t", header=False, index=False)
    return g, ex.edges, pd.DataFrame({"length": pd.Series(g.contigsizes)}), ptf, wd, mgd


def run_dropin(setup, method, extra, use_pt):
    from cphasing_b200.hyperpartition import HyperPartition
    g, edges, sizes, ptf, wd, mgd = setup
    hp = HyperPartition(edges, sizes, k=[mgd.CHROMS, mgd.PLOIDY], prunetable=ptf if use_pt else None, **mgd.KW, **extra)
    if method == "single":
        hp.single_partition(k=mgd.CHROMS * mgd.PLOIDY)
    else:
        hp.incremental_partition([mgd.CHROMS, mgd.PLOIDY])
    return hp


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["single", "single_prune_leiden", "incremental", "incremental_prune"])
def test_dropin_matches_reference_driver_golden(setup, case):
    g, edges, sizes, ptf, wd, mgd = setup
    gold = np.load(os.path.join(GOLDEN, "ref_driver.npz"))
    method, extra, use_pt = mgd.CASES[case]
    hp = run_dropin(setup, method, extra, use_pt)
    assert partition_of(hp, g.names) == partition_of_labels(gold[f"{case}_labels"])
    if method == "incremental":
        # same chromosome -> haplotype structure: groups of one first-round cluster share their inc_chr_idx
        lab = gold[f"{case}_labels"]
        ref_chr = {frozenset(np.where(lab == gi)[0].tolist()): int(c) for gi, c in enumerate(gold[f"{case}_inc_chr_idx"])}
        idx = {nm: i for i, nm in enumerate(g.names)}
        mine = {frozenset(idx[hp.vertices[v]] for v in grp): c for grp, c in zip(hp.K, hp.inc_chr_idx)}
        same_chr = lambda d: {frozenset(k for k in d if d[k] == c) for c in set(d.values())}
        assert same_chr(mine) == same_chr(ref_chr)


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["single_prune_leiden", "incremental_prune", "incremental"])
def test_reference_driver_on_dropin_classes(setup, case):
    rhp, _ = reference_module()
    from cphasing_b200.algorithms import hypergraph as bah
    g, edges, sizes, ptf, wd, mgd = setup
    method, extra, use_pt = mgd.CASES[case]
    saved = (rhp.HyperGraph, rhp.IRMM, rhp.extract_incidence_matrix2)
    rhp.HyperGraph, rhp.IRMM, rhp.extract_incidence_matrix2 = bah.HyperGraph, bah.IRMM, bah.extract_incidence_matrix2
    cwd = os.getcwd()
    os.chdir(wd)
    try:
        with contextlib.redirect_stdout(io.StringIO()):
            hp = rhp.HyperPartition(edges, sizes, k=[mgd.CHROMS, mgd.PLOIDY], prunetable=ptf if use_pt else None,
                                    ultra_complex=0, **mgd.KW, **extra)
            if method == "single":
                hp.single_partition(k=mgd.CHROMS * mgd.PLOIDY)
            else:
                hp.incremental_partition([mgd.CHROMS, mgd.PLOIDY])
    finally:
        os.chdir(cwd)
        rhp.HyperGraph, rhp.IRMM, rhp.extract_incidence_matrix2 = saved
    mine = run_dropin(setup, method, extra, use_pt)
    assert partition_of(hp, g.names) == partition_of(mine, g.names)
    assert len(hp.K) == mgd.CHROMS * mgd.PLOIDY


def test_merge_split_matches_reference():
    rhp, _ = reference_module()
    from cphasing_b200.hyperpartition import HyperPartition as HP
    rng = np.random.default_rng(0)
    names, sizes = [], {}
    for c in range(60):
        L = int(rng.integers(100, 1500))
        for a in range(0, L, 250):
            b = min(L, a + 250)
            names.append(f"ctg{c}|{a}_{b}")
            sizes[names[-1]] = b - a
    verts = np.array(names)
    perm = rng.permutation(len(verts))[:int(0.8 * len(verts))]
    K = [sorted(perm[i::6].tolist()) for i in range(6)] + [[]]

    def mk(cls):
        o = cls.__new__(cls)
        o.vertices, o.K, o.contig_sizes, o.split = verts, [list(x) for x in K], dict(sizes), True
        o.contigsizes = pd.DataFrame({"length": pd.Series(sizes)})
        o._cached_raw_contigsizes = None
        o.inc_chr_idx = [0, 0, 1, 1, 2, 2, 3]
        return o
    a, b = mk(rhp.HyperPartition), mk(HP)
    for secondary in (False, True):
        ra, rb = a.merge_split(secondary=secondary), b.merge_split(secondary=secondary)
        assert ra[0] == rb[0] and ra[1] == rb[1] and ra[2] == rb[2]
        assert list(a.inc_chr_idx) == list(b.inc_chr_idx)


@pytest.mark.gpu
def test_rescue_static_matches_reference(setup):
    """hyperpartition.py:3426-3575 on the same H: the reference's own static method (scipy H and A) against the drop-in's
    (device A = H W H^T, S = A M from cph_group_sums) -- same groups after the rescue, same number of rescued contigs,
    with an allelic constraint table in play."""
    rhp, mgd = reference_module()
    import cphasing.algorithms.hypergraph as rah
    from scipy.sparse import csr_matrix
    from cphasing_b200.algorithms import hypergraph as ah
    from cphasing_b200.hyperpartition import HyperPartition
    g, edges, sizes, ptf, wd, _ = setup
    HG = ah.HyperGraph(edges, min_quality=1)
    H = HG.incidence_matrix(min_contacts=5)
    n = H.shape[0]
    rng = np.random.default_rng(3)
    truth = np.array([g.group[g.contig_idx[c]] for c in HG.nodes])
    # groups = the planted ones with a third of the contigs left unanchored
    loose = rng.random(n) < 0.33
    K0 = [np.where((truth == t) & ~loose)[0].tolist() for t in np.unique(truth)]
    K0 = [k for k in K0 if k]
    un = np.where(loose)[0]
    # constraints: some unanchored contigs may not join the group of an anchored "allele"
    c1 = un[: len(un) // 4]
    c2 = np.array([K0[int(truth[c]) % len(K0)][0] for c in c1])
    ppd = pd.DataFrame({"contig1": c1, "contig2": c2, "mz1": 0, "mz2": 0, "mzShared": 0, "similarity": 0.995, "type": 0})
    sizes_idx = {i: float(g.contigsizes[c]) for i, c in enumerate(HG.nodes)}
    Hs = csr_matrix(H)
    with contextlib.redirect_stderr(io.StringIO()):
        Kr, nr = rhp.HyperPartition._rescue_static([list(k) for k in K0], Hs, dict(sizes_idx), ppd, min_weight=1.0)
    Kd, nd = HyperPartition._rescue_static([list(k) for k in K0], H, dict(sizes_idx), ppd, min_weight=1.0)
    assert nr == nd and nr > 20
    assert [list(map(int, k)) for k in Kr] == [list(map(int, k)) for k in Kd]
    assert any(set(map(int, c1)) - set(x for k in Kd for x in k))        # the constraints did hold something back


@pytest.mark.gpu
@pytest.mark.parametrize("use_mode", ["merge", "conflict"])
def test_merge_with_allele_table_matches_reference(setup, use_mode):
    """_merge(alleletable_df=...) (hyperpartition.py:2288-2471): group overlaps weighed in shared minimizers, the form the
    first round takes with the CLI defaults and an allele table (:1987-1995) -- the reference's own static method on the
    scipy A against the drop-in's (V = M^T A M from the device, rank-1 updates)."""
    rhp, mgd = reference_module()
    from scipy.sparse import csr_matrix
    from cphasing_b200 import synth
    from cphasing_b200.algorithms import hypergraph as ah
    from cphasing_b200.hyperpartition import HyperPartition
    g, edges, sizes, ptf, wd, _ = setup
    HG = ah.HyperGraph(edges, min_quality=1)
    H = HG.incidence_matrix(min_contacts=5)
    A = ah.HyperGraph.clique_expansion_init(H, min_weight=1.0)
    n = A.shape[0]
    name_idx = {c: i for i, c in enumerate(HG.nodes)}
    truth = np.array([g.group[g.contig_idx[c]] for c in HG.nodes])
    rng = np.random.default_rng(7)
    K0 = []
    for t in np.unique(truth):                                  # every planted group in two random pieces
        m = np.where(truth == t)[0]
        cut = rng.random(len(m)) < 0.5
        K0 += [m[cut].tolist(), m[~cut].tolist()]
    K0 = [k for k in K0 if k]
    rows = [(name_idx[g.names[a]], name_idx[g.names[b]], m1, sh) for a, b, m1, m2, sh, sim in synth.allele_pairs(g, seed=3)
            if g.names[a] in name_idx and g.names[b] in name_idx]
    at_df = pd.DataFrame(rows, columns=[1, 2, "mz1", "mzShared"])
    aset = set(map(tuple, at_df[[1, 2]].values))
    vsz = pd.DataFrame({"length": {i: float(g.contigsizes[c]) for i, c in enumerate(HG.nodes)}})
    k = len(np.unique(truth))
    with contextlib.redirect_stderr(io.StringIO()):
        Kr = rhp.HyperPartition._merge(csr_matrix(A), [list(x) for x in K0], vsz, k, allelic_idx_set=aset,
                                       alleletable_df=at_df.copy(), min_allelic_overlap=0.8 if use_mode == "merge" else 0.1,
                                       allele_use_method=use_mode, method="sum")
    Kd = HyperPartition._merge(A, [list(x) for x in K0], vsz, k, allelic_idx_set=aset, alleletable_df=at_df.copy(),
                               min_allelic_overlap=0.8 if use_mode == "merge" else 0.1, allele_use_method=use_mode,
                               method="sum")
    assert len(Kr) == len(Kd) == k
    assert {frozenset(map(int, x)) for x in Kr} == {frozenset(map(int, x)) for x in Kd}
