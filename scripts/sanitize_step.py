#!/usr/bin/env python
"""One small pass over every kernel family of the hot path, for compute-sanitizer (memcheck / racecheck / initcheck):

    compute-sanitizer --tool racecheck python scripts/sanitize_step.py

porec group-by (tile kernel + the unsorted sort path), pairs, incidence, direct / swept / binned expansion, emit, kprune,
Louvain + Leiden, group sums, IRMM rounds.  Results are checked against the oracle so that a "clean" run is also a correct one."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cphasing_b200 import community, device as D, graphops, kprune, synth    # noqa: E402
from oracle import hypergraph_oracle as orc                                   # noqa: E402  (checker only)


def main():
    g = synth.make_genome(240, 4, 3, seed=3)
    t = synth.porec_table(g, 25_000, seed=4)
    raw = D.porec_to_edges(t["read_idx"], t["contig_idx"], t["mapq"], g.n, min_mapq=1)
    perm = np.random.default_rng(0).permutation(len(t["read_idx"]))
    raw_u = D.porec_to_edges(t["read_idx"][perm], t["contig_idx"][perm], t["mapq"][perm], g.n, min_mapq=1)   # sort path
    assert raw_u.nnz == raw.nnz
    r, c, k, q = orc.porec_edges(t["read_idx"], t["contig_idx"], t["mapq"], min_mapq=1, n_contigs=g.n)
    row, col, cnt, mq = D.edges_to_triplets(raw)
    assert np.array_equal(D.to_host(row), r) and np.array_equal(D.to_host(cnt), k)
    H = D.incidence(raw, min_contacts=5, min_mapq=0)
    ref = None
    for form in ("direct", "swept", "binned"):
        os.environ["CPH_EXPAND_FORM"] = form
        os.environ["CPH_BAND_MB"] = "0.05"
        old = D.BINNED_MIN_N, D.BINNED_MIN_NNZ
        D.BINNED_MIN_N, D.BINNED_MIN_NNZ = (0, 0) if form != "direct" else old
        for ho in (True, False):
            acc = D.expand_accumulate(H, use_high_order=ho)
            if form == "direct":
                ref = ref or {}
                ref[ho] = acc
            else:
                assert torch.equal(acc.acc, ref[ho].acc) and torch.equal(acc.diag, ref[ho].diag), form
        D.BINNED_MIN_N, D.BINNED_MIN_NNZ = old
    os.environ.pop("CPH_EXPAND_FORM")
    A = D.expand_emit(ref[True], min_weight=1.0)
    Au = D.expand_emit(ref[False], min_weight=1.0, want32=False)
    pairs = np.array([(a, b) for a, b, *_ in synth.allele_pairs(g, seed=5)], dtype=np.int64)
    keep = H.row_keep.cpu().numpy()
    new = np.cumsum(keep) - 1
    pairs = np.array([(new[a], new[b]) for a, b in pairs if keep[a] and keep[b]], dtype=np.int64).reshape(-1, 2)
    flag, row_of = kprune.kprune_device(Au, pairs, 1)
    a2 = graphops.clone_adjacency(A)
    for method in ("louvain", "leiden"):
        mem, qv, st = community.community_device(a2, 1.0, method)
        assert st["groups"] >= 2
    tp = synth.pairs_table(g, 20_000, seed=6)
    rp = D.pairs_to_edges(tp["chrom1"], tp["chrom2"], tp["mapq"], g.n, min_mapq=1)
    Hp = D.incidence(rp, min_contacts=1, min_mapq=0)
    D.clique_expansion(Hp, use_high_order=False, min_weight=1.0)
    torch.cuda.synchronize()
    print("sanitize_step ok: nnz(H)=%d nnz(A)=%d allelic=%d weak=%d" % (H.nnz, A.nnz, int((flag == 1).sum()), int((flag == 2).sum())))


if __name__ == "__main__":
    main()
