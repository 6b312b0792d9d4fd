#!/usr/bin/env python
"""Golden at a BASELINE size: the REFERENCE's own HyperGraph.incidence_matrix(25) + clique_expansion_init on the first
4 M concatemers of workload C (n = 4 000 contigs; SURVEY.md 8d generator, seed 20242).

Run in the build container only (needs /root/reference):
    python tests/golden/make_golden_large.py

The CSR of A at this size is tens of MB, so the fixture stores digests instead of matrices:
  * sha256 of the canonical (sorted-index) CSR of H and of the INTEGER contact-count matrix
    (clique_expansion_init(..., use_high_order=False, min_weight=2.0)) -- bit-exact targets,
  * for the weighted float32 A (min_weight 2.0): sha256 of the sparsity pattern, float64 row sums, nnz, and a seeded
    sample of 200 000 stored entries (row, col, value) for the tolerance check.
The input is not stored either: tests regenerate the table from the seed (numpy and CUDA generators emit identical
bytes) and compare a sha256 of the hyperedge triplets before looking at any output.
"""
import hashlib
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

M_CONCATEMERS = 4_000_000
SAMPLE = 200_000


def sha(*arrays):
    h = hashlib.sha256()
    for a in arrays:
        a = np.ascontiguousarray(a)
        h.update(str(a.dtype).encode()); h.update(str(a.shape).encode()); h.update(a.tobytes())
    return h.hexdigest()


def canonical(M):
    M = M.tocsr().copy()
    M.sort_indices()
    return M


def digest_csr_int(M):
    """sha256 over (indptr i64, indices i32, data as i64) of a canonical CSR with integer-valued data."""
    M = canonical(M)
    d = np.asarray(M.data)
    di = np.rint(d).astype(np.int64)
    assert np.array_equal(di.astype(d.dtype), d), "data is not integer valued"
    return sha(M.indptr.astype(np.int64), M.indices.astype(np.int32), di)


def main():
    from make_golden import import_reference
    ah = import_reference()
    from cphasing_b200 import synth
    from oracle import hypergraph_oracle as orc
    cfg = synth.CONFIGS["C"]
    g = synth.genome_for("C")
    t0 = time.time()
    t = synth.porec_table(g, M_CONCATEMERS, cfg["seed"], law=cfg["law"])
    r, c, cnt, q = orc.porec_edges(t["read_idx"], t["contig_idx"], t["mapq"], min_mapq=1, min_order=2, max_order=50,
                                   n_contigs=g.n)
    print(f"table: {len(t['read_idx'])} records -> {len(r)} triplets ({time.time() - t0:.1f} s)")
    out = {"m": np.int64(M_CONCATEMERS), "n_records": np.int64(len(t["read_idx"])), "n_triplets": np.int64(len(r)),
           "triplets_sha256": sha(r.astype(np.int32), c.astype(np.int32), cnt.astype(np.uint32), q.astype(np.int8)),
           "params": np.array([1, 25, 2.0])}
    edges = ah.HyperEdges(idx=g.contig_idx, row=r.copy(), col=c.copy(), count=cnt.copy(), contigsizes=g.contigsizes,
                          mapq=q.copy())
    t0 = time.time()
    HG = ah.HyperGraph(edges, min_quality=1, mapq_filter=True)
    H = HG.incidence_matrix(min_contacts=25)
    t_h = time.time() - t0
    Hc = canonical(H)
    out.update(H_shape=np.array(H.shape, dtype=np.int64), H_nnz=np.int64(H.nnz), H_sha256=digest_csr_int(H),
               H_rowsum=np.asarray(Hc.sum(axis=1)).ravel().astype(np.int64),
               H_remove_contig_idx=np.asarray(HG.remove_contig_idx, dtype=np.int64))
    t0 = time.time()
    A = ah.HyperGraph.clique_expansion_init(H, min_weight=2.0)
    t_a = time.time() - t0
    Ac = canonical(A)
    rng = np.random.default_rng(4242)
    pick = np.sort(rng.choice(Ac.nnz, size=min(SAMPLE, Ac.nnz), replace=False))
    rows = np.searchsorted(Ac.indptr, pick, side="right") - 1
    out.update(A_nnz=np.int64(Ac.nnz), A_pattern_sha256=sha(Ac.indptr.astype(np.int64), Ac.indices.astype(np.int32)),
               A_rowsum=np.asarray(Ac.astype(np.float64).sum(axis=1)).ravel(),
               A_row_nnz=np.diff(Ac.indptr).astype(np.int64),
               A_sample_pos=pick.astype(np.int64), A_sample_row=rows.astype(np.int32),
               A_sample_col=Ac.indices[pick].astype(np.int32), A_sample_val=Ac.data[pick].astype(np.float32))
    Au = ah.HyperGraph.clique_expansion_init(H, min_weight=2.0, use_high_order=False)
    Auc = canonical(Au)
    out.update(Au_nnz=np.int64(Auc.nnz), Au_sha256=digest_csr_int(Au),
               Au_rowsum=np.asarray(Auc.astype(np.float64).sum(axis=1)).ravel().astype(np.int64))
    out["ref_seconds"] = np.array([t_h, t_a])
    np.savez_compressed(os.path.join(HERE, "ref_workloadC_4M.npz"), **out)
    print(f"H {H.shape} nnz={H.nnz} ({t_h:.1f} s)   A nnz={A.nnz} ({t_a:.1f} s)   Au nnz={Au.nnz}")
    print("size:", os.path.getsize(os.path.join(HERE, "ref_workloadC_4M.npz")) / 1e6, "MB")


if __name__ == "__main__":
    main()
