"""Parity at a BASELINE size: the first 4 M concatemers of workload C (n = 4 000) against a run of the REFERENCE's own
HyperGraph.incidence_matrix(25) + clique_expansion_init (tests/golden/make_golden_large.py -> ref_workloadC_4M.npz).
Integer work (H, the contact-count matrix) is compared by sha256 of the canonical CSR: bit-exact or fail.  The weighted
A: identical sparsity pattern, sampled values within 3e-6 of the reference's float32 and 1e-6 of the fp64 oracle."""
import hashlib
import os

import numpy as np
import pytest
from scipy.sparse import csr_matrix

from helpers import GOLDEN

M = 4_000_000


def sha(*arrays):
    h = hashlib.sha256()
    for a in arrays:
        a = np.ascontiguousarray(a)
        h.update(str(a.dtype).encode()); h.update(str(a.shape).encode()); h.update(a.tobytes())
    return h.hexdigest()


def digest_csr_int(M_):
    M_ = M_.tocsr().copy()
    M_.sort_indices()
    di = np.rint(np.asarray(M_.data)).astype(np.int64)
    return sha(M_.indptr.astype(np.int64), M_.indices.astype(np.int32), di)


@pytest.fixture(scope="module")
def gold():
    return dict(np.load(os.path.join(GOLDEN, "ref_workloadC_4M.npz")))


def check_A(gold, A, rtol, exact_pattern=False):
    """A: canonical csr (float64 values).  The reference computes A in float32 with sqrt-scaled operands
    (algorithms/hypergraph.py:418-446), so cells whose exact value EQUALS min_weight (1 + 1/3 + 1/3 + 1/3, ...) land on
    either side of the `< min_weight` cut by rounding alone ("entries within rounding of the cut = either").
    Everything else must agree: every sampled reference entry is either present with the same value or sits at the cut,
    and per row the difference of the row sums is explained by (difference of stored entries) x min_weight."""
    A = A.tocsr().copy(); A.sort_indices()
    pat = sha(A.indptr.astype(np.int64), A.indices.astype(np.int32))
    pos, rows, cols, vals = gold["A_sample_pos"], gold["A_sample_row"], gold["A_sample_col"], gold["A_sample_val"]
    rs = np.asarray(A.sum(axis=1)).ravel()
    if pat == str(gold["A_pattern_sha256"]):
        np.testing.assert_allclose(A.data[pos], vals.astype(np.float64), rtol=rtol)
        np.testing.assert_allclose(rs, gold["A_rowsum"], rtol=2e-5)
        return
    assert not exact_pattern, "sparsity pattern differs from the reference run"
    got = np.asarray(A[rows, cols]).ravel()
    miss = got == 0
    assert np.all(np.abs(vals[miss] - 2.0) <= 8e-6)                       # reference-only entries sit at the cut
    np.testing.assert_allclose(got[~miss], vals[~miss].astype(np.float64), rtol=rtol)
    dn = np.diff(A.indptr).astype(np.int64) - gold["A_row_nnz"]          # entries only one side stores, per row
    assert abs(int(dn.sum())) <= 0.06 * int(gold["A_nnz"])                # ties at the cut are a few per cent of A
    np.testing.assert_allclose(rs - 2.0 * dn, gold["A_rowsum"], rtol=2e-5)   # ... and each of them is worth min_weight


def test_oracle_matches_reference_at_4M(gold):
    """the CPU restatement against the reference run (pins the oracle at a BASELINE size)."""
    from cphasing_b200 import synth
    from oracle import hypergraph_oracle as orc
    cfg = synth.CONFIGS["C"]
    g = synth.genome_for("C")
    t = synth.porec_table(g, M, cfg["seed"], law=cfg["law"])
    assert len(t["read_idx"]) == int(gold["n_records"])
    r, c, cnt, q = orc.porec_edges(t["read_idx"], t["contig_idx"], t["mapq"], min_mapq=1, min_order=2, max_order=50,
                                   n_contigs=g.n)
    assert sha(r.astype(np.int32), c.astype(np.int32), cnt.astype(np.uint32), q.astype(np.int8)) == str(gold["triplets_sha256"])
    rr, cc, kk, ncol = orc.hypergraph_get_data(r, c, cnt, q, 1)
    H, _, _ = orc.incidence_matrix(rr, cc, kk, g.n, ncol, 25)
    assert H.shape == tuple(gold["H_shape"]) and digest_csr_int(H) == str(gold["H_sha256"])
    Au = orc.clique_expansion(H, use_high_order=False, min_weight=2.0)
    assert digest_csr_int(Au) == str(gold["Au_sha256"])
    A = orc.clique_expansion(H, min_weight=2.0, dtype=np.float32)
    check_A(gold, A, rtol=3e-6, exact_pattern=True)      # same float32 arithmetic -> same pattern


@pytest.mark.gpu
def test_cuda_path_matches_reference_at_4M(gold):
    """the CUDA path (device-generated table -> libcph) against the same reference run."""
    import torch
    from cphasing_b200 import device as D, synth
    cfg = synth.CONFIGS["C"]
    g = synth.genome_for("C")
    t = synth.porec_table_device(g, M, cfg["seed"], law=cfg["law"])
    assert int(t["read_idx"].numel()) == int(gold["n_records"])
    raw = D.porec_to_edges(t["read_idx"], t["contig_idx"], t["mapq"], g.n, min_mapq=1, min_order=2, max_order=50)
    assert raw.nnz == int(gold["n_triplets"])
    row, col, cnt, mq = D.edges_to_triplets(raw)
    assert sha(D.to_host(row), D.to_host(col), D.to_host(cnt).view(np.uint32), D.to_host(mq)) == str(gold["triplets_sha256"])
    H = D.incidence(raw, min_contacts=25, min_mapq=1)
    assert (H.n_rows, H.n_cols) == tuple(gold["H_shape"]) and H.nnz == int(gold["H_nnz"])
    Hs = csr_matrix(((D.to_host(H.pk).view(np.uint32) & 0xFFFFFF).astype(np.float32), D.to_host(H.indices), D.to_host(H.indptr)),
                    shape=(H.n_rows, H.n_cols))
    assert digest_csr_int(Hs) == str(gold["H_sha256"])
    # integer contact counts: bit-exact
    Au, acc = D.clique_expansion(H, use_high_order=False, min_weight=2.0, want32=False, want64=True)
    assert acc.frac_bits == 0
    Aus = csr_matrix((D.to_host(Au.data64), D.to_host(Au.indices), D.to_host(Au.indptr)), shape=(H.n_rows,) * 2)
    assert digest_csr_int(Aus) == str(gold["Au_sha256"])
    assert np.array_equal(np.asarray(Aus.sum(axis=1)).ravel().astype(np.int64), gold["Au_rowsum"])
    # weighted A: fp64 accumulation against the reference's float32
    A, _ = D.clique_expansion(H, use_high_order=True, min_weight=2.0, want32=True, want64=True)
    As = csr_matrix((D.to_host(A.data64), D.to_host(A.indices), D.to_host(A.indptr)), shape=(H.n_rows,) * 2)
    # against the reference's float32 values: its own accumulation error on long sums (the diagonal adds ~1e4 float32
    # terms) reaches a few 1e-5 relative; the 1e-6 bar of north_star is held against the fp64 restatement below
    check_A(gold, As, rtol=5e-5)
    from oracle import hypergraph_oracle as orc
    A64 = orc.clique_expansion(Hs.astype(np.float64), min_weight=2.0, dtype=np.float64).tocsr()
    A64.sort_indices(); As.sort_indices()
    assert np.array_equal(A64.indptr, As.indptr) and np.array_equal(A64.indices, As.indices)
    np.testing.assert_allclose(As.data, A64.data, rtol=1e-6)
    torch.cuda.synchronize()
