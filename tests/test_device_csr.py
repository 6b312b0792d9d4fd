"""DeviceBackedCSR (cphasing_b200/algorithms/hypergraph.py): the matrices handed back to the reference's callers must behave
as scipy.sparse.csr_matrix for every operation those callers use (H[idx], .power(2), .sum(axis), A[r][:, r],
.dot(M), .diagonal(), .copy(), .tocoo(), pickle), and an in-place edit must never leave a stale device twin behind."""
import pickle

import numpy as np
import pytest
from scipy.sparse import csr_matrix, triu

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mats():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from cphasing_b200 import hypergraph as hg, synth
    from cphasing_b200.algorithms import hypergraph as ah
    g = synth.make_genome(150, 5, 2, seed=3)
    t = synth.porec_table(g, 30000, seed=4)
    ex = hg.HyperExtractor.from_arrays(t["read_idx"], t["contig_idx"], t["mapq"], g.contig_idx, g.contigsizes,
                                       min_quality=1, edge_length=0)
    HG = ah.HyperGraph(ex.edges, min_quality=1)
    H = HG.incidence_matrix(min_contacts=5)
    A = ah.HyperGraph.clique_expansion_init(H, min_weight=1.0)
    return ah, H, A


def plain(M):
    return csr_matrix((np.array(M.data), np.array(M.indices), np.array(M.indptr)), shape=M.shape)


def test_scipy_operations_of_the_reference_callers(mats):
    ah, H, A = mats
    Hp, Ap = plain(H), plain(A)
    assert isinstance(H, csr_matrix) and isinstance(A, csr_matrix)
    idx = np.array([3, 1, 7, 20, 5])
    eq = lambda X, Y: (abs(csr_matrix(X) - csr_matrix(Y)).nnz == 0) and X.shape == Y.shape
    assert eq(H[idx], Hp[idx])                                           # hyperpartition.py:1270 / extract_incidence_matrix2
    assert np.array_equal(np.asarray(H.power(2).sum(axis=1)), np.asarray(Hp.power(2).sum(axis=1)))   # :1276
    assert np.array_equal(np.asarray(H.sum(axis=0)), np.asarray(Hp.sum(axis=0)))
    assert np.array_equal(np.asarray(H.sum(axis=1)), np.asarray(Hp.sum(axis=1)))
    assert H.getnnz() == Hp.getnnz() and np.array_equal(H.getnnz(axis=0), Hp.getnnz(axis=0))
    r = np.arange(0, A.shape[0], 3)
    assert eq(A[r][:, r], Ap[r][:, r])                                   # hyperpartition.py:1765
    assert eq(A[r, :][:, r], Ap[r, :][:, r])
    assert np.array_equal(A.diagonal(), Ap.diagonal())                   # :991
    M = np.random.default_rng(0).random((A.shape[0], 4))
    assert np.allclose(A.dot(M), Ap.dot(M)) and np.allclose(A @ M, Ap @ M)   # :2402-2410 (V = M^T A M)
    assert np.allclose(M.T @ (A @ M), M.T @ (Ap @ M))
    assert eq(A.T, Ap.T) and eq(triu(A, k=1), triu(Ap, k=1)) and eq(abs(A), abs(Ap))
    c, cp = A.tocoo(), Ap.tocoo()
    assert np.array_equal(c.row, cp.row) and np.array_equal(c.col, cp.col) and np.array_equal(c.data, cp.data)
    assert eq(A.copy(), Ap) and eq(A.astype(np.float64), Ap.astype(np.float64)) and eq(A * 2.0, Ap * 2.0)
    assert eq(A + A.T, Ap + Ap.T) and eq(A.multiply(A), Ap.multiply(Ap))
    assert A.max() == Ap.max() and A.nnz == Ap.nnz and A.dtype == Ap.dtype
    B = pickle.loads(pickle.dumps(A))                                    # joblib workers receive H pickled (:2166-2181)
    assert type(B) is csr_matrix and eq(B, Ap)
    # none of the reads above may have cut the matrix loose from its device twin
    assert ah.device_handle(A) is not None and ah.device_handle(H) is not None


def test_in_place_edits_cannot_leave_a_stale_device_twin(mats):
    ah, H, A0 = mats
    A = ah.HyperGraph.clique_expansion_init(H, min_weight=1.0)
    assert ah.device_handle(A) is not None
    with pytest.raises(ValueError):                                      # numpy-level edit of the attached host copy: loud
        A.data[A.data < 5] = 0
    with pytest.raises(ValueError):
        A.data *= 2.0
    # scipy's own in-place methods detach first, then work as usual
    B = ah.HyperGraph.clique_expansion_init(H, min_weight=1.0)
    ref = plain(B)
    B.setdiag(0)
    ref.setdiag(0)
    assert ah.device_handle(B) is None
    assert abs(csr_matrix(B) - ref).nnz == 0
    B.data[B.data < 5] = 0                                               # detached: writable
    B.eliminate_zeros()
    ref.data[ref.data < 5] = 0; ref.eliminate_zeros()
    assert abs(csr_matrix(B) - ref).nnz == 0
    # the detached, edited matrix is what the next device stage sees (uploaded afresh), not the old twin
    from cphasing_b200 import graphops
    up = graphops.upload_adjacency(B)
    assert up.nnz == ref.nnz
    # a copy is a plain writable matrix and leaves the original attached; filter_by_weight no longer corrupts its input
    C = ah.HyperGraph.clique_expansion_init(H, min_weight=1.0)
    before = plain(C)
    F = ah.HyperGraph.filter_by_weight(C, min_weight=10.0)
    assert F.nnz < before.nnz and abs(csr_matrix(C) - before).nnz == 0 and ah.device_handle(C) is not None
    D2 = C.copy()
    D2.data[:] = 0
    assert abs(csr_matrix(C) - before).nnz == 0
    # item assignment and *= detach too
    E = ah.HyperGraph.clique_expansion_init(H, min_weight=1.0)
    E *= 3.0
    assert ah.device_handle(E) is None and np.allclose(E.data, before.data * 3.0)
    # IRMM on an edited matrix uses the edited values: zeroing every off-diagonal entry leaves nothing to cluster on
    G = ah.HyperGraph.clique_expansion_init(H, min_weight=1.0)
    G.setdiag(0)
    G.data[:] = 0
    G.eliminate_zeros()
    assert graphops.upload_adjacency(G).nnz == 0
