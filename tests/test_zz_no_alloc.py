"""cph_incidence_csr_fixed -- the CSR view of an incidence matrix for callers that cannot bring a workspace.

Who needs it: code that runs as one stage inside another program (a workflow engine's operator callback, a plug-in) and
receives its input and output buffers but no scratch, on a GPU whose free memory the host program's pool has already
claimed; and services that keep several assemblies in flight on one GPU and must not call cudaMalloc / cudaFree while
serving, because those calls stall every stream of the device.  For them the temporaries live in a 16 MiB array inside
libcph.so, and the entry point takes inputs and outputs only (at most CPH_INCIDENCE_CSR_FIXED_MAX_NNZ entries).

The checks below go through the C-ABI (ctypes) and compare against cph_incidence_csr, the entry point with a caller-owned
workspace, on the same inputs.  The outputs are integers, so "same" means bit-identical -- the bar of every other
integer test in this suite.  Two calls are made back to back, the second smaller than the first and on a different input
form, so that it finds the first call's leftovers in the shared array; then the limit is probed.

This file sorts last on purpose: `pytest -x` stops at the first failure, and a failure here must not hide the parity tests.
It passes alone in a fresh process (the first call does the per-device set-up) and after the rest of the suite.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_NNZ = 524288          # CPH_INCIDENCE_CSR_FIXED_MAX_NNZ of include/cph.h
HEAD = 0x80000000


def make_pem(n_edges, n_rows, seed, nnz_exact=None):
    """a packed edge-major incidence (include/cph.h, 'Device formats'): hyperedges of order 2..6, head bit on the first entry."""
    rng = np.random.default_rng(seed)
    order = rng.integers(2, 7, n_edges)
    if nnz_exact is not None:
        cs = np.cumsum(order)
        k = int(np.searchsorted(cs, nnz_exact - 1))          # first k edges hold < nnz_exact entries; edge k takes the rest
        order = order[:k + 1].copy()
        order[k] = nnz_exact - (int(cs[k - 1]) if k else 0)
        assert order.sum() == nnz_exact and order[k] >= 1
    nnz = int(order.sum())
    starts = np.concatenate(([0], np.cumsum(order)[:-1]))
    row = rng.integers(0, n_rows, nnz).astype(np.uint32)
    row[starts] |= HEAD
    d_e = np.repeat(np.maximum(order - 1, 1), order).astype(np.uint32)
    pk = (d_e << 24) | rng.integers(1, 40, nnz).astype(np.uint32)
    col = np.repeat(rng.permutation(n_edges)[:len(order)], order).astype(np.uint32)   # explicit, unordered hyperedge ids
    return row, pk, col


def run_both(torch, L, dev, row, pk, col, n_rows):
    """(indptr, indices, pk) from cph_incidence_csr with a torch workspace and from cph_incidence_csr_fixed without one."""
    from cphasing_b200._lib import check, stream
    nnz = len(row)
    d_row = torch.from_numpy(row.view(np.int32).copy()).to(dev)
    d_pk = torch.from_numpy(pk.view(np.int32).copy()).to(dev)
    d_col = None if col is None else torch.from_numpy(col.view(np.int32).copy()).to(dev)
    p = lambda t: None if t is None else t.data_ptr()

    def outputs():
        # filled with a pattern, so an entry point that wrote nothing cannot pass by luck
        return (torch.full((n_rows + 1,), -7, dtype=torch.int64, device=dev),
                torch.full((max(nnz, 1),), -7, dtype=torch.int32, device=dev),
                torch.full((max(nnz, 1),), -7, dtype=torch.int32, device=dev))

    ref = outputs()
    need = C.c_size_t(0)
    check(L.cph_incidence_csr(p(d_row), p(d_pk), p(d_col), nnz, n_rows, None, None, None, None, 0, C.byref(need), None))
    ws = torch.empty(need.value, dtype=torch.uint8, device=dev)
    check(L.cph_incidence_csr(p(d_row), p(d_pk), p(d_col), nnz, n_rows, p(ref[0]), p(ref[1]), p(ref[2]),
                              ws.data_ptr(), ws.numel(), C.byref(need), stream()))
    got = outputs()
    before = torch.cuda.memory_allocated(dev)
    check(L.cph_incidence_csr_fixed(p(d_row), p(d_pk), p(d_col), nnz, n_rows, p(got[0]), p(got[1]), p(got[2]), stream()))
    assert torch.cuda.memory_allocated(dev) == before          # nothing went through torch's allocator either
    torch.cuda.synchronize(dev)
    return [t.cpu().numpy() for t in ref], [t.cpu().numpy() for t in got], nnz


def assert_identical(ref, got, nnz, n_rows):
    assert np.array_equal(ref[0], got[0]), "indptr differs"
    assert np.array_equal(ref[1][:nnz], got[1][:nnz]), "indices differ"
    assert np.array_equal(ref[2][:nnz], got[2][:nnz]), "packed counts differ"
    assert got[0][0] == 0 and got[0][n_rows] == nnz and np.all(np.diff(got[0]) >= 0)


def test_fixed_scratch_entry_point_matches_workspace_entry_point_twice_in_a_row():
    import torch
    from cphasing_b200 import _lib
    L = _lib.lib()
    dev = torch.device("cuda", torch.cuda.current_device())

    # call 1: a large case, hyperedge ids implied by the head bits (pem_col == NULL): uses the head-flag scan as well
    row, pk, _ = make_pem(100_000, 3000, seed=11)
    ref, got, nnz = run_both(torch, L, dev, row, pk, None, 3000)
    assert 300_000 < nnz <= MAX_NNZ
    assert_identical(ref, got, nnz, 3000)

    # call 2, straight after: smaller, explicit hyperedge ids, other row count -- the array still holds call 1's bytes
    row, pk, col = make_pem(9_000, 517, seed=12)
    ref, got, nnz = run_both(torch, L, dev, row, pk, col, 517)
    assert_identical(ref, got, nnz, 517)
    # the CSR really is the matrix: same multiset of (row, hyperedge, packed count) as the input
    rows_out = np.repeat(np.arange(517), np.diff(got[0]))
    a = np.stack([rows_out, got[1][:nnz].astype(np.int64), got[2][:nnz].view(np.uint32).astype(np.int64)])
    b = np.stack([(row & 0x7FFFFFFF).astype(np.int64), col.astype(np.int64), pk.astype(np.int64)])
    assert np.array_equal(a[:, np.lexsort(a[::-1])], b[:, np.lexsort(b[::-1])])

    # call 3: the same input as call 2 again gives the same answer (nothing of call 2 leaked into the result)
    ref3, got3, _ = run_both(torch, L, dev, row, pk, col, 517)
    assert_identical(ref3, got3, nnz, 517)
    assert all(np.array_equal(x, y) for x, y in zip(got, got3))


def test_fixed_scratch_limit():
    import torch
    from cphasing_b200 import _lib
    L = _lib.lib()
    dev = torch.device("cuda", torch.cuda.current_device())

    # exactly the documented maximum is accepted and right
    row, pk, _ = make_pem(200_000, 65_000, seed=13, nnz_exact=MAX_NNZ)
    ref, got, nnz = run_both(torch, L, dev, row, pk, None, 65_000)
    assert nnz == MAX_NNZ
    assert_identical(ref, got, nnz, 65_000)

    # one entry more is refused before anything is launched, with a message that names the limit; outputs untouched
    n = MAX_NNZ + 1
    z = torch.zeros(n, dtype=torch.int32, device=dev)
    indptr = torch.full((11,), -7, dtype=torch.int64, device=dev)
    launches = _lib.launches()
    rc = L.cph_incidence_csr_fixed(z.data_ptr(), z.data_ptr(), None, n, 10, indptr.data_ptr(), z.data_ptr(), z.data_ptr(),
                                   _lib.stream())
    assert rc == _lib.ERR_TOO_LARGE
    msg = L.cph_last_error().decode()
    assert str(MAX_NNZ) in msg and str(n) in msg and "cph_incidence_csr" in msg
    assert _lib.launches() == launches
    torch.cuda.synchronize(dev)
    assert bool((indptr == -7).all())
