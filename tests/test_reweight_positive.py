"""HyperGraph.reweight_adjacency_matrix with a POSITIVE numeric factor (algorithms/hypergraph.py:674-681): the branch the CLI
default `--allelic-positive-factor 3.0` takes in the first round of `incremental_partition` (hyperpartition.py:1903-1907).
Host code on both sides: compared with the reference's own method (baseline/_ref or /root/reference)."""
import os
import sys
import warnings

import numpy as np
import pytest
from scipy.sparse import csr_matrix, random as sprand

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "baseline"))


@pytest.mark.parametrize("factor", [3.0, 1.5])
def test_positive_factor_matches_reference(factor):
    import ref_loader
    if not ref_loader.available():
        pytest.skip("the reference install (baseline/_ref) is not present")
    rah = ref_loader.load()
    from cphasing_b200.algorithms.hypergraph import HyperGraph
    rng = np.random.default_rng(0)
    M = sprand(80, 80, 0.15, random_state=1, dtype=np.float32)
    A = csr_matrix(M + M.T)
    r = rng.integers(0, 80, 60)
    c = rng.integers(0, 80, 60)
    r[-1], c[-1] = r[0], c[0]                                   # a pair listed twice; several pairs have no stored contact
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")                         # scipy's SparseEfficiencyWarning inside the reference
        ref = rah.HyperGraph.reweight_adjacency_matrix(A.copy(), [list(r), list(c)], None, allelic_factor=factor)
    got = HyperGraph.reweight_adjacency_matrix(A.copy(), [list(r), list(c)], None, allelic_factor=factor)
    assert ref.nnz == got.nnz
    d = abs(csr_matrix(ref) - csr_matrix(got))
    assert (d.max() if d.nnz else 0.0) <= 2e-6 * max(abs(ref).max(), 1.0)
    assert abs(A - got).nnz > 0 and got.nnz > A.nnz             # values changed and entries were created
