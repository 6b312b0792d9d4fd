"""DevicePairs (cphasing_b200/algorithms/hypergraph.py): the pair lists of kprune kept as device vectors behind the
two-element list of pandas Series the reference's callers index (hyperpartition.py:438-461).  CPU tensors stand in for
device tensors here -- the container logic is the same."""
import pickle

import numpy as np
import pandas as pd
import torch

from cphasing_b200.algorithms.hypergraph import DevicePairs, HyperGraph


def make():
    r = torch.tensor([0, 5, 2, 7, 7], dtype=torch.int64)
    c = torch.tensor([3, 1, 2, 9, 8], dtype=torch.int64)
    return DevicePairs(r, c), r.numpy(), c.numpy()


def test_behaves_like_the_list_of_two_series():
    P, r, c = make()
    assert len(P) == 2 and bool(P)
    assert isinstance(P[0], pd.Series) and isinstance(P[1], pd.Series)
    assert np.array_equal(P[0].to_numpy(), r) and np.array_equal(P[-1].to_numpy(), c)
    a, b = P
    assert a is P[0] and b is P[1]                      # built once
    rr, cc = HyperGraph._normalize_pair_index(P)        # what host consumers of the reference call
    assert np.array_equal(rr, r) and np.array_equal(cc, c)
    assert isinstance(P, (list, tuple)) and len(P[0:2]) == 2
    Q = pickle.loads(pickle.dumps(P))                   # crosses process boundaries as a plain list of Series
    assert type(Q) is list and np.array_equal(Q[1].to_numpy(), c)


def test_upper_symmetric_matches_the_host_construction():
    from cphasing_b200.algorithms import hypergraph as ah
    P, r, c = make()
    up = P.upper_symmetric()
    want = ah._upper_pairs_symmetric([pd.Series(r), pd.Series(c)])
    assert np.array_equal(up[0].numpy(), want[0]) and np.array_equal(up[1].numpy(), want[1])
    empty = DevicePairs(torch.tensor([4, 4], dtype=torch.int64), torch.tensor([4, 1], dtype=torch.int64))
    assert empty.upper_symmetric() is None              # nothing listed with r < c
