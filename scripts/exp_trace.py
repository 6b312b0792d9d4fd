"""trace of the adaptive phase control on one noisy planted graph (CPH_COMMUNITY_TRACE=1)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from scipy.sparse import csr_matrix
def planted(n, g, p_in, p_out, seed, wmax=5.0):
    rng = np.random.default_rng(seed)
    truth = np.arange(n) % g
    P = np.where(truth[:, None] == truth[None, :], p_in, p_out)
    U = np.triu(rng.random((n, n)) < P, 1)
    W = np.triu(rng.uniform(0.5, wmax, size=(n, n)), 1) * U
    return csr_matrix(W + W.T), truth
from oracle import community_oracle as co
from cphasing_b200 import community
case = tuple(float(x) if "." in x else int(x) for x in sys.argv[1].split(","))
A, truth = planted(*case)
mo, _ = co.community(A, 1.0, "louvain")
print("oracle Q", co.modularity(A, mo, 1.0))
os.environ["CPH_COMMUNITY_TRACE"] = "1"
vc = community.cluster(A, 1.0, "louvain")
print("gpu Q", vc.quality, vc.stats)
