"""diagnostic: where the host blocks inside one e2e step (no extra syncs) -- run on the GPU box."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from bench import PARAMS
from cphasing_b200 import synth
from cphasing_b200 import hypergraph as hg
from cphasing_b200.algorithms import hypergraph as ah

cfg = dict(synth.CONFIGS["C"])
dev = torch.device("cuda:0")
g, t = bench.gen_table(cfg, 0, dev)
host = {k: torch.empty(v.shape, dtype=v.dtype, pin_memory=True).copy_(v) for k, v in t.items()}
del t
idx, sizes = g.contig_idx, g.contigsizes
keep = []
for i in range(14):
    torch.cuda.synchronize()
    T = [time.perf_counter()]
    ex = hg.HyperExtractor.from_arrays(host["read_idx"], host["contig_idx"], host["mapq"], idx, sizes, min_quality=1,
                                       min_order=2, max_order=50, edge_length=0)
    T.append(time.perf_counter())
    HG = ah.HyperGraph(ex.edges, min_quality=1, mapq_filter=False)
    T.append(time.perf_counter())
    H = HG.incidence_matrix(25)
    T.append(time.perf_counter())
    A = ah.HyperGraph.clique_expansion_init(H, min_weight=2.0)
    T.append(time.perf_counter())
    nb = A.data.nbytes + A.indices.nbytes + A.indptr.nbytes
    T.append(time.perf_counter())
    torch.cuda.synchronize()
    T.append(time.perf_counter())
    if len(sys.argv) > 1:
        keep = [ex, HG, H, A]       # previous step's objects die at the NEXT assignment (as in bench.py)
    print("step %2d total %6.1f | extract %5.1f HG %5.1f incidence %5.1f expand %5.1f csr %5.1f sync %5.1f" %
          ((i, (T[-1] - T[0]) * 1e3) + tuple((b - a) * 1e3 for a, b in zip(T[:-1], T[1:]))))
