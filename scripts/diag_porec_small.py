#!/usr/bin/env python
"""porec_to_edges (device-resident columns) against the oracle for one small table; argv: n_concatemers seed."""
import os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cphasing_b200 import device as D, synth
from oracle import hypergraph_oracle as orc
m, seed = int(sys.argv[1]), int(sys.argv[2])
g = synth.make_genome(150, 5, 2, seed=5)
t = synth.porec_table(g, m, seed=seed, law="poisson")
exp = orc.porec_edges(t["read_idx"], t["contig_idx"], t["mapq"], min_mapq=1, min_order=2, max_order=50, n_contigs=g.n)
if len(sys.argv) > 3:      # unaligned device columns -> plain loads instead of bulk copies
    dev = {k: torch.from_numpy(np.concatenate([t[k][:1], t[k]])).cuda()[1:] for k in ("read_idx", "contig_idx", "mapq")}
    raw = D.porec_to_edges(dev["read_idx"], dev["contig_idx"], dev["mapq"], g.n, min_mapq=1, min_order=2, max_order=50)
else:
    raw = D.porec_to_edges(t["read_idx"], t["contig_idx"], t["mapq"], g.n, min_mapq=1, min_order=2, max_order=50)
torch.cuda.synchronize()
row, col, cnt, mq = D.edges_to_triplets(raw)
got = (row.cpu().numpy(), col.cpu().numpy(), cnt.view(torch.int32).cpu().numpy().view(np.uint32), mq.cpu().numpy())
print(m, "records", len(t["read_idx"]), "nnz", raw.nnz, "exp", len(exp[0]), "equal", all(np.array_equal(a, b) for a, b in zip(got, exp)))
if not all(np.array_equal(a, b) for a, b in zip(got, exp)):
    r = t["read_idx"]
    miss = np.setdiff1d(np.unique(exp[1]), np.unique(got[1]))
    extra = np.setdiff1d(np.unique(got[1]), np.unique(exp[1]))
    print("missing reads", len(miss), "extra reads", len(extra))
    for rd in list(miss[:6]) + list(extra[:3]):
        pos = np.flatnonzero(r == rd)
        print(" read", rd, "records", pos[0], "..", pos[-1], "tile", pos[0] // 1792, "offset", pos[0] % 1792, "len", len(pos),
              "contigs", t["contig_idx"][pos].tolist(), "mapq", t["mapq"][pos].tolist())
    # entries that differ among common reads
    import collections
    ge = collections.Counter(zip(got[1].tolist(), got[0].tolist(), got[2].tolist(), got[3].tolist()))
    ee = collections.Counter(zip(exp[1].tolist(), exp[0].tolist(), exp[2].tolist(), exp[3].tolist()))
    d1 = list((ee - ge).items())[:5]; d2 = list((ge - ee).items())[:5]
    print("in exp not got:", d1); print("in got not exp:", d2)
