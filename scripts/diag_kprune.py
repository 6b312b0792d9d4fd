#!/usr/bin/env python
"""Where the kprune step of the partition tail spends its time (workload D shard by default).  Diagnostic."""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench                                                  # noqa: E402
from cphasing_b200 import device as D, kprune, synth          # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "D"
cfg = dict(synth.CONFIGS[name])
if name == "D":
    cfg["m"] //= 8
g, t = bench.gen_table(cfg, 0, "cuda")
raw = D.porec_to_edges(t["read_idx"], t["contig_idx"], t["mapq"], g.n, min_mapq=1)
H = D.incidence(raw, min_contacts=bench.PARAMS["min_contacts"], min_mapq=0)
del t, raw
keep = H.row_keep.cpu().numpy() if H.row_keep is not None else np.ones(g.n, bool)
new = np.cumsum(keep) - 1
pairs = np.array([(new[a], new[b]) for a, b, *_ in synth.allele_pairs(g, seed=cfg["seed"] + 1) if keep[a] and keep[b]],
                 dtype=np.int64).reshape(-1, 2)
print("allele pairs:", len(pairs), flush=True)


def T(label, fn, reps=3):
    best = 1e9
    for _ in range(reps):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        r = fn()
        torch.cuda.synchronize(); best = min(best, (time.perf_counter() - t0) * 1e3)
    print(f"{label}: {best:.2f} ms", flush=True)
    return r


Au, _ = T("unweighted expansion", lambda: D.clique_expansion(H, use_high_order=False, min_weight=1.0, want32=False, want64=True))
print("nnz(Au):", Au.nnz)
T("  accumulate only", lambda: D.expand_accumulate(H, use_high_order=False))
acc = D.expand_accumulate(H, use_high_order=False)
T("  emit only", lambda: D.expand_emit(acc, min_weight=1.0, want32=False, want64=True))
T("allele_groups (host)", lambda: kprune.allele_groups(pairs, Au.n, 1))
flag, row_of = T("kprune_device", lambda: kprune.kprune_device(Au, pairs, 1))


def post():
    al = torch.nonzero(flag == 1).flatten(); wk = torch.nonzero(flag == 2).flatten()
    Pa = (torch.cat([row_of[al], Au.indices[al]]), torch.cat([Au.indices[al], row_of[al]]))
    Pw = (torch.cat([row_of[wk], Au.indices[wk]]), torch.cat([Au.indices[wk], row_of[wk]]))
    return Pa, Pw


T("flag -> pair lists", post)
