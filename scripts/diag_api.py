#!/usr/bin/env python
"""The reference-facing classes end to end on workload C (host table -> clusters), with a cProfile of the host side:
HyperExtractor.from_arrays -> HyperPartition(...).single_partition / incremental_partition.  Diagnostic.

    python scripts/diag_api.py [m] [phasing]
"""
import cProfile
import io
import os
import pstats
import sys
import time

import numpy as np
import pandas as pd
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench                                                       # noqa: E402
from cphasing_b200 import hypergraph as hg, synth                  # noqa: E402
from cphasing_b200.hyperpartition import HyperPartition            # noqa: E402

cfg = dict(synth.CONFIGS["C"])
cfg["m"] = int(sys.argv[1]) if len(sys.argv) > 1 else cfg["m"]
phasing = len(sys.argv) > 2
g, t = bench.gen_table(cfg, 0, "cuda")
host = {k: v.cpu().pin_memory() for k, v in t.items()}
del t
sizes = pd.DataFrame({"length": pd.Series(g.contigsizes)})
al = synth.allele_pairs(g, seed=cfg["seed"] + 1)
names = g.names
at = pd.DataFrame([(0, names[a], names[b], 0, 0, 0, 0.9, 1) for a, b, *_ in al],
                  columns=[0, 1, 2, "mz1", "mz2", "mzShared", "similarity", "strand"])


class AT:
    data = at
    filename = "synthetic"


def run():
    ex = hg.HyperExtractor.from_arrays(host["read_idx"], host["contig_idx"], host["mapq"], g.contig_idx, g.contigsizes,
                                       min_quality=1, min_order=2, max_order=50, edge_length=0)
    torch.cuda.synchronize(); t1 = time.perf_counter()
    chrom, ploidy = cfg["chromosomes"], cfg["ploidy"]
    k = [chrom, ploidy] if phasing else [chrom * ploidy]
    hp = HyperPartition(ex.edges, sizes, k=k, alleletable=AT(), min_contacts=25, min_weight=2.0, min_cis_weight=20,
                        resolution1=-1, resolution2=-1, init_resolution1=1.0, init_resolution2=1.0,
                        min_scaffold_length=2e6, allelic_factor=0, cross_allelic_factor=0)
    torch.cuda.synchronize(); t2 = time.perf_counter()
    if phasing:
        hp.incremental_partition(k)
    else:
        hp.single_partition(k=k[0], is_kprune=True)
    torch.cuda.synchronize(); t3 = time.perf_counter()
    return t1, t2, t3, hp


for rep in range(3):
    t0 = time.perf_counter()
    t1, t2, t3, hp = run()
    print(f"rep {rep}: extract {1e3*(t1-t0):.1f} ms  construct {1e3*(t2-t1):.1f} ms  partition {1e3*(t3-t2):.1f} ms  "
          f"total {1e3*(t3-t0):.1f} ms  groups {len(hp.K)}", flush=True)
pr = cProfile.Profile()
pr.enable()
run()
pr.disable()
s = io.StringIO()
pstats.Stats(pr, stream=s).sort_stats(os.environ.get("SORT", "cumulative")).print_stats(45)
print(s.getvalue()[:9000])

# wall-clock (synchronised) per major call of one more run
import functools
from cphasing_b200 import hyperpartition as _hpm, kprune as _kpm
from cphasing_b200.algorithms import hypergraph as _ahm
acc = {}


def timed(owner, name, label=None):
    raw = getattr(owner, name)            # a staticmethod comes back as the plain function

    @functools.wraps(raw)
    def w(*a, **k):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        r = raw(*a, **k)
        torch.cuda.synchronize()
        acc[label or name] = acc.get(label or name, 0.0) + (time.perf_counter() - t0) * 1e3
        return r
    setattr(owner, name, staticmethod(w) if isinstance(owner.__dict__.get(name), staticmethod) else w)


timed(_hpm.HyperPartition, "kprune")
timed(_hpm, "IRMM")
timed(_ahm.HyperGraph, "clique_expansion_init")
timed(_ahm.HyperGraph, "filter_adjacency_matrix")
timed(_hpm.HyperPartition, "filter_cluster")
timed(_kpm, "kprune_arrays")
timed(_kpm, "kprune_device")
t0 = time.perf_counter()
run()
print("timed run total %.1f ms:" % ((time.perf_counter() - t0) * 1e3), {k: round(v, 1) for k, v in acc.items()})
