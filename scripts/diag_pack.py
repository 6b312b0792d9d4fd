#!/usr/bin/env python
"""Extraction from HOST columns (cph_porec_to_edges_host) for several packing thread counts / chunk sizes.  Diagnostic."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench                                         # noqa: E402
from cphasing_b200 import device as D, synth         # noqa: E402

cfg = dict(synth.CONFIGS["C"])
if len(sys.argv) > 1:
    cfg["m"] = int(sys.argv[1])
g, t = bench.gen_table(cfg, 0, "cuda")
host = [t[k].cpu().pin_memory() for k in ("read_idx", "contig_idx", "mapq")]
print("records:", host[0].numel(), [h.dtype for h in host], flush=True)
del t
for threads, chunk in ((0, 0), (4, 0), (8, 0), (12, 0), (16, 0), (16, 1 << 18), (16, 1 << 22), (32, 0)):
    D.STREAM_PACK_THREADS = threads
    D.STREAM_PACK_CHUNK_RECORDS = chunk
    ts = []
    for _ in range(5):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        raw = D.porec_to_edges(*host, g.n, min_mapq=1)
        torch.cuda.synchronize()
        ts.append((time.perf_counter() - t0) * 1e3)
        del raw
    print(f"threads={threads} chunk={chunk}: {min(ts):.1f} ms (median {sorted(ts)[2]:.1f})", flush=True)
