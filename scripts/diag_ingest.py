#!/usr/bin/env python
"""File -> hyperedges -> groups wall time: a workload-C-shaped Pore-C table of --m concatemers written as
a plain .porec text file (and .gz), decoded by (a) the streaming ingest of this repo feeding the GPU chunk by chunk and
(b) pandas.read_csv of the whole file (what round 1 did, and what the oracle port of the reference does in place of polars).

    python scripts/diag_ingest.py [--m 5000000] [--gz]
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import pandas as pd
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cphasing_b200 import hypergraph as hg, synth                     # noqa: E402
from cphasing_b200.hyperpartition import HyperPartition               # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--m", type=int, default=5_000_000)
    ap.add_argument("--gz", action="store_true")
    ap.add_argument("--dir", default="/tmp")
    a = ap.parse_args()
    cfg = dict(synth.CONFIGS["C"])
    g = synth.make_genome(cfg["n"], cfg["chromosomes"], cfg["ploidy"], cfg["seed"])
    t = synth.porec_table_device(g, a.m, cfg["seed"], law=cfg["law"], first=0, positions=True, device="cuda")
    names = np.array(g.names, dtype=object)
    df = pd.DataFrame({"read_idx": t["read_idx"].cpu().numpy().view(np.uint32), "len": 10000, "rs": 0, "re": 900, "strand": "+",
                       "chrom": names[t["contig_idx"].cpu().numpy().view(np.uint32).astype(np.int64)],
                       "start": t["start"].cpu().numpy().view(np.uint32), "end": t["end"].cpu().numpy().view(np.uint32),
                       "mapq": t["mapq"].cpu().numpy(), "identity": 0.97, "filter": "pass"})
    path = os.path.join(a.dir, "diag_ingest.porec" + (".gz" if a.gz else ""))
    t0 = time.perf_counter()
    df.to_csv(path, sep="\t", header=False, index=False, compression="gzip" if a.gz else None)
    out = {"concatemers": a.m, "records": len(df), "file_bytes": os.path.getsize(path), "gz": a.gz,
           "write_s": round(time.perf_counter() - t0, 2), "host_cores": len(os.sched_getaffinity(0))}
    del df, t
    sizes = pd.DataFrame({"length": pd.Series(g.contigsizes)})

    def groups(ex):
        hp = HyperPartition(ex.edges, sizes, k=[g.G], min_contacts=25, min_weight=2.0, min_cis_weight=20, resolution1=-1,
                            init_resolution1=1.0, min_scaffold_length=2e6, allelic_factor=0, cross_allelic_factor=0)
        hp.single_partition(k=g.G)
        return len(hp.K)
    for rep in range(2):                                              # second pass: page cache + warm workspaces
        torch.cuda.synchronize(); t0 = time.perf_counter()
        ex = hg.HyperExtractor([path], g.contig_idx, g.contigsizes, min_quality=1, edge_length=0, threads=out["host_cores"])
        torch.cuda.synchronize(); t1 = time.perf_counter()
        k = groups(ex)
        torch.cuda.synchronize(); t2 = time.perf_counter()
        out["streaming"] = {"file_to_hyperedges_s": round(t1 - t0, 3), "hyperedges_to_groups_s": round(t2 - t1, 3), "groups": k,
                            "MB_per_s": round(out["file_bytes"] / (t1 - t0) / 1e6)}
    t0 = time.perf_counter()
    d2 = pd.read_csv(path, sep="\t", header=None, comment="#", usecols=[0, 5, 6, 7, 8],
                     names=["read_idx", "chrom", "start", "end", "mapq"],
                     dtype={"read_idx": np.uint32, "start": np.uint32, "end": np.uint32, "mapq": np.int8})
    codes = hg._name_codes(d2["chrom"], g.contig_idx)
    t1 = time.perf_counter()
    ex2 = hg.HyperExtractor.from_arrays(d2["read_idx"].to_numpy(), codes, d2["mapq"].to_numpy(), g.contig_idx, g.contigsizes,
                                        min_quality=1, edge_length=0)
    torch.cuda.synchronize(); t2 = time.perf_counter()
    out["pandas_whole_table"] = {"read_csv_s": round(t1 - t0, 3), "arrays_to_hyperedges_s": round(t2 - t1, 3)}
    same = all(np.array_equal(np.asarray(getattr(ex.edges, f)), np.asarray(getattr(ex2.edges, f))) for f in ("row", "col", "count", "mapq"))
    out["identical_hyperedges"] = bool(same)
    os.remove(path)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
