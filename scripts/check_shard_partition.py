#!/usr/bin/env python
"""The whole hyperpartition through the public classes on concatemer-column shards (north_star's 8-GPU form), under torchrun:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 scripts/check_shard_partition.py

Every rank extracts the hyperedges of ITS concatemer range and runs HyperPartition(column_shard=True): single_partition with
the in-process kprune, and the two-round incremental_partition with an allele table.  Rank 0 repeats both on the whole
table in one process; the groups must be identical.  One JSON line."""
import json
import os
import sys
import time

import pandas as pd
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    from cphasing_b200 import hypergraph as hg, parallel as par, synth
    from cphasing_b200.hyperpartition import HyperPartition
    chrom, ploidy, n, m = 6, 4, 1200, 2_400_000
    g = synth.make_genome(n, chrom, ploidy, seed=15)
    sizes = pd.DataFrame({"length": pd.Series(g.contigsizes)})
    names = g.names
    at = pd.DataFrame([(0, names[a], names[b], m1, m2, sh, sim, 1) for a, b, m1, m2, sh, sim in synth.allele_pairs(g, seed=16)],
                      columns=[0, 1, 2, "mz1", "mz2", "mzShared", "similarity", "strand"])

    class AT:
        data = at
        filename = "synthetic"
    kw = dict(min_contacts=5, min_weight=1.0, min_cis_weight=1, resolution1=-1, resolution2=-1, init_resolution1=1.0,
              init_resolution2=1.0, min_scaffold_length=1e5, allelic_factor=0, cross_allelic_factor=0)

    def edges_of(first, count):
        t = synth.porec_table(g, count, seed=17, first=first)
        return hg.HyperExtractor.from_arrays(t["read_idx"], t["contig_idx"], t["mapq"], g.contig_idx, g.contigsizes,
                                             min_quality=1, edge_length=0).edges

    def run(edges, shard):
        out = {}
        hp = HyperPartition(edges, sizes, k=[chrom * ploidy], alleletable=AT(), column_shard=shard, **kw)
        hp.single_partition(k=chrom * ploidy, is_kprune=True)
        out["single"] = sorted(sorted(hp.vertices[v] for v in grp) for grp in hp.K)
        torch.cuda.synchronize(); t0 = time.perf_counter()
        hp = HyperPartition(edges, sizes, k=[chrom, ploidy], alleletable=AT(), column_shard=shard, merge_use_allele=True, **kw)
        hp.fan_out_groups = False         # (the one-process reference run on rank 0 must not wait for the other ranks)
        hp.incremental_partition([chrom, ploidy])
        torch.cuda.synchronize(); out["incremental_s"] = time.perf_counter() - t0
        out["incremental"] = sorted(sorted(hp.vertices[v] for v in grp) for grp in hp.K)
        return out
    first, count = par.shard_range(m, rank, world)
    mine = run(edges_of(first, count), True)
    res = {"world": world, "groups_single": len(mine["single"]), "groups_incremental": len(mine["incremental"]),
           "sharded_incremental_s": round(mine["incremental_s"], 4)}
    ok = True
    if rank == 0:
        whole = run(edges_of(0, m), False)
        res["single_identical"] = mine["single"] == whole["single"]
        res["incremental_identical"] = mine["incremental"] == whole["incremental"]
        res["one_process_incremental_s"] = round(whole["incremental_s"], 4)
        ok = res["single_identical"] and res["incremental_identical"]
    # every rank must hold the same groups
    import zlib
    same = torch.tensor([zlib.crc32(json.dumps(mine["incremental"]).encode())], dtype=torch.int64, device=f"cuda:{local}")
    lo, hi = same.clone(), same.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN); dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    res["all_ranks_agree"] = bool(lo.item() == hi.item())
    ok = ok and res["all_ranks_agree"]
    dist.barrier()
    if rank == 0:
        res["ok"] = bool(ok)
        print(json.dumps(res), flush=True)
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
