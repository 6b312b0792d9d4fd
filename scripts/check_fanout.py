#!/usr/bin/env python
"""Second-round fan-out check (SURVEY.md 8e "replicas"), under torchrun with one rank per GPU:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 scripts/check_fanout.py

Every rank builds the same planted genome (chromosomes x haplotypes) and Pore-C table and runs
HyperPartition.incremental_partition: the first round everywhere (deterministic kernels: same groups on every rank), the
second round with the first-round groups dealt round-robin over the ranks and all-gathered.  Rank 0 repeats the run with
the fan-out switched off and compares the groups; prints one JSON line."""
import json
import os
import sys
import time

import numpy as np
import pandas as pd
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    from cphasing_b200 import hypergraph as hg, synth
    from cphasing_b200.hyperpartition import HyperPartition
    chrom, ploidy, n, m = 8, 4, 1600, 3_000_000
    g = synth.make_genome(n, chrom, ploidy, seed=5)
    t = synth.porec_table(g, m, seed=6)
    sizes = pd.DataFrame({"length": pd.Series(g.contigsizes)})

    def run(fan):
        ex = hg.HyperExtractor.from_arrays(t["read_idx"], t["contig_idx"], t["mapq"], g.contig_idx, g.contigsizes,
                                           min_quality=1, edge_length=0)
        hp = HyperPartition(ex.edges, sizes, k=[chrom, ploidy], min_contacts=5, min_weight=1.0, min_cis_weight=1,
                            resolution1=-1, resolution2=-1, init_resolution1=1.0, init_resolution2=1.0,
                            min_scaffold_length=1e5, allelic_factor=0, cross_allelic_factor=0)
        hp.fan_out_groups = fan
        torch.cuda.synchronize(); t0 = time.perf_counter()
        hp.incremental_partition([chrom, ploidy])
        torch.cuda.synchronize()
        return [sorted(x) for x in hp.K], time.perf_counter() - t0
    run(True)                                            # warm-up (workspaces, NCCL)
    K_fan, t_fan = run(True)
    out = {"world": world, "groups": len(K_fan), "fan_out_s": round(t_fan, 4)}
    ok = True
    if rank == 0:
        K_one, t_one = run(False)
        out["single_rank_s"] = round(t_one, 4)
        out["same_groups"] = sorted(K_fan) == sorted(K_one)
        mem = np.full(g.n, -1)
        for gi, grp in enumerate(K_fan):
            mem[grp] = gi
        from oracle import community_oracle as co            # checker only
        okv = mem >= 0
        out["ari_vs_planted"] = float(co.ari(mem[okv], g.group[okv])) if okv.any() else None
        ok = out["same_groups"]
    # what a row-partitioned sweep would pay per phase: an all-gather of the label vector (n = 20 000 x int32, workload D)
    lab = torch.zeros(20_000 // world, dtype=torch.int32, device=f"cuda:{local}")
    full = torch.empty(20_000 // world * world, dtype=torch.int32, device=f"cuda:{local}")
    for _ in range(20):
        dist.all_gather_into_tensor(full, lab)
    torch.cuda.synchronize(); dist.barrier(); t0 = time.perf_counter()
    for _ in range(500):
        dist.all_gather_into_tensor(full, lab)
    torch.cuda.synchronize()
    out["label_allgather_us"] = round((time.perf_counter() - t0) / 500 * 1e6, 1)
    dist.barrier()
    if rank == 0:
        out["ok"] = bool(ok)
        print(json.dumps(out), flush=True)
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
