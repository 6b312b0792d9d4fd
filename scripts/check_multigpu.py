#!/usr/bin/env python
"""Multi-GPU parity check, run under torchrun with one rank per GPU:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 scripts/check_multigpu.py

Every rank generates its concatemer-column shard of one synthetic Pore-C table, runs stage 1 + the sharded expansion
(row-sum all-reduce, triangle reduce-scatter + all-gather over NCCL) and rank 0 compares the result with the
single-GPU run over the whole table: the integer contact-count matrix must be BIT-EXACT, the weighted A equal to
1e-9 relative (fixed-point sums are order-free, so they are in fact identical), H shards must tile the global H.
Prints one JSON line on rank 0 and exits non-zero on any mismatch."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device(f"cuda:{local}")
    dist.init_process_group("nccl", device_id=dev)
    from cphasing_b200 import device as D, parallel as par, synth
    import argparse
    ap = argparse.ArgumentParser()
    ap.add_argument("--contigs", type=int, default=3000, help="contigs; >= 5600 takes the swept expansion with the per-band "
                                                        "exchange under the next band's kernels (CPH_BAND_MB sets the band size)")
    ap.add_argument("--concatemers", type=int, default=4_000_000)
    a = ap.parse_args()
    n, m = a.contigs, a.concatemers
    g = synth.make_genome(n, 10, 4, seed=77)
    first, count = par.shard_range(m, rank, world)
    t = synth.porec_table_device(g, count, 78, law="poisson", first=first, device=dev)
    raw = D.porec_to_edges(t["read_idx"], t["contig_idx"], t["mapq"], g.n, min_mapq=1)
    out = {"world": world, "n": n, "m": m}
    ok = True
    for use_high_order, mw in ((False, 0.0), (True, 2.0)):
        H, A = par.sharded_clique_expansion(raw, min_contacts=25, min_weight=mw, use_high_order=use_high_order)
        nnz_h = torch.tensor([H.nnz], dtype=torch.int64, device=dev)
        dist.all_reduce(nnz_h)
        if rank == 0:
            tf = synth.porec_table_device(g, m, 78, law="poisson", first=0, device=dev)
            rawf = D.porec_to_edges(tf["read_idx"], tf["contig_idx"], tf["mapq"], g.n, min_mapq=1)
            Hf = D.incidence(rawf, min_contacts=25, min_mapq=0)
            Af, _ = D.clique_expansion(Hf, use_high_order=use_high_order, min_weight=mw)
            same_pattern = bool(torch.equal(A.indptr, Af.indptr) and torch.equal(A.indices, Af.indices))
            if use_high_order:
                close = same_pattern and bool(torch.allclose(A.data64, Af.data64, rtol=1e-9, atol=0))
                exact = same_pattern and bool(torch.equal(A.data64, Af.data64))
                out["weighted"] = {"same_pattern": same_pattern, "rtol_1e-9": close, "bit_identical": exact,
                                   "nnz_A": int(Af.nnz)}
                ok &= close
            else:
                exact = same_pattern and bool(torch.equal(A.data64, Af.data64))
                out["contact_counts"] = {"bit_exact": exact, "nnz_A": int(Af.nnz),
                                         "sum": float(Af.data64.sum().item())}
                ok &= exact
            out["H_shards_tile_global_H"] = bool(int(nnz_h.item()) == Hf.nnz)
            ok &= out["H_shards_tile_global_H"]
            del tf, rawf, Hf, Af
        dist.barrier()
    # the same through the public classes in column-shard mode (host tables in, scipy CSR out)
    from cphasing_b200 import hypergraph as hgm
    from cphasing_b200.algorithms import hypergraph as ah
    mc = min(m, 1_000_000)
    f2, c2 = par.shard_range(mc, rank, world)
    th = synth.porec_table(g, c2, seed=79, first=f2)
    ex = hgm.HyperExtractor.from_arrays(th["read_idx"], th["contig_idx"], th["mapq"], g.contig_idx, g.contigsizes,
                                        min_quality=1, edge_length=0)
    HG = ah.HyperGraph(ex.edges, min_quality=1)
    HG.column_shard = True
    Hs = HG.incidence_matrix(min_contacts=10)
    As = ah.HyperGraph.clique_expansion_init(Hs, min_weight=1.0)
    if rank == 0:
        tf = synth.porec_table(g, mc, seed=79)
        exf = hgm.HyperExtractor.from_arrays(tf["read_idx"], tf["contig_idx"], tf["mapq"], g.contig_idx, g.contigsizes,
                                             min_quality=1, edge_length=0)
        HGf = ah.HyperGraph(exf.edges, min_quality=1)
        Af = ah.HyperGraph.clique_expansion_init(HGf.incidence_matrix(min_contacts=10), min_weight=1.0)
        same = (As.shape == Af.shape and np.array_equal(As.indptr, Af.indptr) and np.array_equal(As.indices, Af.indices)
                and np.array_equal(As.data, Af.data))
        out["class_layer_column_shard"] = {"A_identical": bool(same), "nnz_A": int(Af.nnz), "nodes_equal":
                                           bool(np.array_equal(HG.nodes, HGf.nodes))}
        ok &= same
    dist.barrier()
    if rank == 0:
        out["ok"] = bool(ok)
        print(json.dumps(out), flush=True)
    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.broadcast(flag, 0)
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) else 1)


if __name__ == "__main__":
    main()
