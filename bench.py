#!/usr/bin/env python
"""bench.py -- the hyperpartition hot path on synthetic contact tables of BASELINE.json's shapes.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload C|B|E|A|D] [--impl reference]

One "step" = one pass of the hot path over one table: contact table -> hyperedges (filters,
group-by, order filter) -> filtered incidence matrix H -> clique expansion A = H W H^T (CSR out)
[-> kprune scoring -> Louvain/Leiden when --full].  `value` = concatemers/s with the table already
resident in HBM (CUDA events, max over ranks); `e2e` = the same through the reference-facing
classes with HOST (pinned) buffers in and the scipy CSR of A out.  Prints ONE JSON line.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# default workload at every N: config C (potato-scale Pore-C, 4k contigs, 50M concatemers PER GPU, weak scaling over
# concatemer-column shards; N=1 is exactly the named config).  `--workload D` runs config D's genome (20k contigs) with
# 62.5M concatemers per GPU (the named 500M at N=8); its numbers are recorded in profiles/ and DESIGN.md.
# dram bytes per launch of the dominant kernel from the committed ncu --set full capture, keyed by (workload, m per GPU)
NCU_TRAFFIC = {("C", 50_000_000): 2721182000 + 2306597000}      # profiles/r1j_ncu_full_top3.txt
PARAMS = dict(min_mapq=1, min_order=2, max_order=50, min_contacts=25, min_weight=2.0)   # cphasing/_config.py:38-45


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=None, help="A|B|C|D|E (default C; every GPU takes a full-size shard, D: 1/8 of its 500M)")
    ap.add_argument("--m", type=int, default=None, help="override the number of concatemers per GPU")
    ap.add_argument("--cpu-sample", type=int, default=10_000_000, help="concatemers in the CPU-baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-partition", action="store_true", help="skip the kprune + clustering tail (hyperpartition wall s)")
    return ap.parse_args()


def workload(args):
    from cphasing_b200 import synth
    name = args.workload or "C"
    cfg = dict(synth.CONFIGS[name])
    if name == "D":
        cfg["m"] = cfg["m"] // 8                 # per GPU: 62.5M (the named 500M at 8 GPUs)
    elif args.gpus > 1:
        pass                                     # other shapes: every GPU takes a full-size shard
    if args.m:
        cfg["m"] = args.m
    return name, cfg


class Clocks:
    """nvidia-smi sampler for the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.p = None
        self.path = f"/tmp/cph_clocks_{os.getpid()}.csv"

    def start(self):
        try:
            self.f = open(self.path, "w")
            self.p = subprocess.Popen(["nvidia-smi", f"--id={self.idx}", f"--query-gpu={self.Q}",
                                       "--format=csv,noheader,nounits", "-lms", "100"], stdout=self.f,
                                      stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.close()
        sm, mx, reasons = [], [], set()
        for line in open(self.path):
            f = [x.strip() for x in line.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if sm:
            out = {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}
        try:
            os.remove(self.path)
        except OSError:
            pass
        return out


# ------------------------------------------------------------------------------------------------
# own arm
# ------------------------------------------------------------------------------------------------
def gen_table(cfg, rank, dev):
    from cphasing_b200 import synth
    g = synth.make_genome(cfg["n"], cfg["chromosomes"], cfg["ploidy"], cfg["seed"])
    first = rank * cfg["m"]
    if cfg["kind"] == "pairs":
        t = synth.pairs_table_device(g, cfg["m"], cfg["seed"], first=first, device=dev)
    else:
        t = synth.porec_table_device(g, cfg["m"], cfg["seed"], law=cfg["law"], first=first, device=dev)
    return g, t


def device_step(D, cfg, g, t, ev, world, rank):
    """one pass with the table resident in HBM; ev: list of (name, cuda event) appended in order."""
    import torch
    import torch.distributed as dist

    def mark(name):
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        ev.append((name, e))
    mark("start")
    if cfg["kind"] == "pairs":
        raw = D.pairs_to_edges(t["chrom1"], t["chrom2"], t["mapq"], g.n, min_mapq=PARAMS["min_mapq"])
    else:
        raw = D.porec_to_edges(t["read_idx"], t["contig_idx"], t["mapq"], g.n, min_mapq=PARAMS["min_mapq"],
                               min_order=PARAMS["min_order"], max_order=PARAMS["max_order"])
    mark("edges")
    from cphasing_b200 import parallel as par
    # exchange 1 (world > 1): global per-contig contact sums + largest count for the min_contacts row filter
    H = D.incidence(raw, min_contacts=PARAMS["min_contacts"], min_mapq=0,
                    rowsum_hook=par.allreduce_rowsums if world > 1 else None)
    mark("incidence")
    # the fixed-point scale derives from the (all-reduced) row sums, so every rank picks the same one
    acc = D.expand_accumulate(H, use_high_order=True)
    mark("accumulate")
    if world > 1:
        par.sum_triangle(acc.acc, acc.n)     # exchange 2: reduce-scatter over contig row blocks + all-gather (NCCL)
        par.sum_vector(acc.diag)
        mark("exchange")
    A = D.expand_emit(acc, min_weight=PARAMS["min_weight"], want32=True, want64=True)
    mark("emit")
    return raw, H, A


def partition_tail(D, cfg, g, H, A, reps=3, world=1, rank=0):
    """the rest of `hyperpartition` on the device-resident H / A of the last step: unweighted expansion -> kprune
    (allelic + cross-allelic pairs from the synthetic allele table) -> pair re-weighting -> Louvain at resolution 1.
    Wall-clock (synchronised) per stage, best of `reps`; ARI of the clustering against the planted haplotype groups."""
    import torch
    from cphasing_b200 import community, graphops, kprune, synth
    keep = H.row_keep.cpu().numpy() if H.row_keep is not None else np.ones(g.n, bool)
    new = np.cumsum(keep) - 1
    pairs = np.array([(new[a], new[b]) for a, b, *_ in synth.allele_pairs(g, seed=cfg["seed"] + 1) if keep[a] and keep[b]],
                     dtype=np.int64).reshape(-1, 2)
    best = {}
    out = {}
    G = g.G
    for _ in range(reps):
        tm = {}
        torch.cuda.synchronize(); t0 = time.perf_counter()
        if world > 1:       # H is this rank's column shard: the unweighted counts need the same exchange as A
            from cphasing_b200 import parallel as par
            accu = D.expand_accumulate(H, use_high_order=False)
            par.sum_triangle(accu.acc, accu.n)
            par.sum_vector(accu.diag)
            Au = D.expand_emit(accu, min_weight=1.0, want32=False, want64=True)
            del accu
            if rank != 0:   # kprune and the clustering run on one GPU (latency-bound at n <= 2e4)
                continue
        else:
            Au, _ = D.clique_expansion(H, use_high_order=False, min_weight=1.0, want32=False, want64=True)
        flag, row_of = kprune.kprune_device(Au, pairs, 1)
        al = torch.nonzero(flag == 1).flatten(); wk = torch.nonzero(flag == 2).flatten()
        Pa = (torch.cat([row_of[al], Au.indices[al]]), torch.cat([Au.indices[al], row_of[al]]))     # symmetrised, on device
        Pw = (torch.cat([row_of[wk], Au.indices[wk]]), torch.cat([Au.indices[wk], row_of[wk]]))
        torch.cuda.synchronize(); tm["kprune_ms"] = (time.perf_counter() - t0) * 1e3; t0 = time.perf_counter()
        a = graphops.clone_adjacency(A)
        graphops.reweight_device(a, Pa, Pw, 0.0, 0.0)
        # resolution auto-search of hyperpartition.py:1049-1078: start at 1.0, +0.2 until at least G groups
        gamma, rounds = 1.0, 0
        while True:
            mem, q, st = community.community_device(a, gamma, "louvain")
            rounds += 1
            if st["groups"] >= G or rounds >= 50:
                break
            gamma += 0.2
        torch.cuda.synchronize(); tm["cluster_ms"] = (time.perf_counter() - t0) * 1e3
        for k_, v in tm.items():
            best[k_] = min(best.get(k_, v), v)
        out = {"allelic_pairs": int(al.numel()), "weak_pairs": int(wk.numel()), "groups": st["groups"], "planted_groups": G,
               "modularity": q, "resolution": round(gamma, 2), "clusterings": rounds, "cluster_ms_each": tm["cluster_ms"] / rounds}
    if rank != 0:
        return None
    from oracle import community_oracle as co          # checker only: ARI against the planted groups
    truth = g.group[keep]
    out["ari_vs_planted"] = float(co.ari(mem.cpu().numpy(), truth))
    out.update(best)
    return out


def e2e_step(cfg, g, host, names, trace=None):
    """host (pinned) buffers -> reference-facing classes -> scipy CSR of A on the host."""
    import torch
    from cphasing_b200 import hypergraph as hg
    from cphasing_b200.algorithms import hypergraph as ah
    idx, sizes = names
    if cfg["kind"] == "pairs":
        ex = hg.Extractor.from_arrays(host["chrom1"], host["chrom2"], host["mapq"], idx, sizes,
                                      min_quality=PARAMS["min_mapq"], edge_length=0)
    else:
        ex = hg.HyperExtractor.from_arrays(host["read_idx"], host["contig_idx"], host["mapq"], idx, sizes,
                                           min_quality=PARAMS["min_mapq"], min_order=PARAMS["min_order"],
                                           max_order=PARAMS["max_order"], edge_length=0)
    tr = trace is not None
    if tr:
        torch.cuda.synchronize(); trace["extract (H2D + group-by)"] = time.perf_counter()
    HG = ah.HyperGraph(ex.edges, min_quality=PARAMS["min_mapq"], mapq_filter=False)
    H = HG.incidence_matrix(PARAMS["min_contacts"])
    if tr:
        torch.cuda.synchronize(); trace["incidence"] = time.perf_counter()
    A = ah.HyperGraph.clique_expansion_init(H, min_weight=PARAMS["min_weight"])
    if tr:
        torch.cuda.synchronize(); trace["expansion"] = time.perf_counter()
    d2h = A.data.nbytes + A.indices.nbytes + A.indptr.nbytes        # touching .data materialises the scipy CSR
    torch.cuda.synchronize()
    if tr:
        trace["D2H + scipy csr"] = time.perf_counter()
    return A, d2h


def api_partition(cfg, g, host):
    """`hyperpartition` as a user of the reference runs it, through the public classes only (kind = porec): host table ->
    HyperExtractor -> HyperPartition(alleletable) -> single_partition(k, is_kprune=True) with the resolution auto-search
    -> groups.  Wall seconds, best of 2, everything (pandas bookkeeping included) inside."""
    import pandas as pd
    import torch
    from cphasing_b200 import hypergraph as hg, synth
    from cphasing_b200.hyperpartition import HyperPartition
    from oracle import community_oracle as co                    # checker only (ARI against the planted groups)
    sizes = pd.DataFrame({"length": pd.Series(g.contigsizes)})
    names = g.names
    at = pd.DataFrame([(0, names[a], names[b], 0, 0, 0, 0.9, 1) for a, b, *_ in synth.allele_pairs(g, seed=cfg["seed"] + 1)],
                      columns=[0, 1, 2, "mz1", "mz2", "mzShared", "similarity", "strand"])

    class AT:
        data = at
        filename = "synthetic"
    G = g.G
    best, hp = None, None
    for _ in range(2):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        ex = hg.HyperExtractor.from_arrays(host["read_idx"], host["contig_idx"], host["mapq"], g.contig_idx, g.contigsizes,
                                           min_quality=PARAMS["min_mapq"], min_order=PARAMS["min_order"],
                                           max_order=PARAMS["max_order"], edge_length=0)
        hp = HyperPartition(ex.edges, sizes, k=[G], alleletable=AT(), min_contacts=PARAMS["min_contacts"],
                            min_weight=PARAMS["min_weight"], min_cis_weight=20, resolution1=-1, init_resolution1=1.0,
                            min_scaffold_length=2e6, allelic_factor=0, cross_allelic_factor=0)
        hp.single_partition(k=G, is_kprune=True)
        torch.cuda.synchronize(); dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    idx = {nm: i for i, nm in enumerate(names)}
    mem = np.full(len(names), -1)
    for gi, grp in enumerate(hp.K):
        for v in grp:
            mem[idx[hp.vertices[v]]] = gi
    ok = mem >= 0
    return {"wall_s": best, "groups": len(hp.K), "ari_vs_planted": float(co.ari(mem[ok], g.group[ok])),
            "api": "HyperExtractor.from_arrays -> HyperPartition(alleletable=...).single_partition(k, is_kprune=True), "
                   "resolution auto-search; host table in, groups out"}


def cpu_baseline(cfg, g, t, n_sample):
    """the oracle (numpy/scipy port of the reference path) on the first n_sample concatemers of the same
    table, one core."""
    import torch
    from oracle import hypergraph_oracle as orc
    if cfg["kind"] == "pairs":
        k = min(n_sample, cfg["m"])
        c1 = t["chrom1"][:k].cpu().numpy(); c2 = t["chrom2"][:k].cpu().numpy(); q = t["mapq"][:k].cpu().numpy()
        t0 = time.perf_counter()
        r, c, cnt, mq = orc.pairs_edges(c1, c2, q, min_mapq=PARAMS["min_mapq"], n_contigs=g.n)
    else:
        rid = t["read_idx"].view(torch.int32)
        first = int(rid[0].item())
        k = min(n_sample, cfg["m"])
        cut = int(torch.searchsorted(rid, torch.tensor([first + k], dtype=torch.int32, device=rid.device)).item())
        ri = t["read_idx"][:cut].cpu().numpy(); ci = t["contig_idx"][:cut].cpu().numpy(); q = t["mapq"][:cut].cpu().numpy()
        t0 = time.perf_counter()
        r, c, cnt, mq = orc.porec_edges(ri, ci, q, min_mapq=PARAMS["min_mapq"], min_order=PARAMS["min_order"],
                                        max_order=PARAMS["max_order"], n_contigs=g.n)
    rr, cc, kk, ncol = orc.hypergraph_get_data(r, c, cnt, mq, PARAMS["min_mapq"], mapq_filter=False)
    H, _, _ = orc.incidence_matrix(rr, cc, kk, g.n, ncol, PARAMS["min_contacts"])
    A = orc.clique_expansion(H, min_weight=PARAMS["min_weight"], dtype=np.float32)     # the reference's own float32 path
    dt = time.perf_counter() - t0
    return {"value": k / dt, "unit": "concatemers/s", "cores": 1, "kind": "port",
            "sample": f"first {k} concatemers of the same table ({len(r)} H entries), oracle/hypergraph_oracle.py "
                      f"(numpy+scipy restatement; polars/MKL absent), {dt:.1f} s"}, A


def run_own(args):
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback); use --impl reference for the CPU arm")
    torch.cuda.set_device(local)
    dev = torch.device(f"cuda:{local}")
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    from cphasing_b200 import _lib, device as D
    name, cfg = workload(args)
    g, t = gen_table(cfg, rank, dev)
    n_rec = int(next(iter(t.values())).numel())
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        ev = []
        raw, H, A = device_step(D, cfg, g, t, ev, world, rank)
    barrier()
    clocks = Clocks(local)
    if rank == 0:
        clocks.start()
    l0 = _lib.launches()
    all_ev = []
    t_start = torch.cuda.Event(enable_timing=True); t_end = torch.cuda.Event(enable_timing=True)
    barrier()
    t_start.record()
    for _ in range(args.steps):
        ev = []
        raw, H, A = device_step(D, cfg, g, t, ev, world, rank)
        all_ev.append(ev)
    t_end.record()
    barrier()
    launches = _lib.launches() - l0
    ms_total = t_start.elapsed_time(t_end)
    if world > 1:
        mt = torch.tensor([ms_total], dtype=torch.float64, device=dev)
        dist.all_reduce(mt, op=dist.ReduceOp.MAX)
        ms_total = float(mt.item())
    clk = clocks.stop() if rank == 0 else None
    ms_step = ms_total / args.steps
    stages = {}
    for ev in all_ev:
        for (n0, e0), (n1, e1) in zip(ev[:-1], ev[1:]):
            stages.setdefault(n1, []).append(e0.elapsed_time(e1))
    stages = {k: sum(v) / len(v) for k, v in stages.items()}

    # roofline of the DOMINANT kernel (profiles/r1j_launch_summary.txt: porec_tile_kernel = 42 % of the step on Pore-C
    # tables; on pairs tables the expansion's expand_diag_kernel).  ALGORITHMIC bytes:
    #   porec_tile_kernel: 9 B per alignment record read + 8 B per surviving (read, contig) entry written
    #   expand_diag_kernel: 8 B per H entry read (row + packed count/degree); its work figure is the pair REDs
    # duration = CUDA events around the C-ABI call that launches it, averaged over the timed steps.
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    nnzH, nnzA, n_rows = H.nnz, A.nnz, H.n_rows
    acc_ms = stages.get("accumulate", float("nan"))
    edges_ms = stages.get("edges", float("nan"))
    stage1_bytes = 9 * n_rec + 8 * raw.nnz + 8 * (g.n + 1)
    stage2_bytes = 8 * nnzH + 4 * H.n_cols + 12 * nnzA + 16 * (n_rows + 1)
    s1_ms = stages.get("edges", 0) + stages.get("incidence", 0)
    s2_ms = stages.get("accumulate", 0) + stages.get("emit", 0)
    per_kernel = {"expand_diag_kernel (cph_expand_accumulate_diag)":
                  {"bytes": 8 * nnzH, "ms": acc_ms, "GBps": 8 * nnzH / acc_ms / 1e6, "frac": 8 * nnzH / acc_ms / 1e6 / peak}}
    if cfg["kind"] == "pairs":
        kname, kbytes, kms, traffic = "expand_diag_kernel (cph_expand_accumulate_diag)", 8 * nnzH, acc_ms, None
    else:
        kname, kbytes, kms = "porec_tile_kernel (cph_porec_to_edges)", 9 * n_rec + 8 * raw.nnz, edges_ms
        # ncu --set full, one launch on this workload: dram__bytes_read + dram__bytes_write (profiles/r1j_ncu_full_top3.txt)
        traffic = NCU_TRAFFIC.get((name, cfg["m"]))
        per_kernel[kname] = {"bytes": kbytes, "ms": kms, "GBps": kbytes / kms / 1e6, "frac": kbytes / kms / 1e6 / peak}
    roof = {"bound": "hbm", "kernel": kname, "achieved": kbytes / kms / 1e6, "peak": peak, "unit": "GB/s",
            "frac": kbytes / kms / 1e6 / peak, "traffic": traffic,
            "peak_source": "MEASURED_PEAKS.json hbm_gbs (measured copy)" if peaks else "fallback 6650",
            "algorithmic_bytes_per_launch": kbytes, "kernel_ms": kms, "kernels": per_kernel,
            "stage1_GBps": stage1_bytes / s1_ms / 1e6 if s1_ms else None,
            "stage2_GBps": stage2_bytes / s2_ms / 1e6 if s2_ms else None}
    out = {"metric": "concatemers/sec (H build + H*W*H^T)", "value": cfg["m"] * world / (ms_step / 1e3),
           "unit": "concatemers/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
           "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64 fixed-point / f64",
           "data": "synthetic", "gpu_launches": launches,
           "config": {"workload": f"{name}: n={cfg['n']} contigs, m={cfg['m']} concatemers/GPU x {world} GPU, law={cfg['law']}, "
                                  f"kind={cfg['kind']}", "records_per_gpu": n_rec, "nnz_H": nnzH, "nnz_A": nnzA,
                      "params": PARAMS, "l2": "inputs (>2 GB) and intermediates exceed the 126 MB L2; no flush needed",
                      "parallelism": f"concatemer-column shards x{world}" if world > 1 else "single GPU"},
           "stages_ms": stages, "roofline": roof, "clocks": clk}
    if not args.no_partition:
        tail = partition_tail(D, cfg, g, H, A, world=world, rank=rank)
        barrier()
    if rank == 0 and not args.no_partition:
        tail["wall_s"] = (ms_step + tail["kprune_ms"] + tail["cluster_ms"]) / 1e3
        tail["what"] = ("table -> H -> A (ms_per_step) + unweighted expansion + kprune + pair re-weighting + Louvain "
                        "inside the resolution auto-search (1.0, +0.2 until >= planted groups), device-resident")
        out["hyperpartition"] = tail
    if rank == 0 and world == 1 and not args.no_e2e:
        import torch
        host = {k: torch.empty(v.shape, dtype=v.dtype, pin_memory=True).copy_(v) for k, v in t.items()}
        names = (g.contig_idx, g.contigsizes)
        h2d = sum(v.numel() * v.element_size() for v in host.values())
        del t
        _lib.release_workspace(); torch.cuda.empty_cache()
        for _ in range(2):
            Ah, d2h = e2e_step(cfg, g, host, names)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        ne = max(2, min(args.steps, 10))
        each = []
        for _ in range(ne):
            t_s = time.perf_counter()
            Ah, d2h = e2e_step(cfg, g, host, names)
            each.append((time.perf_counter() - t_s) * 1e3)
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / ne
        trace = {}
        torch.cuda.synchronize(); t1 = time.perf_counter()
        e2e_step(cfg, g, host, names, trace)           # one extra, untimed pass with a sync after every stage
        stages_e2e, prev = {}, t1
        for k_, v in trace.items():
            stages_e2e[k_] = (v - prev) * 1e3; prev = v
        from cphasing_b200 import device as Dm
        sent = int(Dm.LAST_H2D_BYTES) or h2d       # what crossed PCIe (host threads re-encode the table to ~4 B per record)
        out["e2e"] = {"value": cfg["m"] / dt, "unit": "concatemers/s", "h2d_bytes_per_step": sent, "d2h_bytes_per_step": d2h,
                      "host_input_bytes_per_step": h2d, "host_pack_threads": Dm._pack_threads(),
                      "ms_per_step": dt * 1e3, "ms_each": each, "stages_ms_synchronised": stages_e2e, "api": "HyperExtractor.from_arrays -> HyperGraph.incidence_matrix -> "
                                                      "HyperGraph.clique_expansion_init -> scipy csr (host)"}
        if not args.no_partition and cfg["kind"] == "porec" and "hyperpartition" in out:
            out["hyperpartition"]["api"] = api_partition(cfg, g, host)
        t = {k: v.to(dev) for k, v in host.items()} if not args.no_cpu_baseline else None
    if world > 1 and not args.no_e2e:
        # every rank feeds its own shard from pinned host memory through the device-module API (chunked H2D under the
        # group-by kernel, the two NCCL exchanges, CSR emit) and rank 0 reads A back; max over ranks
        import torch
        from cphasing_b200 import device as Dm
        host = {k: torch.empty(v.shape, dtype=v.dtype, pin_memory=True).copy_(v) for k, v in t.items()}
        h2d = sum(v.numel() * v.element_size() for v in host.values())
        del t, raw, H, A
        _lib.release_workspace(); torch.cuda.empty_cache()

        def e2e_multi():
            evs = []
            _, _, A2 = device_step(D, cfg, g, host, evs, world, rank)
            nb = 0
            if rank == 0:
                parts = [Dm.to_host(A2.indptr), Dm.to_host(A2.indices), Dm.to_host(A2.data32)]
                nb = sum(x.nbytes for x in parts)
            torch.cuda.synchronize()
            return nb
        for _ in range(2):
            e2e_multi()
        ne = max(2, min(args.steps, 10))
        barrier(); t0 = time.perf_counter()
        for _ in range(ne):
            d2h = e2e_multi()
        barrier()
        dtm = torch.tensor([(time.perf_counter() - t0) / ne], dtype=torch.float64, device=dev)
        dist.all_reduce(dtm, op=dist.ReduceOp.MAX)
        dt = float(dtm.item())
        sent = torch.tensor([int(Dm.LAST_H2D_BYTES) or h2d], dtype=torch.int64, device=dev)
        dist.all_reduce(sent)
        out["e2e"] = {"value": cfg["m"] * world / dt, "unit": "concatemers/s", "h2d_bytes_per_step": int(sent.item()),
                      "host_input_bytes_per_step": h2d * world, "host_pack_threads": Dm._pack_threads(),
                      "d2h_bytes_per_step": d2h, "ms_per_step": dt * 1e3,
                      "api": "per rank: device.porec_to_edges(host columns) -> device.incidence (row-sum all-reduce) -> "
                             "device.expand_accumulate -> triangle reduce-scatter/all-gather -> device.expand_emit; "
                             "rank 0 reads CSR(A) back"}
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cb, _ = cpu_baseline(cfg, g, t, args.cpu_sample)
        out["cpu_baseline"] = cb
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


# ------------------------------------------------------------------------------------------------
# reference arm: the reference's CPU algorithm (oracle port; the reference itself is pure Python over
# polars / scipy / igraph and cannot be installed or travel -- DESIGN.md) on the host cores
# ------------------------------------------------------------------------------------------------
def _ref_worker(conn, cfg, first, m):
    from cphasing_b200 import synth
    from oracle import hypergraph_oracle as orc
    from scipy.sparse import csr_matrix
    g = synth.make_genome(cfg["n"], cfg["chromosomes"], cfg["ploidy"], cfg["seed"])
    if cfg["kind"] == "pairs":
        t = synth.pairs_table(g, m, cfg["seed"], first=first)
    else:
        t = synth.porec_table(g, m, cfg["seed"], law=cfg["law"], first=first)
    conn.send(("ready", len(next(iter(t.values())))))
    st = {}
    while True:
        cmd = conn.recv()
        if cmd[0] == "stop":
            break
        if cmd[0] == "phase1":      # table -> triplets -> local row sums
            if cfg["kind"] == "pairs":
                r, c, cnt, mq = orc.pairs_edges(t["chrom1"], t["chrom2"], t["mapq"], min_mapq=PARAMS["min_mapq"], n_contigs=g.n)
            else:
                r, c, cnt, mq = orc.porec_edges(t["read_idx"], t["contig_idx"], t["mapq"], min_mapq=PARAMS["min_mapq"],
                                                min_order=PARAMS["min_order"], max_order=PARAMS["max_order"], n_contigs=g.n)
            c = c - first if len(c) else c
            H, _, _ = orc.incidence_matrix(r, c, cnt, g.n, int(c.max()) + 1 if len(c) else 1, 0)
            st["H"] = H
            conn.send(np.asarray(H.sum(axis=1)).ravel())
        elif cmd[0] == "phase2":    # global row filter, local column filter, partial A
            keep = cmd[1]
            H = st["H"][keep]
            H = H[:, np.asarray(H.sum(axis=0)).ravel() >= 2]
            A = orc.clique_expansion(H, min_weight=0, dtype=np.float32)
            conn.send(A)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import multiprocessing as mp
    name, cfg = workload(args)
    cores = min(os.cpu_count() or 1, 32)
    per = 400_000
    m = per * cores
    ctx = mp.get_context("spawn")
    procs = []
    for i in range(cores):
        a, b = ctx.Pipe()
        p = ctx.Process(target=_ref_worker, args=(b, cfg, i * per, per), daemon=True)
        p.start()
        procs.append((p, a))
    n_rec = sum(a.recv()[1] for _, a in procs)

    def step():
        for _, a in procs:
            a.send(("phase1",))
        rs = sum(a.recv() for _, a in procs)
        keep = rs >= PARAMS["min_contacts"]
        for _, a in procs:
            a.send(("phase2", keep))
        A = None
        for _, a in procs:
            P = a.recv()
            A = P if A is None else A + P
        A = A.tocsr()
        A.data[A.data < np.float32(PARAMS["min_weight"])] = 0
        A.eliminate_zeros()
        return A
    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        A = step()
    dt = (time.perf_counter() - t0) / args.steps
    for p, a in procs:
        a.send(("stop",))
    v = m / dt
    sample = (f"{m} concatemers ({n_rec} records) of workload {name} per step, {cores} worker processes each running "
              f"oracle/hypergraph_oracle.py on a {per}-concatemer shard (phase 1 triplets+row sums, phase 2 partial A), "
              f"partials summed on rank 0")
    print(json.dumps({"impl": "reference", "metric": "concatemers/sec (H build + H*W*H^T)", "value": v, "unit": "concatemers/s",
                      "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3,
                      "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32 (reference arithmetic)",
                      "data": "synthetic",
                      "config": {"workload": f"{name}: n={cfg['n']} contigs, law={cfg['law']}, kind={cfg['kind']} (bounded sample)",
                                 "params": PARAMS},
                      "cpu_baseline": {"value": v, "unit": "concatemers/s", "cores": cores, "kind": "port", "sample": sample},
                      "e2e": {"value": v, "unit": "concatemers/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def _json_only_stdout():
    """stdout must carry exactly ONE JSON line: libraries that write banners to fd 1 (NCCL's version line, torchrun
    notices) are sent to stderr; the JSON line goes to the saved descriptor."""
    sys.stdout.flush()
    saved = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(saved, "w")


if __name__ == "__main__":
    a = parse()
    _json_only_stdout()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_own(a)
