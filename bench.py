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
NCU_TRAFFIC = {("C", 50_000_000): 2724058000 + 2344288000}      # profiles/r2w_ncu_full_C.txt (porec_tile_kernel<1, 8>)
PARAMS = dict(min_mapq=1, min_order=2, max_order=50, min_contacts=25, min_weight=2.0)   # cphasing/_config.py:38-45


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=None, help="A|B|C|D|E (default C; every GPU takes a full-size shard, D and E: 1/8 of the named size)")
    ap.add_argument("--m", type=int, default=None, help="override the number of concatemers per GPU")
    ap.add_argument("--cpu-sample", type=int, default=10_000_000, help="concatemers in the CPU-baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-partition", action="store_true", help="skip the kprune + clustering tail (hyperpartition wall s)")
    ap.add_argument("--no-extra", action="store_true", help="skip the extra workloads (D and E shards) after the headline one")
    ap.add_argument("--extra", default="B,Cf,D,E", help="comma-separated extra workloads: D and E at 1/8 of the named size per GPU, "
                                                        "B (Hi-C pairs) at its full 100 M pairs per GPU, Cf = C with the "
                                                        "reference's default edge filter (--edge-length 5m, 17 B records)")
    ap.add_argument("--extra-parity-sample", type=int, default=2_000_000,
                    help="concatemers of each extra workload that the CPU path recomputes for the parity check (E: half of it)")
    ap.add_argument("--extra-steps", type=int, default=5)
    return ap.parse_args()


def workload(args):
    from cphasing_b200 import synth
    name = args.workload or "C"
    cfg = dict(synth.CONFIGS[name])
    if name in ("D", "E"):
        cfg["m"] = cfg["m"] // 8                 # per GPU: D 62.5M (the named 500M at 8 GPUs), E 25M (200M at 8 GPUs)
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
        t = synth.porec_table_device(g, cfg["m"], cfg["seed"], law=cfg["law"], first=first, device=dev,
                                     positions=bool(cfg.get("edge_length")))
    return g, t


_LEN_DEV = {}


def edge_filter_args(cfg, g, t, dev=None):
    """keyword arguments of the reference's edge filter (hypergraph.py:1438-1457; `--edge-length 5m` is the default of
    `cphasing hyperpartition`, _config.py:37) for D.porec_to_edges / the oracle: {} when the config has it off"""
    el = int(cfg.get("edge_length") or 0)
    if not el:
        return {}
    if dev is None:
        return dict(start=t["start"], end=t["end"], contig_len=g.length, edge_length=el)
    key = (id(g), str(dev))
    if key not in _LEN_DEV:
        import torch
        _LEN_DEV.clear()
        _LEN_DEV[key] = torch.from_numpy(np.ascontiguousarray(g.length.astype(np.int64))).to(dev)
    return dict(start=t["start"], end=t["end"], contig_len=_LEN_DEV[key], edge_length=el)


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
                               min_order=PARAMS["min_order"], max_order=PARAMS["max_order"],
                               **edge_filter_args(cfg, g, t, t["mapq"].device))
    mark("edges")
    from cphasing_b200 import parallel as par
    # exchange 1 (world > 1): global per-contig contact sums + largest count for the min_contacts row filter
    H = D.incidence(raw, min_contacts=PARAMS["min_contacts"], min_mapq=0,
                    rowsum_hook=par.allreduce_rowsums if world > 1 else None)
    mark("incidence")
    # the fixed-point scale derives from the (all-reduced) row sums, so every rank picks the same one.  Exchange 2
    # (world > 1): the partial accumulators are summed over the ranks (NCCL) -- band by band under the kernels of the next
    # band when the triangle is swept (n >= 5.6k), so "exchange" is what stays exposed after the last local kernel.
    acc = par.expand_accumulate_sum(H, use_high_order=True, on_local_done=lambda: mark("accumulate"))
    if world > 1:
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
            accu = par.expand_accumulate_sum(H, use_high_order=False)
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


def reference_module():
    """the UNMODIFIED reference's ``cphasing.algorithms.hypergraph`` (baseline/_ref, polars/igraph/... stubbed -- see
    baseline/ref_loader.py) or None when the install did not travel."""
    try:
        sys.path.insert(0, os.path.join(ROOT, "baseline"))
        import ref_loader
        if not ref_loader.available():
            return None
        import logging
        ah = ref_loader.load()
        logging.getLogger().setLevel(logging.ERROR)
        logging.getLogger("cphasing").setLevel(logging.ERROR)
        return ah
    except Exception as e:                                   # noqa: BLE001
        print(f"[bench] reference import failed ({e!r}); CPU legs fall back to the oracle port", file=sys.stderr)
        return None


def cpu_path(cfg, g, t, ah, counts=False, high_order=True):
    """the reference's CPU path on a host table `t` (numpy columns): table -> triplets (oracle port of the polars
    group-by, hypergraph.py:1395-1481 -- polars is not installable) -> HyperGraph.incidence_matrix(min_contacts) ->
    clique_expansion_init(min_weight) with the REFERENCE's own classes when `ah` is given, else the oracle's
    restatement.  -> (H, A, seconds per stage)"""
    from oracle import hypergraph_oracle as orc
    tm = {}
    t0 = time.perf_counter()
    if cfg["kind"] == "pairs":
        r, c, cnt, mq = orc.pairs_edges(t["chrom1"], t["chrom2"], t["mapq"], min_mapq=PARAMS["min_mapq"], n_contigs=g.n)
    else:
        r, c, cnt, mq = orc.porec_edges(t["read_idx"], t["contig_idx"], t["mapq"], min_mapq=PARAMS["min_mapq"],
                                        min_order=PARAMS["min_order"], max_order=PARAMS["max_order"], n_contigs=g.n,
                                        **edge_filter_args(cfg, g, t))
    tm["triplets"] = time.perf_counter() - t0; t0 = time.perf_counter()
    if ah is not None:
        edges = ah.HyperEdges(idx=g.contig_idx, row=r, col=c, count=cnt, contigsizes=g.contigsizes, mapq=mq)
        HG = ah.HyperGraph(edges, min_quality=PARAMS["min_mapq"], mapq_filter=False)     # hyperpartition.py:249
        H = HG.incidence_matrix(min_contacts=PARAMS["min_contacts"])
        tm["incidence"] = time.perf_counter() - t0; t0 = time.perf_counter()
        A = ah.HyperGraph.clique_expansion_init(H, min_weight=PARAMS["min_weight"], use_high_order=high_order)
    else:
        rr, cc, kk, ncol = orc.hypergraph_get_data(r, c, cnt, mq, PARAMS["min_mapq"], mapq_filter=False)
        H, _, _ = orc.incidence_matrix(rr, cc, kk, g.n, ncol, PARAMS["min_contacts"])
        tm["incidence"] = time.perf_counter() - t0; t0 = time.perf_counter()
        A = orc.clique_expansion(H, min_weight=PARAMS["min_weight"], dtype=np.float32, use_high_order=high_order)
    tm["expansion"] = time.perf_counter() - t0
    return H, A, tm, len(r)


def sample_of(cfg, t, n_sample):
    """first n_sample concatemers of the device table -> (host numpy columns, device slices, k)"""
    import torch
    k = min(n_sample, cfg["m"])
    if cfg["kind"] == "pairs":
        cut = k
    else:
        rid = t["read_idx"].view(torch.int32)
        first = int(rid[0].item())
        cut = int(torch.searchsorted(rid, torch.tensor([first + k], dtype=torch.int32, device=rid.device)).item())
    dev = {kk: v[:cut] for kk, v in t.items()}
    host = {kk: v.cpu().numpy() for kk, v in dev.items()}
    return host, dev, k


def cpu_baseline(cfg, g, host, k, ah):
    """the reference's CPU path on the first k concatemers of the benched table, one core (scipy's csr_matmat is
    single-threaded; MKL and polars are absent)."""
    H, A, tm, ntrip = cpu_path(cfg, g, host, ah)
    dt = sum(tm.values())
    kind = "reference" if ah is not None else "port"
    what = ("the reference's own HyperGraph.incidence_matrix + clique_expansion_init (baseline/_ref, scipy float32) after "
            "the oracle port of its polars group-by (polars not installable)") if ah is not None else \
           "oracle/hypergraph_oracle.py (numpy+scipy restatement; the reference install is absent)"
    return {"value": k / dt, "unit": "concatemers/s", "cores": 1, "kind": kind,
            "seconds": {a: round(b, 3) for a, b in tm.items()},
            "sample": f"first {k} concatemers of the same table ({ntrip} H entries): {what}, {dt:.1f} s"}, H, A


def _canon(M):
    M = M.tocsr().copy()
    M.sort_indices()
    return M


def parity_check(D, cfg, g, host, dev_t, H_cpu, A_cpu, ah):
    """the CUDA path on the SAME sample the CPU baseline just computed, compared output by output:
    H and the integer contact-count matrix bit for bit, the weighted A against the fp64 oracle (1e-6, north_star) and
    against the reference's float32 result (pattern; values to float32 accumulation error)."""
    from scipy.sparse import csr_matrix
    from oracle import hypergraph_oracle as orc
    if cfg["kind"] == "pairs":
        raw = D.pairs_to_edges(dev_t["chrom1"], dev_t["chrom2"], dev_t["mapq"], g.n, min_mapq=PARAMS["min_mapq"])
    else:
        raw = D.porec_to_edges(dev_t["read_idx"], dev_t["contig_idx"], dev_t["mapq"], g.n, min_mapq=PARAMS["min_mapq"],
                               min_order=PARAMS["min_order"], max_order=PARAMS["max_order"],
                               **edge_filter_args(cfg, g, dev_t, dev_t["mapq"].device))
    H = D.incidence(raw, min_contacts=PARAMS["min_contacts"], min_mapq=0)
    Hs = _canon(csr_matrix(((D.to_host(H.pk).view(np.uint32) & 0xFFFFFF).astype(np.float32), D.to_host(H.indices),
                            D.to_host(H.indptr)), shape=(H.n_rows, H.n_cols)))
    Hc = _canon(H_cpu)
    out = {"sample_concatemers": None, "H_shape": list(Hs.shape), "H_nnz": int(Hs.nnz)}
    out["expansion_form"] = "swept row bands (sw_prep + sw_band)" if (H.n_rows >= D.BINNED_MIN_N and H.nnz >= D.BINNED_MIN_NNZ) \
        else "direct (expand_diag_kernel)"
    out["H_bit_exact"] = bool(Hs.shape == Hc.shape and np.array_equal(Hs.indptr, Hc.indptr)
                              and np.array_equal(Hs.indices, Hc.indices) and np.array_equal(Hs.data, Hc.data))
    # integer contact counts (use_high_order=False): bit-exact against the CPU integer matrix
    Au, acc = D.clique_expansion(H, use_high_order=False, min_weight=PARAMS["min_weight"], want32=False, want64=True)
    Aus = _canon(csr_matrix((D.to_host(Au.data64), D.to_host(Au.indices), D.to_host(Au.indptr)), shape=(H.n_rows,) * 2))
    if ah is not None:
        Auc = ah.HyperGraph.clique_expansion_init(H_cpu, min_weight=PARAMS["min_weight"], use_high_order=False)
    else:
        Auc = orc.clique_expansion(H_cpu, use_high_order=False, min_weight=PARAMS["min_weight"])
    Auc = _canon(Auc)
    out["counts_bit_exact"] = bool(acc.frac_bits == 0 and np.array_equal(Aus.indptr, Auc.indptr)
                                   and np.array_equal(Aus.indices, Auc.indices)
                                   and np.array_equal(Aus.data.astype(np.int64), np.rint(Auc.data).astype(np.int64)))
    out["counts_nnz"] = int(Aus.nnz)
    del Au, acc
    # weighted A
    A, _ = D.clique_expansion(H, use_high_order=True, min_weight=PARAMS["min_weight"], want32=False, want64=True)
    As = _canon(csr_matrix((D.to_host(A.data64), D.to_host(A.indices), D.to_host(A.indptr)), shape=(H.n_rows,) * 2))
    A64 = _canon(orc.clique_expansion(Hc.astype(np.float64), min_weight=PARAMS["min_weight"], dtype=np.float64))
    thr = PARAMS["min_weight"]

    def compare(X, Y, near):
        """-> (pattern_equal, entries present on one side only that are NOT within `near` of the threshold, max rel)"""
        same = X.nnz == Y.nnz and np.array_equal(X.indptr, Y.indptr) and np.array_equal(X.indices, Y.indices)
        if same:
            rel = float(np.max(np.abs(X.data - Y.data) / np.maximum(np.abs(Y.data), 1e-300))) if X.nnz else 0.0
            return True, 0, rel
        d = (X - Y).tocsr()
        Xp = X.copy(); Xp.data[:] = 1
        Yp = Y.copy(); Yp.data[:] = 1
        only = (Xp - Yp).tocsr(); only.eliminate_zeros()       # +1: only in X, -1: only in Y
        oc = only.tocoo()
        vals = np.where(oc.data > 0, np.asarray(X[oc.row, oc.col]).ravel(), np.asarray(Y[oc.row, oc.col]).ravel())
        bad = int(np.sum(np.abs(vals - thr) > near * thr))
        both = Xp.multiply(Yp).tocsr()
        Xi = X.multiply(both).tocsr(); Yi = Y.multiply(both).tocsr()
        Xi.sort_indices(); Yi.sort_indices()
        rel = float(np.max(np.abs(Xi.data - Yi.data) / np.maximum(np.abs(Yi.data), 1e-300))) if Xi.nnz else 0.0
        return False, bad, rel
    pe, bad, rel = compare(As, A64, 1e-9)
    out.update(weighted_nnz=int(As.nnz), pattern_equal=bool(pe), weighted_max_rel=rel, weighted_tol=1e-6,
               weighted_against="oracle fp64 restatement of algorithms/hypergraph.py:372-455 on the CPU H")
    pe32, bad32, rel32 = compare(As, _canon(A_cpu).astype(np.float64), 1e-4)
    out.update(vs_cpu_float32={"pattern_equal": bool(pe32), "entries_off_threshold": bad32, "max_rel": rel32,
                               "nnz_cpu": int(A_cpu.nnz),
                               "note": "the CPU side multiplies sqrt-scaled float32 operands and accumulates in float32 "
                                       "(scipy csr_matmat, algorithms/hypergraph.py:418-446): cells whose exact value equals "
                                       "min_weight fall on either side of the cut by rounding, and long sums carry ~1e-4 "
                                       "relative float32 error; informational except entries_off_threshold"})
    out["ok"] = bool(out["H_bit_exact"] and out["counts_bit_exact"] and pe and bad == 0 and rel <= 1e-6 and bad32 == 0
                     and rel32 <= 1e-3)
    return out


def timed_run(D, _lib, name, cfg, g, t, steps, warmup, world, rank, local, dev, barrier, clocks_on=True):
    """W untimed + K timed passes of the device-resident step; CUDA events, max over ranks -> (json fields, raw, H, A)"""
    import torch
    import torch.distributed as dist
    n_rec = int(next(iter(t.values())).numel())
    for _ in range(max(warmup, 3)):
        ev = []
        raw, H, A = device_step(D, cfg, g, t, ev, world, rank)
    barrier()
    clocks = Clocks(local)
    if rank == 0 and clocks_on:
        clocks.start()
    l0 = _lib.launches()
    all_ev = []
    t_start = torch.cuda.Event(enable_timing=True); t_end = torch.cuda.Event(enable_timing=True)
    barrier()
    t_start.record()
    for _ in range(steps):
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
    clk = clocks.stop() if (rank == 0 and clocks_on) else None
    ms_step = ms_total / steps
    stages = {}
    for ev in all_ev:
        for (n0, e0), (n1, e1) in zip(ev[:-1], ev[1:]):
            stages.setdefault(n1, []).append(e0.elapsed_time(e1))
    stages = {k: sum(v) / len(v) for k, v in stages.items()}

    # roofline of the DOMINANT kernel (profiles/: the group-by kernel of cph_porec_to_edges on Pore-C tables; on pairs
    # tables and at high order the expansion kernel).  ALGORITHMIC bytes:
    #   group-by kernel: 9 B per alignment record read + 8 B per surviving (read, contig) entry written
    #   expansion kernel: 8 B per H entry read (row + packed count/degree); its work figure is the pair updates
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
    rec_b = 17 if cfg.get("edge_length") else 9        # + start, end (u32 each) when the edge filter reads them
    stage1_bytes = rec_b * n_rec + 8 * raw.nnz + 8 * (g.n + 1)
    stage2_bytes = 8 * nnzH + 4 * H.n_cols + 12 * nnzA + 16 * (n_rows + 1)
    s1_ms = stages.get("edges", 0) + stages.get("incidence", 0)
    s2_ms = stages.get("accumulate", 0) + stages.get("emit", 0)
    xname = "expansion kernel(s) (cph_expand_accumulate_*)"
    per_kernel = {xname: {"bytes": 8 * nnzH, "ms": acc_ms, "GBps": 8 * nnzH / acc_ms / 1e6, "frac": 8 * nnzH / acc_ms / 1e6 / peak}}
    gname = "porec_tile_kernel (cph_porec_to_edges)" if cfg["kind"] != "pairs" else "pairs_flag/emit_kernel (cph_pairs_to_edges)"
    per_kernel[gname] = {"bytes": rec_b * n_rec + 8 * raw.nnz, "ms": edges_ms, "GBps": (rec_b * n_rec + 8 * raw.nnz) / edges_ms / 1e6,
                         "frac": (rec_b * n_rec + 8 * raw.nnz) / edges_ms / 1e6 / peak}
    traffic = None
    if cfg["kind"] == "pairs" or acc_ms > edges_ms:
        kname, kbytes, kms = xname, 8 * nnzH, acc_ms
    else:
        kname, kbytes, kms = gname, rec_b * n_rec + 8 * raw.nnz, edges_ms
        # ncu --set full, one launch on this workload: dram__bytes_read + dram__bytes_write (profiles/)
        traffic = NCU_TRAFFIC.get((name, cfg["m"]))
    roof = {"bound": "hbm", "kernel": kname, "achieved": kbytes / kms / 1e6, "peak": peak, "unit": "GB/s",
            "frac": kbytes / kms / 1e6 / peak, "traffic": traffic,
            "peak_source": "MEASURED_PEAKS.json hbm_gbs (measured copy)" if peaks else "fallback 6650",
            "algorithmic_bytes_per_launch": kbytes, "kernel_ms": kms, "kernels": per_kernel,
            "stage1_GBps": stage1_bytes / s1_ms / 1e6 if s1_ms else None,
            "stage2_GBps": stage2_bytes / s2_ms / 1e6 if s2_ms else None}
    out = {"metric": "concatemers/sec (H build + H*W*H^T)", "value": cfg["m"] * world / (ms_step / 1e3),
           "unit": "concatemers/s", "n_gpus": world, "steps": steps, "warmup": max(warmup, 3),
           "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64 fixed-point / f64",
           "data": "synthetic", "gpu_launches": launches,
           "config": {"workload": f"{name}: n={cfg['n']} contigs, m={cfg['m']} concatemers/GPU x {world} GPU, law={cfg['law']}, "
                                  f"kind={cfg['kind']}" + (f", edge_length={cfg['edge_length']}" if cfg.get("edge_length") else ""),
                      "records_per_gpu": n_rec, "nnz_H": nnzH, "nnz_A": nnzA,
                      "params": PARAMS, "l2": "inputs (>2 GB) and intermediates exceed the 126 MB L2; no flush needed",
                      "parallelism": f"concatemer-column shards x{world}" if world > 1 else "single GPU"},
           "stages_ms": stages, "roofline": roof, "clocks": clk}
    return out, raw, H, A


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
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    out, raw, H, A = timed_run(D, _lib, name, cfg, g, t, args.steps, args.warmup, world, rank, local, dev, barrier)
    ms_step = out["ms_per_step"]
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
    if world > 1 and not args.no_e2e and cfg["kind"] == "porec":
        # every rank feeds ITS shard from pinned host memory through the public classes in column-shard mode
        # (HyperExtractor -> HyperGraph(column_shard) .incidence_matrix -> clique_expansion_init: chunked H2D under the
        # group-by kernel, row-sum all-reduce, the accumulator exchange, CSR emit) and rank 0 reads A back; max over ranks.
        # The process is first bound to the NUMA node of its GPU so that the pinned staging buffers sit next to it.
        import torch
        from cphasing_b200 import device as Dm, hypergraph as hg
        from cphasing_b200.algorithms import hypergraph as ah
        node = Dm.bind_host_to_gpu_node(local)
        host = {k: torch.empty(v.shape, dtype=v.dtype, pin_memory=True).copy_(v) for k, v in t.items()}
        h2d = sum(v.numel() * v.element_size() for v in host.values())
        del t, raw, H, A
        _lib.release_workspace(); torch.cuda.empty_cache()

        def e2e_multi():
            ex = hg.HyperExtractor.from_arrays(host["read_idx"], host["contig_idx"], host["mapq"], g.contig_idx, g.contigsizes,
                                               min_quality=PARAMS["min_mapq"], min_order=PARAMS["min_order"],
                                               max_order=PARAMS["max_order"], edge_length=0)
            HG = ah.HyperGraph(ex.edges, min_quality=PARAMS["min_mapq"], mapq_filter=False)
            HG.column_shard = True
            Hs = HG.incidence_matrix(PARAMS["min_contacts"])
            A2 = ah.HyperGraph.clique_expansion_init(Hs, min_weight=PARAMS["min_weight"])
            nb = 0
            if rank == 0:
                nb = A2.data.nbytes + A2.indices.nbytes + A2.indptr.nbytes     # touching .data materialises the scipy CSR
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
                      "host_numa_node_rank0": node, "d2h_bytes_per_step": d2h, "ms_per_step": dt * 1e3,
                      "api": "per rank: HyperExtractor.from_arrays(host shard) -> HyperGraph(column_shard=True)"
                             ".incidence_matrix (row-sum all-reduce) -> HyperGraph.clique_expansion_init (accumulator "
                             "exchange over NCCL) ; rank 0 reads the scipy CSR of A back"}
    parity_ok = True
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        ah = reference_module()
        host_s, dev_s, k = sample_of(cfg, t, args.cpu_sample)
        cb, Hc, Ac = cpu_baseline(cfg, g, host_s, k, ah)
        out["cpu_baseline"] = cb
        # parity INSIDE the bench: the CUDA path on the sample the CPU leg just computed (rc != 0 on mismatch)
        par_ = parity_check(D, cfg, g, host_s, dev_s, Hc, Ac, ah)
        par_["sample_concatemers"] = k
        par_["cpu_side"] = cb["kind"]
        out["parity"] = par_
        parity_ok = par_["ok"]
    # ---- extra workloads, driver-timed next to the headline one: B (Hi-C pairs: n = 1k, 100 M pairs per GPU), D
    # (north_star's 8-GPU config: n = 20k, 62.5 M concatemers per GPU = 500 M at N = 8) and E (high-order HiFi-C: n = 10k,
    # order <= 50, 25 M per GPU = 200 M at N = 8), each with its own parity check against the CPU path at N = 1
    if not args.no_extra and args.workload is None and not args.m:
        import gc
        from cphasing_b200 import synth
        raw = H = A = t = host = host_s = dev_s = Hc = Ac = Ah = None       # drop the headline workload's buffers
        gc.collect()
        _lib.release_workspace(); torch.cuda.empty_cache()
        extra = {}
        for xn in args.extra.split(","):
            xcfg = dict(synth.CONFIGS["C" if xn == "Cf" else xn])
            if xn == "Cf":                 # workload C with the reference's DEFAULT extraction filter (--edge-length 5m)
                xcfg["edge_length"] = 5_000_000
            if xn in ("D", "E"):
                xcfg["m"] = xcfg["m"] // 8
            try:
                xg, xt = gen_table(xcfg, rank, dev)
                xo, xraw, xH, xA = timed_run(D, _lib, xn, xcfg, xg, xt, args.extra_steps, 3, world, rank, local, dev, barrier,
                                             clocks_on=False)
                del xraw
                if rank == 0 and world == 1 and not args.no_cpu_baseline:
                    # parity at the benched shape: the CPU path recomputes a sample of THIS table (n = 20k swept bands
                    # for D, orders up to 50 for E, duplicated intra-contig pairs for B); rc != 0 on mismatch as for C
                    ah_x = reference_module()
                    ks = args.extra_parity_sample // 2 if xn == "E" else args.extra_parity_sample
                    xhost, xdev, xk = sample_of(xcfg, xt, ks)
                    xcb, xHc, xAc = cpu_baseline(xcfg, xg, xhost, xk, ah_x)
                    xpar = parity_check(D, xcfg, xg, xhost, xdev, xHc, xAc, ah_x)
                    xpar["sample_concatemers"] = xk
                    xpar["cpu_side"] = xcb["kind"]
                    xo["cpu_baseline"] = xcb
                    xo["parity"] = xpar
                    parity_ok = parity_ok and xpar["ok"]
                    del xhost, xdev, xHc, xAc
                del xt
                if not args.no_partition and xcfg["kind"] != "pairs" and xn != "Cf":   # tail: Pore-C configs (Cf would repeat C's)
                    tail = partition_tail(D, xcfg, xg, xH, xA, reps=1, world=world, rank=rank)
                    barrier()
                    if rank == 0:
                        tail["wall_s"] = (xo["ms_per_step"] + tail["kprune_ms"] + tail["cluster_ms"]) / 1e3
                        xo["hyperpartition"] = tail
                for k_ in ("metric", "unit", "higher_is_better", "scaling", "vs_baseline", "data", "clocks", "warmup"):
                    xo.pop(k_, None)
                extra[xn] = xo
                del xH, xA, xg
            except Exception as e:                                   # noqa: BLE001
                extra[xn] = {"error": repr(e)[:300]}
            gc.collect()
            _lib.release_workspace(); torch.cuda.empty_cache()
        out["extra"] = extra
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if not parity_ok:
        print("[bench] PARITY MISMATCH between the CUDA path and the CPU path on the sample", file=sys.stderr)
        sys.exit(3)


# ------------------------------------------------------------------------------------------------
# reference arm: the reference's own CPU implementation of the path on the host cores.  The reference is pure Python
# over polars / scipy / igraph; it is installed UNMODIFIED under baseline/_ref (pip --no-deps) and its
# HyperEdges / HyperGraph.incidence_matrix / clique_expansion_init run as they are (scipy float32, single-threaded
# csr_matmat -- MKL absent).  The one piece that cannot run is the polars group-by of process_pore_c_table
# (hypergraph.py:1395-1481; polars is not in the wheelhouse): the oracle's numpy port stands in for it.  The reference
# has no multi-core mode for this path without polars/MKL, so "all the host threads" = one independent replica of the
# stock single-process path per core, each on its own table of the workload's shape; value = concatemers of all
# replicas / wall time.  The same fixed sample (REF_SAMPLE concatemers per replica) is used at every N.
# ------------------------------------------------------------------------------------------------
REF_SAMPLE = int(os.environ.get("CPH_REF_SAMPLE", 1_000_000))


def _ref_worker(conn, cfg, first, m):
    from cphasing_b200 import synth
    ah = reference_module()
    g = synth.make_genome(cfg["n"], cfg["chromosomes"], cfg["ploidy"], cfg["seed"])
    if cfg["kind"] == "pairs":
        t = synth.pairs_table(g, m, cfg["seed"], first=first)
    else:
        t = synth.porec_table(g, m, cfg["seed"], law=cfg["law"], first=first)
        t["read_idx"] = t["read_idx"] - np.uint32(first) if first else t["read_idx"]
    conn.send(("ready", len(next(iter(t.values()))), ah is not None))
    while True:
        cmd = conn.recv()
        if cmd[0] == "stop":
            break
        H, A, tm, ntrip = cpu_path(cfg, g, t, ah)
        conn.send((tm, int(H.nnz), int(A.nnz)))


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import multiprocessing as mp
    name, cfg = workload(args)
    try:
        cores = len(os.sched_getaffinity(0))
    except AttributeError:
        cores = os.cpu_count() or 1
    cores = max(1, min(cores, 64))
    per = REF_SAMPLE
    m = per * cores
    ctx = mp.get_context("spawn")
    procs = []
    for i in range(cores):
        a, b = ctx.Pipe()
        p = ctx.Process(target=_ref_worker, args=(b, cfg, i * per, per), daemon=True)
        p.start()
        procs.append((p, a))
    ready = [a.recv() for _, a in procs]
    n_rec = sum(r[1] for r in ready)
    kind = "reference" if all(r[2] for r in ready) else "port"

    def step():
        for _, a in procs:
            a.send(("step",))
        return [a.recv() for _, a in procs]
    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        res = step()
    dt = (time.perf_counter() - t0) / args.steps
    for p, a in procs:
        a.send(("stop",))
    v = m / dt
    stage_s = {k_: round(float(np.mean([r[0][k_] for r in res])), 3) for k_ in res[0][0]}
    for p, _ in procs:
        p.join(30)
    # The timed replicas above hold REF_SAMPLE concatemers each, not the 50 M of the named table (hours of CPU time).  To show
    # what that does to the number, time ONE step of the same path on tables of 1 M, 4 M and 10 M concatemers per replica at
    # a fixed concurrency of 4 replicas (N = 1 arm only: the other arms of the scaling run would repeat the same figures).
    size_sweep = None
    if args.gpus == 1 and not os.environ.get("CPH_REF_NO_SWEEP"):
        size_sweep = {"replicas": min(4, cores), "unit": "concatemers/s per replica", "per_replica_concatemers": {}}
        for per_s in (1_000_000, 4_000_000, 10_000_000):
            ws_ = []
            for i in range(size_sweep["replicas"]):
                a, b = ctx.Pipe()
                p = ctx.Process(target=_ref_worker, args=(b, cfg, i * per_s, per_s), daemon=True)
                p.start()
                ws_.append((p, a))
            for _, a in ws_:
                a.recv()
            t1 = time.perf_counter()
            for _, a in ws_:
                a.send(("step",))
            for _, a in ws_:
                a.recv()
            d1 = time.perf_counter() - t1
            for p, a in ws_:
                a.send(("stop",))
                p.join(30)
            size_sweep["per_replica_concatemers"][str(per_s)] = round(per_s / d1, 1)
    what = ("the UNMODIFIED reference (baseline/_ref): HyperEdges -> HyperGraph(mapq_filter=False).incidence_matrix(25) -> "
            "clique_expansion_init(min_weight=2.0), float32 scipy; its polars group-by replaced by the oracle's numpy port "
            "(polars not installable)") if kind == "reference" else "oracle/hypergraph_oracle.py port (reference install absent)"
    sample = (f"{cores} independent replicas (one per core; the path is single-threaded without polars/MKL), each "
              f"{per} concatemers of workload {name} per step ({n_rec} records in all): {what}; mean seconds per stage "
              f"per replica {stage_s}")
    print(json.dumps({"impl": "reference", "metric": "concatemers/sec (H build + H*W*H^T)", "value": v, "unit": "concatemers/s",
                      "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3,
                      "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32 (reference arithmetic)",
                      "data": "synthetic",
                      "config": {"workload": f"{name}: n={cfg['n']} contigs, law={cfg['law']}, kind={cfg['kind']} "
                                             f"(bounded sample: {per} concatemers x {cores} replicas per step)",
                                 "params": PARAMS},
                      "cpu_baseline": {"value": v, "unit": "concatemers/s", "cores": cores, "kind": kind, "sample": sample,
                                       "size_sweep": size_sweep},
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
