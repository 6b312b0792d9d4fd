#!/usr/bin/env python
"""Times the swept expansion (prep pass + one pass per row band) against the direct and the binned kernels on a shard of
workload D / E, for several band sizes (CPH_BAND_MB), and checks that all forms give the same accumulator.  Diagnostic.

    python scripts/diag_swept.py [--workload D] [--mb 32,48,64,96] [--reps 5]
"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench                                         # noqa: E402
from cphasing_b200 import device as D, synth         # noqa: E402
from cphasing_b200._lib import check, lib, ptr, stream   # noqa: E402


def timed(fn, reps):
    ts = []
    for _ in range(reps):
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="D")
    ap.add_argument("--mb", default="32,48,64,96")
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--m", type=int, default=0)
    ap.add_argument("--no-binned", action="store_true")
    a = ap.parse_args()
    cfg = dict(synth.CONFIGS[a.workload]); cfg["m"] = a.m or cfg["m"] // 8
    g, t = bench.gen_table(cfg, 0, "cuda")
    raw = D.porec_to_edges(t["read_idx"], t["contig_idx"], t["mapq"], g.n, min_mapq=1)
    H = D.incidence(raw, min_contacts=bench.PARAMS["min_contacts"], min_mapq=0)
    del t, raw
    n = H.n_rows
    fb = D.frac_bits_for(H.value_bound())
    heads = (H.pem_row[:H.nnz].view(torch.int32) < 0)
    k = torch.bincount(torch.cumsum(heads, 0, dtype=torch.int64) - 1).double()
    pairs = float((k * (k - 1) / 2).sum().item())
    del heads, k
    print(f"workload {a.workload}: n={n} nnz={H.nnz} frac_bits={fb} pairs={pairs:.4g}", flush=True)
    res = {}
    for form in ["direct"] + ([] if a.no_binned else ["binned"]):
        os.environ["CPH_EXPAND_FORM"] = form
        res[form] = D.expand_accumulate(H, use_high_order=True)
        print(f"{form} ms:", round(timed(lambda: D.expand_accumulate(H, use_high_order=True), a.reps), 3), flush=True)
    ref = res["direct"]
    os.environ["CPH_EXPAND_FORM"] = "swept"
    pe4 = torch.empty(H.nnz, dtype=torch.int32, device="cuda")
    flag = torch.empty(1, dtype=torch.int32, device="cuda")
    prep = lambda: check(lib().cph_expand_sweep_prep(ptr(H.pem_row), ptr(H.pem_pk), H.nnz, n, 1, ptr(pe4), ptr(flag), stream()))
    prep()
    print("prep ms:", round(timed(prep, a.reps), 3), "flag", int(flag.item()), flush=True)
    for mb in [float(x) for x in a.mb.split(",")]:
        os.environ["CPH_BAND_MB"] = str(mb)
        r0 = D.sweep_bands(n)
        acc = D.expand_accumulate(H, use_high_order=True)
        same = bool(torch.equal(acc.acc, ref.acc) and torch.equal(acc.diag, ref.diag))
        tt = timed(lambda: D.expand_accumulate(H, use_high_order=True), a.reps)

        def bands():
            for k in range(len(r0) - 1):
                check(lib().cph_expand_sweep_band(ptr(pe4), H.nnz, n, r0[k], r0[k + 1], fb, 1, ptr(acc.acc), ptr(acc.diag), stream()))
        tb = timed(bands, a.reps)
        print(f"band {mb} MB: {len(r0) - 1} bands, whole call {tt:.2f} ms (bands only {tb:.2f} ms = {pairs / tb / 1e6:.1f} G pairs/s), "
              f"identical={same}", flush=True)
        del acc


if __name__ == "__main__":
    main()
