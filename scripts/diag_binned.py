#!/usr/bin/env python
"""Times the three passes of the binned expansion apart (count / fill / apply) against the direct kernel on the
workload-D shard, for several block sizes (CPH_XB_BS).  Diagnostic only.

    python scripts/diag_binned.py [--bs 887,1774,2508] [--reps 5]
"""
import argparse
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench                                         # noqa: E402
from cphasing_b200 import _lib, device as D, synth   # noqa: E402
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
    ap.add_argument("--bs", default="887,1774,2508")
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--m", type=int, default=0)
    a = ap.parse_args()
    cfg = dict(synth.CONFIGS["D"]); cfg["m"] = a.m or cfg["m"] // 8
    g, t = bench.gen_table(cfg, 0, "cuda")
    raw = D.porec_to_edges(t["read_idx"], t["contig_idx"], t["mapq"], g.n, min_mapq=1)
    H = D.incidence(raw, min_contacts=bench.PARAMS["min_contacts"], min_mapq=0)
    del t, raw
    n = H.n_rows
    fb = D.frac_bits_for(H.value_bound())
    tri = n * (n + 1) // 2
    acc = torch.empty(tri, dtype=torch.int64, device="cuda")
    diag = torch.empty(n, dtype=torch.int64, device="cuda")
    print(f"n={n} nnz={H.nnz} frac_bits={fb}", flush=True)
    fn = lib().cph_expand_accumulate_binned
    args = (ptr(H.pem_row), ptr(H.pem_pk), H.nnz, n, 1, fb, 1, ptr(acc), ptr(diag))
    old = D.BINNED_MIN_N
    D.BINNED_MIN_N = 1 << 30
    print("direct ms:", timed(lambda: D.expand_accumulate(H, use_high_order=True), a.reps), flush=True)
    D.BINNED_MIN_N = old
    for bs in [int(x) for x in a.bs.split(",")]:
        os.environ["CPH_XB_BS"] = str(bs)
        need = C.c_size_t(0); hp = C.c_uint64(0)
        check(fn(1, *args, None, C.byref(hp), None, 0, C.byref(need), None))
        ws = torch.empty(need.value, dtype=torch.uint8, device="cuda")
        wsa = (ws.data_ptr(), ws.numel(), C.byref(need), stream())
        p1 = lambda: check(fn(1, *args, None, C.byref(hp), *wsa))
        p1()
        rec = torch.empty(max(hp.value, 1), dtype=torch.int64, device="cuda")
        t1 = timed(p1, a.reps)
        t3 = timed(lambda: check(fn(3, *args, ptr(rec), C.byref(hp), *wsa)), a.reps)
        t4 = timed(lambda: check(fn(4, *args, ptr(rec), C.byref(hp), *wsa)), a.reps)
        print(f"bs={bs}: pairs={hp.value} count+zero {t1:.2f} ms  fill {t3:.2f} ms  apply {t4:.2f} ms  sum {t1+t3+t4:.2f}", flush=True)
        del rec, ws


if __name__ == "__main__":
    main()
