#!/usr/bin/env python
"""Turns ncu output into the text summaries kept under profiles/ (runs on the CPU box, ncu -i only).

    python scripts/summarize_ncu.py launches gpurun_out/x_launches.csv "<command that was profiled>" > profiles/x_launch_summary.txt
    python scripts/summarize_ncu.py full gpurun_out/x.ncu-rep [max kernels per name] > profiles/x_ncu_full.txt
"""
import collections
import csv
import re
import subprocess
import sys

STEP = ("porec_tile", "incidence_tile", "expand_diag", "row_sums", "epi_", "scan_", "sw_", "bin_", "pairs_")


def short(name):
    name = re.sub(r"\(.*", "", name)
    return name.replace("<unnamed>::", "").replace("void ", "")[:70]


def launches(path, cmd):
    rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"'))]
    hdr = rows[0]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    tot = collections.defaultdict(float)
    cnt = collections.Counter()
    for r in rows[1:]:
        ns = float(r[vi].replace(",", "")) * {"ns": 1, "us": 1e3, "ms": 1e6}.get(r[ui], 1)
        tot[short(r[ki])] += ns / 1e6
        cnt[short(r[ki])] += 1
    step = sum(v for k, v in tot.items() if any(s in k for s in STEP))
    print(f"ncu --metrics gpu__time_duration.sum --clock-control none: {cmd}")
    print("(cold-cache, serialised per-launch times: compare SHARES, not absolutes)")
    print(f"launches {sum(cnt.values())}, total {sum(tot.values()):.2f} ms, kernels of the H/A step {step:.2f} ms\n")
    for k, v in sorted(tot.items(), key=lambda kv: -kv[1])[:40]:
        sh = f"{100 * v / step:5.1f} % of the step kernels" if any(s in k for s in STEP) else ""
        print(f"{v:9.3f} ms {cnt[k]:5d} launches {v / cnt[k]:9.3f} ms each  {k:70s} {sh}")


KEYS = [("gpu__time_duration.sum", "duration"), ("dram__bytes_read.sum", "DRAM read"), ("dram__bytes_write.sum", "DRAM written"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput % of peak"),
        ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2 throughput %"),
        ("l1tex__throughput.avg.pct_of_peak_sustained_active", "L1/TEX (LSU) throughput %"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue slots busy %"),
        ("smsp__thread_inst_executed_per_inst_executed.ratio", "active threads per warp instruction"),
        ("smsp__inst_executed.sum", "warp instructions"), ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
        ("launch__registers_per_thread", "registers/thread"), ("launch__grid_size", "grid"), ("launch__block_size", "block"),
        ("launch__shared_mem_per_block_dynamic", "dynamic smem/block"), ("launch__shared_mem_per_block_static", "static smem/block"),
        ("l1tex__t_sectors_pipe_lsu_mem_global_op_red.sum", "global RED sectors"),
        ("l1tex__t_requests_pipe_lsu_mem_global_op_red.sum", "global RED requests")]


def full(rep, per_name=2):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(l for l in out.splitlines() if l.startswith('"')))
    hdr, units = rows[0], rows[1]
    seen = collections.Counter()
    print(f"ncu --set full --clock-control none, {rep} (one launch per block below)\n")
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        nm = short(d["Kernel Name"])
        seen[nm] += 1
        if seen[nm] > per_name:
            continue
        print(f"== {nm}   (launch id {d.get('ID')})")
        for k, label in KEYS:
            if k in d and d[k] != "":
                print(f"   {label:38s} {d[k]:>18s} {units[hdr.index(k)]}")
        rd, wr = d.get("dram__bytes_read.sum"), d.get("dram__bytes_write.sum")
        print()
    print("kernels captured:", dict(seen))


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else "")
    else:
        full(sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else 2)
