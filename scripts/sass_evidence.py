#!/usr/bin/env python
"""cuobjdump -sass of the shipped library -> profiles/r2_sass_mnemonics.txt: static instruction counts per kernel for the
mnemonics that prove what the kernels use (TMA bulk copies, mbarriers, reductions, shared-memory atomics, match.any)."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "cphasing_b200", "lib", "libcph.so")
sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
cur, counts = None, collections.defaultdict(collections.Counter)
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        continue
    m = re.search(r"^\s+/\*[0-9a-f]+\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and cur:
        counts[cur][m.group(2).split(".")[0]] += 1
names = list(counts)
dem = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.split("\n")


def clean(n):
    return re.sub(r"\(.*", "", n.replace("(anonymous namespace)::", "").replace("void ", ""))


want = ["UBLKCP", "SYNCS", "REDG", "ATOMG", "ATOMS", "MATCH", "VOTE", "REDUX", "LDG", "STG"]
out = ["cuobjdump -sass cphasing_b200/lib/libcph.so  (sm_100a only): static instruction counts per kernel",
       "UBLKCP = cp.async.bulk (TMA bulk copy), SYNCS = mbarrier ops, REDG = global reduction without return, ATOMG = with return,",
       "ATOMS = shared-memory atomic, MATCH = match.any.  No UTC*MMA / HMMA anywhere: the path has no contraction (SURVEY.md 8d).", ""]
rows = []
for k, d in zip(names, dem):
    c = counts[k]
    sel = {w: c[w] for w in want if c.get(w)}
    if any(w in sel for w in ("UBLKCP", "SYNCS", "REDG", "ATOMG", "ATOMS", "MATCH")):
        rows.append(f"{clean(d)[:58]:58s} total {sum(c.values()):6d}  " + "  ".join(f"{a}={b}" for a, b in sel.items()))
out += sorted(rows)
tot = collections.Counter()
for c in counts.values():
    tot.update(c)
out += ["", "whole library: " + "  ".join(f"{w}={tot[w]}" for w in want + ["UTCHMMA", "HMMA", "UTMALDG"])]
open(os.path.join(ROOT, "profiles", "r2_sass_mnemonics.txt"), "w").write("\n".join(out) + "\n")
print("\n".join(out[-12:]))
