"""diagnostic: variance of the pinned H2D copy alone, the streamed extract alone, and the full e2e step."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from cphasing_b200 import synth, device as D

cfg = dict(synth.CONFIGS["C"])
dev = torch.device("cuda:0")
g, t = bench.gen_table(cfg, 0, dev)
host = {k: torch.empty(v.shape, dtype=v.dtype, pin_memory=True).copy_(v) for k, v in t.items()}
torch.cuda.synchronize()


def timed(f, n=10):
    out = []
    for _ in range(n):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        r = f()
        torch.cuda.synchronize(); out.append(round((time.perf_counter() - t0) * 1e3, 1))
        del r
    return out


print("plain copy 3 columns :", timed(lambda: [t[k].copy_(host[k], non_blocking=True) for k in host]))
for ch in (2 << 20, 8 << 20, 32 << 20):
    D.STREAM_CHUNK_RECORDS = ch
    print(f"streamed extract chunk={ch >> 20}M :", timed(lambda: D.porec_to_edges(host["read_idx"], host["contig_idx"], host["mapq"], g.n)))
D.STREAM_CHUNK_RECORDS = 8 << 20
names = (g.contig_idx, g.contigsizes)
del t
print("e2e step             :", timed(lambda: bench.e2e_step(cfg, g, host, names)))
def slp():
    time.sleep(0.05)
    t0 = time.perf_counter(); bench.e2e_step(cfg, g, host, names); return (time.perf_counter() - t0) * 1e3
print("e2e step after sleep :", [round(slp(), 1) for _ in range(8)])
