"""Multi-GPU form of stages 1-2: one process per GPU, concatemer-column shards, NCCL through
``torch.distributed``.

Why it shards: W is diagonal and column-local (``D_e`` depends only on the hyperedge's own entries,
algorithms/hypergraph.py:385), so with rank r owning a contiguous range of concatemer ids

    A = sum_r  H_r W_r H_r^T .

Exchanges per pass (everything else is local):
  1. all-reduce of the per-contig contact sums (n x int64) and of the largest count (1 x int32): the ``min_contacts`` row
     filter (algorithms/hypergraph.py:286-300) is global, and every rank must pick the same fixed-point scale;
     the column filter (sum of counts >= 2, :305-306) is column-local and needs nothing.
  2. sum of the partial accumulators (packed upper triangle, 64-bit fixed point, integer => order independent and
     bit-identical on every rank).  On NCCL this is a reduce-scatter over contiguous CONTIG ROW BLOCKS of the triangle
     followed by an all-gather of the reduced blocks; on other backends (gloo in the CPU tests) one all-reduce.
Clustering (stage 4) then runs replicated: the kernels are deterministic, so every rank computes the same membership
from the same A without any exchange (n <= 2*10^4 vertices is latency-bound, SURVEY.md 8e caveat).

The functions take tensors on whatever device the process group works on, so the collective logic is exercised on CPU
with gloo (tests/test_parallel_cpu.py) and on NVLink with NCCL (bench.py --gpus N).
"""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist


def world():
    return (dist.get_rank(), dist.get_world_size()) if dist.is_available() and dist.is_initialized() else (0, 1)


def shard_range(m, rank, world_size):
    """contiguous concatemer-id range [first, first + count) of `rank`; the ranges tile [0, m) exactly."""
    base, rem = divmod(int(m), int(world_size))
    first = rank * base + min(rank, rem)
    return first, base + (1 if rank < rem else 0)


def tri_offset(i, n):
    """index of (i, i) in the packed upper triangle acc[i(2n-i+1)/2 + (j-i)]."""
    return i * (2 * n - i + 1) // 2


def row_blocks(n, world_size):
    """contig row blocks [lo, hi) whose packed-triangle slices have (nearly) equal element counts: rank r reduces the
    rows of block r.  Returns a list of (row_lo, row_hi)."""
    total = n * (n + 1) // 2
    cuts = [0]
    for r in range(1, world_size):
        target = total * r // world_size
        # smallest i with tri_offset(i) >= target
        lo, hi = cuts[-1], n
        while lo < hi:
            mid = (lo + hi) // 2
            if tri_offset(mid, n) >= target:
                hi = mid
            else:
                lo = mid + 1
        cuts.append(lo)
    cuts.append(n)
    return [(cuts[r], cuts[r + 1]) for r in range(world_size)]


def allreduce_rowsums(rowsum, max_count, group=None):
    """exchange 1: global contact sums and largest count (used as the `rowsum_hook` of device.incidence)."""
    if world()[1] > 1:
        dist.all_reduce(rowsum, op=dist.ReduceOp.SUM, group=group)
        dist.all_reduce(max_count, op=dist.ReduceOp.MAX, group=group)
    return rowsum, max_count


def sum_triangle(acc, n, group=None):
    """exchange 2: acc (int64 packed upper triangle of n rows, fixed point) <- sum over ranks, in place.

    NCCL: reduce-scatter over the row blocks of ``row_blocks`` (padded to equal length), then all-gather.  Each rank
    therefore owns the reduced rows of its block between the two steps -- the hook for a row-partitioned consumer."""
    rank, ws = world()
    if ws == 1:
        return acc
    backend = dist.get_backend(group)
    if backend != "nccl":
        dist.all_reduce(acc, op=dist.ReduceOp.SUM, group=group)
        return acc
    blocks = row_blocks(n, ws)
    offs = [tri_offset(lo, n) for lo, _ in blocks] + [n * (n + 1) // 2]
    sizes = [offs[r + 1] - offs[r] for r in range(ws)]
    chunk = max(sizes)
    if all(s == chunk for s in sizes):
        send = acc[:chunk * ws]
    else:
        send = torch.zeros(chunk * ws, dtype=acc.dtype, device=acc.device)
        for r in range(ws):
            send[r * chunk:r * chunk + sizes[r]] = acc[offs[r]:offs[r + 1]]
    mine = torch.empty(chunk, dtype=acc.dtype, device=acc.device)
    dist.reduce_scatter_tensor(mine, send, op=dist.ReduceOp.SUM, group=group)
    dist.all_gather_into_tensor(send, mine, group=group)
    if send.data_ptr() != acc.data_ptr():
        for r in range(ws):
            acc[offs[r]:offs[r + 1]] = send[r * chunk:r * chunk + sizes[r]]
    return acc


def merge_group_results(local, group=None):
    """second-round fan-out ("replicas"; replaces the loky workers of hyperpartition.py:2150-2189): every
    rank contributes the {group number: result} of the first-round groups it ran; returns the union, same on every rank."""
    rank, ws = world()
    if ws == 1:
        return dict(local)
    parts = [None] * ws
    dist.all_gather_object(parts, dict(local), group=group)
    out = {}
    for p in parts:
        dup = set(out) & set(p)
        if dup:
            raise RuntimeError(f"first-round groups {sorted(dup)} were run by more than one rank")
        out.update(p)
    return out


def sum_vector(v, group=None):
    if world()[1] > 1:
        dist.all_reduce(v, op=dist.ReduceOp.SUM, group=group)
    return v


class RankConsensus:
    """device.LocalConsensus over a process group.  The accumulate stage picks its form (direct / swept), its band size
    and its fallback from shard-local numbers (entries, hyperedges, the prep kernel's flag); were each rank to decide
    alone, ranks near a threshold would issue different collectives (per-band all-reduces of different sizes against a
    reduce-scatter) and NCCL would hang or add mismatched cells.  Here the numbers are all-reduced first, so the sequence
    and the sizes of the collectives are the same everywhere."""

    def __init__(self, device, group=None):
        self.group = group
        self.world_size = world()[1]
        self.device = device if dist.get_backend(group) == "nccl" else torch.device("cpu")

    def sums(self, *vals):
        t = torch.tensor([int(v) for v in vals], dtype=torch.int64, device=self.device)
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        return tuple(int(v) for v in t.tolist())

    def any(self, flag):
        t = torch.tensor([1 if flag else 0], dtype=torch.int32, device=self.device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
        return bool(int(t.item()))


def expand_accumulate_sum(H, use_high_order=True, group=None, on_local_done=None):
    """this rank's partial accumulator of A = H W H^T summed over the ranks, in place and identical everywhere
    (exchange 2).  When the accumulator is swept band by band (device.expand_accumulate_swept, n >= 5.6k) the sum of band
    k travels on a side stream while band k + 1 is being accumulated, so only the last band's exchange is exposed;
    otherwise one reduce-scatter + all-gather of the whole triangle follows the kernel.  `on_local_done()` is called when
    the last local kernel has been launched (benchmarks mark the end of the local work there)."""
    from . import device as D
    rank, ws = world()
    if ws == 1:
        acc = D.expand_accumulate(H, use_high_order=use_high_order)
        if on_local_done:
            on_local_done()
        return acc
    overlap = dist.get_backend(group) == "nccl"
    main = torch.cuda.current_stream() if overlap else None
    comm = _comm_stream(H.device) if overlap else None
    holder, bands = {}, []

    def hook(k, lo, hi, c_lo, c_hi):
        if k < 0:               # the swept attempt was abandoned (packed form does not fit): forget its exchanges
            if overlap and bands:
                done = torch.cuda.Event()
                done.record(comm)
                main.wait_event(done)
            bands.clear()
            return
        bands.append((c_lo, c_hi))
        if not overlap:
            return
        ev = torch.cuda.Event()
        ev.record(main)
        comm.wait_event(ev)
        with torch.cuda.stream(comm):
            dist.all_reduce(holder["acc"][c_lo:c_hi], op=dist.ReduceOp.SUM, group=group)

    n = H.n_rows
    holder["acc"] = torch.empty(max(n * (n + 1) // 2, 1), dtype=torch.int64, device=H.device)
    acc = D.expand_accumulate(H, use_high_order=use_high_order, band_hook=hook, acc_out=holder["acc"],
                              consensus=RankConsensus(H.device, group))
    if on_local_done:
        on_local_done()
    if bands and overlap:
        done = torch.cuda.Event()
        done.record(comm)
        main.wait_event(done)
    elif bands:
        for c_lo, c_hi in bands:
            dist.all_reduce(acc.acc[c_lo:c_hi], op=dist.ReduceOp.SUM, group=group)
    else:
        sum_triangle(acc.acc, acc.n, group)
    sum_vector(acc.diag, group)
    return acc


_COMM = {}


def _comm_stream(device):
    key = str(device)
    if key not in _COMM:
        _COMM[key] = torch.cuda.Stream(device=device)
    return _COMM[key]


def sharded_clique_expansion(raw, min_contacts=25, min_weight=2.0, use_high_order=True, node_w=None, group=None,
                             want32=True, want64=True, min_mapq=0, mapq_filter=True, row_drop=None):
    """stage 1c + 2 on this rank's shard `raw` (device.RawEdges of its concatemers) with the two exchanges.
    `min_mapq` / `mapq_filter` / `row_drop` are HyperGraph's entry filter (min_quality) and remove_rows, as on one GPU.
    Returns (DeviceH of the shard, DeviceCSR of the GLOBAL A, identical on every rank)."""
    from . import device as D
    H = D.incidence(raw, min_contacts=min_contacts, min_mapq=min_mapq, row_drop=row_drop, mapq_filter=mapq_filter,
                    rowsum_hook=lambda rs, mx: allreduce_rowsums(rs, mx, group))
    acc = expand_accumulate_sum(H, use_high_order=use_high_order, group=group)   # scale from the all-reduced bound
    A = D.expand_emit(acc, min_weight=min_weight, node_w=node_w, want32=want32, want64=want64)
    return H, A
