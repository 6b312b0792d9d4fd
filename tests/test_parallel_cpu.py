"""World-size-2 gloo tests (CPU) of the multi-GPU host logic in cphasing_b200/parallel.py: shard ranges, row blocks,
and that the two exchanges reproduce the single-process result.  The per-shard arithmetic here is done by the ORACLE
(test infrastructure) -- on the GPU box the same collectives move the CUDA kernels' partial accumulators."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from cphasing_b200 import parallel as par


def test_shard_ranges_tile_exactly():
    for m in (0, 1, 7, 1000, 500_000_001):
        for ws in (1, 2, 3, 8):
            pos = 0
            for r in range(ws):
                first, cnt = par.shard_range(m, r, ws)
                assert first == pos and cnt >= 0
                pos += cnt
            assert pos == m
            sizes = [par.shard_range(m, r, ws)[1] for r in range(ws)]
            assert max(sizes) - min(sizes) <= 1


def test_row_blocks_cover_and_balance():
    for n in (1, 5, 400, 4000, 20000):
        for ws in (1, 2, 4, 8):
            blocks = par.row_blocks(n, ws)
            assert blocks[0][0] == 0 and blocks[-1][1] == n
            assert all(blocks[i][1] == blocks[i + 1][0] for i in range(ws - 1))
            if n >= 400:
                tot = n * (n + 1) // 2
                sz = [par.tri_offset(hi, n) - par.tri_offset(lo, n) for lo, hi in blocks]
                assert sum(sz) == tot
                assert max(sz) - min(sz) <= 2 * n            # within one row of perfect balance


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, ws, port, m, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=ws)
    try:
        from cphasing_b200 import synth
        from oracle import hypergraph_oracle as orc
        g = synth.make_genome(60, 3, 2, seed=11)
        first, cnt = par.shard_range(m, rank, ws)
        t = synth.porec_table(g, cnt, seed=12, first=first)
        r, c, k, mq = orc.porec_edges(t["read_idx"], t["contig_idx"], t["mapq"], min_mapq=1, n_contigs=g.n)
        # exchange 1: global row sums / max count
        rs = torch.from_numpy(np.bincount(r, weights=k, minlength=g.n).astype(np.int64))
        mx = torch.tensor([int(k.max()) if len(k) else 0], dtype=torch.int32)
        par.allreduce_rowsums(rs, mx)
        keep = rs.numpy() >= 25
        c = c - first if len(c) else c
        H, _, _ = orc.incidence_matrix(r, c, k, g.n, max(int(c.max()) + 1 if len(c) else 1, 1), 0)
        H = H[keep]
        H = H[:, np.asarray(H.sum(axis=0)).ravel() >= 2]                 # the column filter is column-local
        n = H.shape[0]
        C = orc.contact_counts(H).toarray().astype(np.int64)             # integer partial of this shard
        iu = np.triu_indices(n)
        acc = torch.from_numpy(np.ascontiguousarray(C[iu]))              # packed upper triangle, row-major
        par.sum_triangle(acc, n)                                         # exchange 2
        q.put((rank, rs.numpy(), int(mx.item()), n, acc.numpy()))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_two_rank_column_shards_reproduce_the_single_process_counts():
    from cphasing_b200 import synth
    from oracle import hypergraph_oracle as orc
    m, ws = 4000, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, ws, port, m, q)) for r in range(ws)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=240) for _ in range(ws)], key=lambda x: x[0])
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    # single process
    g = synth.make_genome(60, 3, 2, seed=11)
    t = synth.porec_table(g, m, seed=12)
    r, c, k, mq = orc.porec_edges(t["read_idx"], t["contig_idx"], t["mapq"], min_mapq=1, n_contigs=g.n)
    rs = np.bincount(r, weights=k, minlength=g.n).astype(np.int64)
    H, _, _ = orc.incidence_matrix(r, c, k, g.n, int(c.max()) + 1, 25)
    C = orc.contact_counts(H).toarray().astype(np.int64)
    iu = np.triu_indices(H.shape[0])
    for rank, rs_r, mx_r, n_r, acc_r in res:
        assert np.array_equal(rs_r, rs)
        assert mx_r == int(k.max())
        assert n_r == H.shape[0]
        assert np.array_equal(acc_r, C[iu]), "summed integer partials must be bit-exact"


def _fan_worker(rank, ws, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=ws)
    try:
        # 5 first-round groups; rank r runs the jobs r, r + ws, ... exactly as HyperPartition.incremental_partition does
        jobs = [(num, list(range(10 * num, 10 * num + 4))) for num in range(1, 6)]
        local = {num: (None, [grp[:2], grp[2:]]) for j, (num, grp) in enumerate(jobs) if j % ws == rank}
        merged = par.merge_group_results(local)
        q.put((rank, sorted(merged), merged[3][1]))
        try:
            par.merge_group_results({1: (None, [])})               # every rank claims group 1: must be refused
            q.put((rank, "no error"))
        except RuntimeError as e:
            q.put((rank, "refused" if "more than one rank" in str(e) else repr(e)))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_second_round_fan_out_gathers_every_group_on_every_rank():
    """the replicas form of the second round (hyperpartition.py:2150-2189 loky fan-out): round-robin jobs, all-gather"""
    ws = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_fan_worker, args=(r, ws, port, q)) for r in range(ws)]
    for p in procs:
        p.start()
    got = [q.get(timeout=240) for _ in range(2 * ws)]
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    full = [g for g in got if len(g) == 3]
    assert len(full) == ws
    for rank, nums, grp3 in full:
        assert nums == [1, 2, 3, 4, 5]
        assert grp3 == [[30, 31], [32, 33]]
    assert all(g[1] == "refused" for g in got if len(g) == 2)


def test_sweep_bands_partition_the_triangle():
    from cphasing_b200 import device as D
    for n, mb in ((1, 1), (7, 0.0001), (6000, 3), (20000, 128), (10000, 96), (100000, 96)):
        r0 = D.sweep_bands(n, band_bytes=int(mb * (1 << 20)))
        assert r0[0] == 0 and r0[-1] == n and all(a < b for a, b in zip(r0, r0[1:]))
        cells = [b * (2 * n - b + 1) // 2 - a * (2 * n - a + 1) // 2 for a, b in zip(r0, r0[1:])]
        assert sum(cells) == n * (n + 1) // 2
        limit = max(int(mb * (1 << 20)) // 8, n)
        assert all(c <= limit or b - a == 1 for c, a, b in zip(cells, r0, r0[1:]))     # a band is at most the limit (or one row)


def _band_sum_worker(rank, ws, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=ws)
    try:
        from cphasing_b200 import device as D

        class H:                       # stands in for a DeviceH shard; the accumulate below is a host fake
            n_rows, device = 50, torch.device("cpu")
        n = H.n_rows
        tri = n * (n + 1) // 2
        real = D.expand_accumulate

        def fake(Hs, use_high_order=True, band_hook=None, acc_out=None, **kw):
            acc = acc_out if acc_out is not None else torch.empty(tri, dtype=torch.int64)
            r0 = D.sweep_bands(n, band_bytes=8 * 200)          # ~7 bands
            for k in range(len(r0) - 1):
                lo, hi = r0[k], r0[k + 1]
                c_lo, c_hi = lo * (2 * n - lo + 1) // 2, hi * (2 * n - hi + 1) // 2
                acc[c_lo:c_hi] = torch.arange(c_lo, c_hi, dtype=torch.int64) * (rank + 1)
                band_hook(k, lo, hi, c_lo, c_hi)
            return D.Accumulator(acc, torch.full((n,), rank + 1, dtype=torch.int64), n, 0, bool(use_high_order))
        D.expand_accumulate = fake
        marks = []
        acc = par.expand_accumulate_sum(H, use_high_order=False, on_local_done=lambda: marks.append(1))
        D.expand_accumulate = real
        q.put((rank, acc.acc.clone(), acc.diag.clone(), len(marks)))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_band_by_band_exchange_sums_every_cell_once():
    """parallel.expand_accumulate_sum with a swept accumulator: every band all-reduced exactly once (gloo: after the local
    work; NCCL runs the same calls on a side stream under the next band), diagonal summed, hook for the benchmark called"""
    ws = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_band_sum_worker, args=(r, ws, port, q)) for r in range(ws)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in range(ws)]
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    n = 50
    tri = n * (n + 1) // 2
    for rank, acc, diag, nm in res:
        assert torch.equal(acc, torch.arange(tri, dtype=torch.int64) * 3)      # (1 + 2) x the cell's own index
        assert torch.equal(diag, torch.full((n,), 3, dtype=torch.int64))
        assert nm == 1


def _consensus_worker(rank, ws, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=ws)
    try:
        c = par.RankConsensus(torch.device("cpu"))
        # shard sizes on either side of a threshold, a flag raised on one rank only
        q.put((rank, c.world_size, c.sums(100 + rank, 7), c.any(rank == 1), c.any(False)))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_rank_consensus_gives_every_rank_the_same_answer():
    """the form / band-size / fallback choices of the accumulate stage are taken on all-reduced numbers (ADVICE r2:
    rank-local choices could issue mismatched collectives)"""
    ws = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_consensus_worker, args=(r, ws, port, q)) for r in range(ws)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in range(ws)]
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    for rank, w, sums, flagged, clean in res:
        assert w == 2 and sums == (201, 14) and flagged is True and clean is False


def test_local_consensus_is_the_identity():
    from cphasing_b200 import device as D
    c = D.LocalConsensus()
    assert c.world_size == 1 and c.sums(5, 6) == (5, 6) and c.any(0) is False and c.any(3) is True
