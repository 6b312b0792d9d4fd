"""Seeded synthetic contact tables of the shapes BASELINE.json names.

Everything is *counter based*: every random draw is ``mix64(h_e + j*K)`` with ``h_e`` a hash of
(seed, concatemer id) and ``j`` a fixed draw slot, so any slice of a table can be regenerated
independently (per GPU shard, or a bounded sample for the CPU baseline) and the numpy code here
and the CUDA twin in ``csrc/synth.cu`` produce identical bytes.

The tables mimic what the reference ingests:
  * Pore-C table columns the path uses (cphasing/hypergraph.py:1500-1504, 1706-1717):
    ``read_idx, chrom(->contig idx), start, end, mapping_quality``
  * 4DN pairs columns (cphasing/hypergraph.py:904-909): ``chrom1, pos1, chrom2, pos2, mapq``
  * allele table pairs (cphasing/core.py:289-300) for kprune.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

U64 = np.uint64
_PHI = U64(0x9E3779B97F4A7C15)
_K = U64(0xD1B54A32D192ED03)
_M1 = U64(0xBF58476D1CE4E5B9)
_M2 = U64(0x94D049BB133111EB)

# draw slots (shared with csrc/synth.cu)
SLOT_ORDER = 0
SLOT_GROUP = 1
SLOT_MONO = 8          # + 8*s + {0: class, 1: hap, 2: contig, 3: multiplicity, 4: start, 5: span}
SLOT_REC = 1024        # + 16*s + r  -> mapq of record r of monomer s
MAX_MULT = 8
MAX_ORDER = 50


def mix64(x):
    """splitmix64 finaliser over uint64 arrays (wraps mod 2**64)."""
    with np.errstate(over="ignore"):
        z = (np.asarray(x, dtype=U64) + _PHI).astype(U64)
        z = (z ^ (z >> U64(30))) * _M1
        z = (z ^ (z >> U64(27))) * _M2
        return z ^ (z >> U64(31))


def edge_hash(seed, e):
    with np.errstate(over="ignore"):
        return mix64(U64(seed) ^ ((np.asarray(e, dtype=U64) + U64(1)) * _PHI))


def draw(h, j):
    """uniform double in [0,1) for draw slot j of the concatemer with hash h."""
    with np.errstate(over="ignore"):
        z = mix64(h + U64(j) * _K)
    return (z >> U64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)


@dataclass
class Genome:
    """Contigs, their lengths and the planted haplotype groups."""
    n: int
    chromosomes: int
    ploidy: int
    length: np.ndarray            # int64[n]
    group: np.ndarray             # int32[n]  planted group (chromosome*ploidy + hap)
    order: np.ndarray             # int32[n]  contigs sorted by group (stable)
    cum: np.ndarray               # float64[n+1] cumulative length over `order`
    gstart: np.ndarray            # int32[G+1] range of each group in `order`
    names: list = field(default_factory=list)

    @property
    def G(self):
        return self.chromosomes * self.ploidy

    @property
    def contig_idx(self):
        return {c: i for i, c in enumerate(self.names)}

    @property
    def contigsizes(self):
        return {c: int(l) for c, l in zip(self.names, self.length)}


def make_genome(n, chromosomes, ploidy, seed):
    """log-normal(ln 2 Mb, 0.8) lengths clipped to [50 kb, 40 Mb]; round-robin groups."""
    rng = np.random.default_rng(seed)
    length = np.exp(rng.normal(np.log(2e6), 0.8, size=n))
    length = np.clip(length, 5e4, 4e7).astype(np.int64)
    G = chromosomes * ploidy
    group = (np.arange(n) % G).astype(np.int32)
    order = np.argsort(group, kind="stable").astype(np.int32)
    cum = np.concatenate(([0.0], np.cumsum(length[order].astype(np.float64))))
    gstart = np.searchsorted(group[order], np.arange(G + 1)).astype(np.int32)
    names = [f"{(g // ploidy) + 1}{chr(ord('A') + g % ploidy)}.ctg{i}" for i, g in enumerate(group)]
    return Genome(n, chromosomes, ploidy, length, group, order, cum, gstart, names)


def _poisson2_cdf():
    # P(X<=x) for Poisson(2), x = 0..47
    from math import exp, factorial
    p = np.array([exp(-2.0) * 2.0 ** x / factorial(x) for x in range(48)])
    return np.cumsum(p)


_POIS2 = _poisson2_cdf()


def concatemer_orders(h, law="poisson"):
    u = draw(h, SLOT_ORDER)
    if law == "poisson":          # k = min(2 + Poisson(2), 50)
        k = 2 + np.searchsorted(_POIS2, u, side="right")
    elif law == "exp":            # k = min(2 + floor(Exp(mean 8)), 50)  (HiFi-C / CiFi heavy tail)
        k = 2 + np.floor(-8.0 * np.log1p(-u)).astype(np.int64)
    elif law == "pair":
        k = np.full(u.shape, 2, dtype=np.int64)
    else:
        raise ValueError(law)
    return np.minimum(k, MAX_ORDER).astype(np.int32)


def _pick_in_range(g: Genome, lo, hi, u):
    """contig (position in g.order) in [lo,hi) with probability proportional to length."""
    x = g.cum[lo] + u * (g.cum[hi] - g.cum[lo])
    p = np.searchsorted(g.cum, x, side="right") - 1
    return np.clip(p, lo, hi - 1)


def monomer_contigs(g: Genome, h, s, grp):
    """contig of monomer slot s of each concatemer (hash h, home group grp)."""
    uc = draw(h, SLOT_MONO + 8 * s + 0)
    uh = draw(h, SLOT_MONO + 8 * s + 1)
    up = draw(h, SLOT_MONO + 8 * s + 2)
    chrom = grp // g.ploidy
    hap = grp % g.ploidy
    if g.ploidy > 1:
        ohap = (hap + 1 + np.floor(uh * (g.ploidy - 1)).astype(np.int64)) % g.ploidy
    else:
        ohap = hap
    homolog = chrom * g.ploidy + ohap
    tgt = np.where(uc < 0.85, grp, homolog)
    lo = g.gstart[tgt].astype(np.int64)
    hi = g.gstart[tgt + 1].astype(np.int64)
    anywhere = uc >= 0.95
    lo = np.where(anywhere, 0, lo)
    hi = np.where(anywhere, g.n, hi)
    return g.order[_pick_in_range(g, lo, hi, up)]


def mapq_from_u(u):
    """MAPQ in {0 (10 %), 1..59 uniform (20 %), 60 (70 %)}."""
    q = np.where(u < 0.1, 0, np.where(u < 0.3, 1 + np.floor((u - 0.1) / 0.2 * 59).astype(np.int64), 60))
    return np.minimum(q, 60).astype(np.int8)


def home_group(g: Genome, h):
    glen = g.cum[g.gstart[1:]] - g.cum[g.gstart[:-1]]
    gc = np.concatenate(([0.0], np.cumsum(glen)))
    x = draw(h, SLOT_GROUP) * gc[-1]
    return np.clip(np.searchsorted(gc, x, side="right") - 1, 0, g.G - 1).astype(np.int64)


def porec_table(g: Genome, m, seed, law="poisson", first=0):
    """Pore-C alignment records of concatemers [first, first+m), sorted by read_idx.

    Returns dict of numpy columns: read_idx u32, contig_idx u32, start u32, end u32, mapq i8.
    """
    e = np.arange(first, first + m, dtype=np.int64)
    h = edge_hash(seed, e)
    k = concatemer_orders(h, law)
    grp = home_group(g, h)
    cols = {"read_idx": [], "contig_idx": [], "start": [], "end": [], "mapq": []}
    kmax = int(k.max()) if m else 0
    for s in range(kmax):
        act = np.nonzero(k > s)[0]
        ha = h[act]
        ctg = monomer_contigs(g, ha, s, grp[act])
        um = draw(ha, SLOT_MONO + 8 * s + 3)
        # Geometric(0.7): P(r) = 0.7 * 0.3**(r-1)
        mult = 1 + np.floor(np.log1p(-um) / np.log(0.3)).astype(np.int64)
        mult = np.clip(mult, 1, MAX_MULT)
        L = g.length[ctg]
        st = np.floor(draw(ha, SLOT_MONO + 8 * s + 4) * L).astype(np.int64)
        span = 100 + np.floor(draw(ha, SLOT_MONO + 8 * s + 5) * 4900).astype(np.int64)
        en = np.minimum(st + span, L)
        for r in range(MAX_MULT):
            ar = np.nonzero(mult > r)[0]
            if ar.size == 0:
                break
            q = mapq_from_u(draw(ha[ar], SLOT_REC + 16 * s + r))
            cols["read_idx"].append(e[act][ar])
            cols["contig_idx"].append(ctg[ar])
            cols["start"].append(st[ar])
            cols["end"].append(en[ar])
            cols["mapq"].append(q)
    out = {}
    if not cols["read_idx"]:
        return {"read_idx": np.zeros(0, np.uint32), "contig_idx": np.zeros(0, np.uint32),
                "start": np.zeros(0, np.uint32), "end": np.zeros(0, np.uint32), "mapq": np.zeros(0, np.int8)}
    rid = np.concatenate(cols["read_idx"])
    # canonical record order = (read, slot, record) -- the loops above emit slot-major, so sort stably
    perm = np.argsort(rid, kind="stable")
    out["read_idx"] = rid[perm].astype(np.uint32)
    out["contig_idx"] = np.concatenate(cols["contig_idx"])[perm].astype(np.uint32)
    out["start"] = np.concatenate(cols["start"])[perm].astype(np.uint32)
    out["end"] = np.concatenate(cols["end"])[perm].astype(np.uint32)
    out["mapq"] = np.concatenate(cols["mapq"])[perm].astype(np.int8)
    return out


def pairs_table(g: Genome, m, seed, first=0, intra=0.05):
    """Hi-C pairs [first, first+m): chrom1, pos1, chrom2, pos2, mapq (k = 2 exactly)."""
    e = np.arange(first, first + m, dtype=np.int64)
    h = edge_hash(seed, e)
    grp = home_group(g, h)
    c1 = monomer_contigs(g, h, 0, grp)
    c2 = monomer_contigs(g, h, 1, grp)
    same = draw(h, SLOT_MONO + 8 * 0 + 3) < intra
    c2 = np.where(same, c1, c2)
    p1 = np.floor(draw(h, SLOT_MONO + 8 * 0 + 4) * g.length[c1]).astype(np.uint32)
    p2 = np.floor(draw(h, SLOT_MONO + 8 * 1 + 4) * g.length[c2]).astype(np.uint32)
    q = mapq_from_u(draw(h, SLOT_REC))
    return {"chrom1": c1.astype(np.uint32), "pos1": p1, "chrom2": c2.astype(np.uint32), "pos2": p2, "mapq": q}


def allele_pairs(g: Genome, seed, per_contig=1):
    """Allele table rows (contig1, contig2, mz1, mz2, mzShared, similarity) in the allele2 layout.

    Every contig is paired with the co-located contigs (same rank inside the group) of the other
    haplotypes of its chromosome; similarity ~ U(0.85, 0.99); mz fields derive from the lengths.
    """
    rng = np.random.default_rng(seed)
    rows = []
    rank = np.empty(g.n, dtype=np.int64)
    for gi in range(g.G):
        members = g.order[g.gstart[gi]:g.gstart[gi + 1]]
        rank[members] = np.arange(len(members))
    by_group = [g.order[g.gstart[gi]:g.gstart[gi + 1]] for gi in range(g.G)]
    for i in range(g.n):
        gi = int(g.group[i]); c = gi // g.ploidy
        for hp in range(g.ploidy):
            gj = c * g.ploidy + hp
            if gj == gi:
                continue
            mem = by_group[gj]
            if len(mem) == 0:
                continue
            j = int(mem[min(rank[i], len(mem) - 1)])
            sim = float(rng.uniform(0.85, 0.99))
            mz1 = int(g.length[i] // 1000); mz2 = int(g.length[j] // 1000)
            rows.append((i, j, mz1, mz2, int(min(mz1, mz2) * sim), sim))
    return rows


# ---- the named configurations (BASELINE.json "configs") -------------------------
CONFIGS = {
    # name: (n contigs, chromosomes, ploidy, m concatemers, order law, kind, seed)
    "A": dict(n=400, chromosomes=5, ploidy=4, m=200_000, law="poisson", kind="porec", seed=20240),
    "B": dict(n=1_000, chromosomes=10, ploidy=2, m=100_000_000, law="pair", kind="pairs", seed=20241),
    "C": dict(n=4_000, chromosomes=12, ploidy=4, m=50_000_000, law="poisson", kind="porec", seed=20242),
    "D": dict(n=20_000, chromosomes=10, ploidy=6, m=500_000_000, law="poisson", kind="porec", seed=20243),
    "E": dict(n=10_000, chromosomes=10, ploidy=4, m=200_000_000, law="exp", kind="porec", seed=20244),
}


def genome_for(name):
    c = CONFIGS[name]
    return make_genome(c["n"], c["chromosomes"], c["ploidy"], c["seed"])


# ---- device twins (csrc/synth.cu): the bench generates its tables straight into HBM ---------------
_LAW = {"poisson": 0, "exp": 1, "pair": 2}


def _genome_dev(g: Genome, dev):
    import torch
    glen = g.cum[g.gstart[1:]] - g.cum[g.gstart[:-1]]
    gcum = np.concatenate(([0.0], np.cumsum(glen)))
    t = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a.astype(dt))).to(dev)
    return (t(g.cum, np.float64), t(gcum, np.float64), t(g.gstart, np.int32), t(g.order, np.int32), t(g.length, np.int64))


def porec_table_device(g: Genome, m, seed, law="poisson", first=0, positions=False, device=None):
    """dict of device tensors: read_idx u32, contig_idx u32, mapq i8 [, start u32, end u32]."""
    import ctypes as C
    import torch
    from . import _lib
    from .device import _dev
    dev = _dev(device)
    L = _lib.lib()
    off = torch.empty(m + 1, dtype=torch.int64, device=dev)
    tot = C.c_uint64(0)
    _lib.call_ws("cph_synth_porec_count", dev, int(seed), int(first), int(m), _LAW[law], off.data_ptr(), C.byref(tot))
    N = int(tot.value)
    cum, gcum, gstart, order, length = _genome_dev(g, dev)
    out = {"read_idx": torch.empty(max(N, 1), dtype=torch.uint32, device=dev)[:N],
           "contig_idx": torch.empty(max(N, 1), dtype=torch.uint32, device=dev)[:N],
           "mapq": torch.empty(max(N, 1), dtype=torch.int8, device=dev)[:N]}
    if positions:
        out["start"] = torch.empty(max(N, 1), dtype=torch.uint32, device=dev)[:N]
        out["end"] = torch.empty(max(N, 1), dtype=torch.uint32, device=dev)[:N]
    _lib.check(L.cph_synth_porec_fill(int(seed), int(first), int(m), _LAW[law], g.n, g.G, g.ploidy, cum.data_ptr(),
                                      gcum.data_ptr(), gstart.data_ptr(), order.data_ptr(), length.data_ptr(),
                                      off.data_ptr(), out["read_idx"].data_ptr(), out["contig_idx"].data_ptr(),
                                      out["start"].data_ptr() if positions else None,
                                      out["end"].data_ptr() if positions else None, out["mapq"].data_ptr(), _lib.stream()))
    torch.cuda.current_stream().synchronize()
    return out


def pairs_table_device(g: Genome, m, seed, first=0, intra=0.05, positions=False, device=None):
    import torch
    from . import _lib
    from .device import _dev
    dev = _dev(device)
    cum, gcum, gstart, order, length = _genome_dev(g, dev)
    out = {"chrom1": torch.empty(m, dtype=torch.uint32, device=dev), "chrom2": torch.empty(m, dtype=torch.uint32, device=dev),
           "mapq": torch.empty(m, dtype=torch.int8, device=dev)}
    if positions:
        out["pos1"] = torch.empty(m, dtype=torch.uint32, device=dev)
        out["pos2"] = torch.empty(m, dtype=torch.uint32, device=dev)
    _lib.check(_lib.lib().cph_synth_pairs(int(seed), int(first), int(m), float(intra), g.n, g.G, g.ploidy, cum.data_ptr(),
                                          gcum.data_ptr(), gstart.data_ptr(), order.data_ptr(), length.data_ptr(),
                                          out["chrom1"].data_ptr(), out["chrom2"].data_ptr(),
                                          out["pos1"].data_ptr() if positions else None,
                                          out["pos2"].data_ptr() if positions else None, out["mapq"].data_ptr(),
                                          _lib.stream()))
    torch.cuda.current_stream().synchronize()
    return out
