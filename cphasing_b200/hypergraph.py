"""Drop-in for ``cphasing.hypergraph``: ``Extractor`` (4DN pairs) and ``HyperExtractor`` (Pore-C
tables) -> ``HyperEdges`` (reference: cphasing/hypergraph.py:621-1306, 1395-1911).

Text decoding (gzip / TSV -> columns) stays on the host, as SURVEY.md 8(a1) scopes it; the name ->
index join is a pandas categorical lookup; every filter, the per-read order filter, the
``group_by(read, contig)`` and the sort by contig run on the device (csrc/edges.cu).  Both classes
also accept in-memory columnar input (``from_arrays``), which is what bench.py's end-to-end leg and
the parity tests use.  The HCR-BED midpoint filter (hypergraph.py:1721-1752, 786-893) and the split-contig row index
(:1406-1424, 994-1023) run inside the same record-filter step of the kernels (``cph_prefilter``): the host only
decodes the BED file and builds the per-contig interval table.
"""
from __future__ import annotations

import logging
import sys

import numpy as np
import pandas as pd
import torch

from . import device as D
from .algorithms.hypergraph import HyperEdges, LazyTriplets

logger = logging.getLogger(__name__)


def _edges_from_raw(raw: D.RawEdges, contig_idx, contigsizes, lazy=True):
    """RawEdges -> HyperEdges in the reference's wire order (sorted by row, hypergraph.py:1877-1889, 1044)."""
    cache = {}

    def trip(i):
        def get():
            if "t" not in cache:
                cache["t"] = D.edges_to_triplets(raw)
            return cache["t"][i]
        return get
    cols = (LazyTriplets(trip(0), raw.nnz, np.int32), LazyTriplets(trip(1), raw.nnz, np.int32),
            LazyTriplets(trip(2), raw.nnz, np.uint32), LazyTriplets(trip(3), raw.nnz, np.int8))
    if not lazy:
        cols = tuple(np.asarray(c) for c in cols)
    e = HyperEdges(idx=contig_idx, row=cols[0], col=cols[1], count=cols[2], contigsizes=contigsizes, mapq=cols[3])
    e._raw = raw
    return e


def _materialise(edges):
    for f in ("row", "col", "count", "mapq"):
        v = getattr(edges, f)
        if isinstance(v, LazyTriplets):
            setattr(edges, f, np.asarray(v))
    return edges


def _save(edges, output):
    """hypergraph.py:1302-1306, 1907-1911: the same msgpack document (readable by the reference) plus the binary sidecar
    `output + '.bin'` that `hgio.load` prefers (0.1 s instead of 8 s per 8e7 ints on the way back in)."""
    from . import hgio
    _materialise(edges)
    hgio.save(edges, output)
    logger.info(f"Successful output graph into `{output}`")


def _lengths(contig_idx, contigsizes):
    n = len(contig_idx)
    ln = np.zeros(n, dtype=np.int64)
    for c, i in contig_idx.items():
        ln[i] = int(contigsizes.get(c, 0)) if hasattr(contigsizes, "get") else int(contigsizes[c])
    return ln


def _name_codes(series, contig_idx):
    """contig names -> index (n_contigs for names outside the index: the inner join of :1435 drops them)."""
    cat = pd.Categorical(series)
    lut = np.array([contig_idx.get(c, len(contig_idx)) for c in cat.categories], dtype=np.uint32)
    codes = np.asarray(cat.codes)
    out = np.where(codes >= 0, lut[np.maximum(codes, 0)], len(contig_idx)).astype(np.uint32)
    return out


def _name_space(contig_idx, split_length, split_contig_boundarys):
    """code space of the contig column(s): the contig index itself, or -- with splitting on -- the ORIGINAL contig names
    of split_contig_boundarys (cli.py:4776-4779: name -> first split bin), in its own order."""
    if split_length and split_contig_boundarys is not None:
        return {c: i for i, c in enumerate(split_contig_boundarys.keys())}
    return contig_idx


def _hcr_table(hcr_bed, names):
    """BED (chrom, start, end) -> (ptr int64[n + 1], start u32, end u32) ordered by (name code, start); stable, so that
    among equal starts the last listed interval is the one join_asof(backward) lands on (hypergraph.py:1721-1731)."""
    if isinstance(hcr_bed, (str, bytes)) or hasattr(hcr_bed, "__fspath__"):
        bed = pd.read_csv(hcr_bed, sep="\t", header=None, usecols=[0, 1, 2], names=["chrom", "start", "end"], comment="#")
    elif isinstance(hcr_bed, pd.DataFrame):
        bed = hcr_bed.iloc[:, :3].copy()
        bed.columns = ["chrom", "start", "end"]
    else:
        c, a, b = hcr_bed
        bed = pd.DataFrame({"chrom": list(c), "start": np.asarray(a), "end": np.asarray(b)})
    n = len(names)
    code = bed["chrom"].map(names.get)
    ok = code.notna().to_numpy()
    code = code[ok].to_numpy().astype(np.int64)
    st = bed["start"].to_numpy()[ok].astype(np.uint32)
    en = bed["end"].to_numpy()[ok].astype(np.uint32)
    order = np.lexsort((st, code))          # lexsort is stable: file order survives among equal (code, start)
    code, st, en = code[order], st[order], en[order]
    ptr = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(np.bincount(code, minlength=n), out=ptr[1:])
    return ptr, st, en


def _prefilter(contig_idx, split_length, split_contig_boundarys, hcr_bed, hcr_invert):
    """-> (D.PreFilter or None, name -> code dict of the contig column)"""
    names = _name_space(contig_idx, split_length, split_contig_boundarys)
    if not (hcr_bed is not None or (split_length and split_contig_boundarys is not None)):
        return None, names
    pf = D.PreFilter(n_names=len(names), hcr_invert=bool(hcr_invert))
    if hcr_bed is not None:
        pf.bed_ptr, pf.bed_start, pf.bed_end = _hcr_table(hcr_bed, names)
    if split_length and split_contig_boundarys is not None:
        pf.split_length = int(split_length)
        pf.split_init = np.array([int(v) for v in split_contig_boundarys.values()], dtype=np.uint32)
    return pf, names


class _NotGrouped(Exception):
    pass


class HyperExtractor:
    """hypergraph.py:1483-1911."""

    def __init__(self, pore_c_table_pathes, contig_idx, contigsizes, min_order=2, max_order=50, min_alignments=30,
                 min_quality=1, edge_length=2e6, split_length=None, split_contig_boundarys=None, hcr_bed=None,
                 hcr_invert=False, threads=4, is_parquet=False, log_dir="logs", _tables=None):
        self.pore_c_table_pathes = list(pore_c_table_pathes) if pore_c_table_pathes is not None else []
        self.contig_idx = contig_idx
        self.contigsizes = contigsizes
        self.min_order, self.max_order = int(min_order), int(max_order)
        self.min_quality = int(min_quality)
        self.edge_length = int(edge_length) if edge_length else 0
        self.split_length, self.split_contig_boundarys = split_length, split_contig_boundarys
        self.hcr_bed, self.hcr_invert = hcr_bed, hcr_invert
        self.prefilter, self.names = _prefilter(contig_idx, split_length, split_contig_boundarys, hcr_bed, hcr_invert)
        self.threads = threads
        self.is_parquet = is_parquet
        tables = _tables if _tables is not None else list(self.pore_c_table_pathes)
        self.edges = self.generate_edges(tables)

    @classmethod
    def from_arrays(cls, read_idx, contig_idx_col, mapq, contig_idx, contigsizes, start=None, end=None, **kw):
        """in-memory columnar input: one table of (read_idx u32, contig index u32, mapq i8[, start, end])."""
        t = {"read_idx": read_idx, "contig": contig_idx_col, "mapq": mapq, "start": start, "end": end}
        return cls(None, contig_idx, contigsizes, _tables=[t], **kw)

    def _load(self, path):
        """whole table in memory (parquet input, and the fallback for text tables that are not grouped by read)"""
        # porec table columns (hypergraph.py:1500-1504, 1706-1717): read_idx(0) chrom(5) start(6) end(7) mapq(8)
        if self.is_parquet:
            df = pd.read_parquet(path, columns=["read_idx", "chrom", "start", "end", "mapping_quality"])
            df.columns = ["read_idx", "chrom", "start", "end", "mapq"]
            return {"read_idx": df["read_idx"].to_numpy(), "contig": _name_codes(df["chrom"], self.names),
                    "mapq": df["mapq"].to_numpy(), "start": df["start"].to_numpy(), "end": df["end"].to_numpy()}
        from . import ingest
        cols, bad = ingest.read_table(path, ingest.POREC_FIELDS, self.names, threads=max(int(self.threads or 1), 1))
        self._check_table(path, len(cols["read_idx"]), bad)
        return cols

    @staticmethod
    def _check_table(path, n_rows, bad):
        """stop (exit status -1 like hypergraph.py:1621-1630) on a table that cannot be used, saying which of the two it is"""
        if bad:
            logger.error(f"The pore-c table `{path}` holds {bad} malformed line(s) (fewer than 9 tab-separated columns, or a "
                         "non-numeric read_idx / start / end / mapping_quality), please check it.")
            sys.exit(-1)
        if n_rows == 0:
            logger.error(f"The pore-c table `{path}` is empty, please check it.")
            sys.exit(-1)

    def _stream(self, path):
        """text table -> chunks of whole reads, decoded block by block while the previous chunk is on the GPU
        (cphasing_b200/ingest.py).  Yields column dicts; raises _NotGrouped when the file is not ordered by read id (the
        group-by of chunk k would then miss records that come later) -- the caller falls back to `_load`."""
        from . import ingest
        carry, last, total = None, -1, 0
        for cols, bad in ingest.read_table_chunks(path, ingest.POREC_FIELDS, self.names,
                                                  threads=max(int(self.threads or 1), 1)):
            if bad:                      # fail at the first bad block, before more chunks are grouped on the GPU
                self._check_table(path, 1, bad)
            r = cols["read_idx"]
            if len(r) == 0:
                continue
            if int(r[0]) < last or np.any(r[1:] < r[:-1]):
                raise _NotGrouped()
            last = int(r[-1])
            if carry is not None:
                cols = {k: np.concatenate([carry[k], v]) for k, v in cols.items()}
                r = cols["read_idx"]
            cut = int(np.searchsorted(r, r[-1], side="left"))          # the last read may continue in the next block
            carry = {k: v[cut:] for k, v in cols.items()}
            if cut:
                total += cut
                yield {k: v[:cut] for k, v in cols.items()}
        if carry is not None and len(carry["read_idx"]):
            total += len(carry["read_idx"])
            yield carry
        self._check_table(path, total, 0)

    def _edges_of(self, t, n, ln, offset):
        return D.porec_to_edges(t["read_idx"], t["contig"], t["mapq"], n, start=t.get("start"), end=t.get("end"),
                                contig_len=ln, min_mapq=self.min_quality, min_order=self.min_order,
                                max_order=self.max_order, edge_length=self.edge_length, read_offset=offset,
                                prefilter=self.prefilter)

    def generate_edges(self, tables):
        """tables: in-memory column dicts (from_arrays) or paths (decoded here, streaming where the format allows)"""
        n = len(self.contig_idx)
        ln = _lengths(self.contig_idx, self.contigsizes) if self.edge_length else None
        parts = []
        offset = 0
        for t in tables:
            if isinstance(t, dict):
                mine = [self._edges_of(t, n, ln, offset)]
            elif self.is_parquet:
                mine = [self._edges_of(self._load(t), n, ln, offset)]
            else:
                try:
                    mine = [self._edges_of(c, n, ln, offset) for c in self._stream(t)]
                except _NotGrouped:
                    logger.info(f"`{t}` is not ordered by read id: decoding it as one table.")
                    mine = [self._edges_of(self._load(t), n, ln, offset)]
            parts.extend(mine)
            # disjoint read ids per input file (hypergraph.py:1871-1875): the next file starts after the largest
            # SURVIVING (already shifted) read id of this one
            if len(tables) > 1:
                offset = max([offset] + [p.n_cols for p in mine if p.nnz])
        raw = parts[0] if len(parts) == 1 else _concat_raw(parts, n)
        logger.info(f"Only retained Pore-C concatemer that: {self.min_order} <= contig order <= {self.max_order}.")
        e = _edges_from_raw(raw, self.contig_idx, self.contigsizes)
        logger.info(f"Result of {raw.nnz:,} raw hyperedges of {raw.n_cols:,} concatemers.")
        return e

    def save(self, output):
        _save(self.edges, output)


def _concat_raw(parts, n):
    cat = lambda f: torch.cat([getattr(p, f) for p in parts])
    packed = all(p.packed and (p.n_edges or not p.nnz) for p in parts)     # every part straight from the group-by kernel
    count = cat("count") if packed else torch.cat([p.plain_count() for p in parts])
    return D.RawEdges(cat("row"), cat("col"), count, cat("mapq"), sum(p.nnz for p in parts),
                      max(p.n_cols for p in parts), n, packed, sum(p.n_edges for p in parts) if packed else 0,
                      min(p.min_order for p in parts) if packed else 0)


class Extractor:
    """hypergraph.py:621-1306 (the live definition)."""

    def __init__(self, pairs_pathes, contig_idx, contigsizes, min_quality=1, hcr_bed=None, hcr_invert=False, threads=4,
                 edge_length=2e6, split_length=None, split_contig_boundarys=None, max_q0_ratio=0.0, low_memory=True,
                 log_dir="logs", _tables=None):
        if max_q0_ratio:
            raise NotImplementedError("max_q0_ratio (per-contig MAPQ-0 ratio filter) is not implemented")
        self.pairs_pathes = list(pairs_pathes) if pairs_pathes is not None else []
        self.contig_idx = contig_idx
        self.contigsizes = contigsizes
        self.min_mapq = int(min_quality)
        self.edge_length = int(edge_length) if edge_length else 0
        split_on = bool(split_length and split_contig_boundarys is not None)
        if split_on and hcr_bed is not None:
            # _filter_pairs_hcr (:786-893) looks the ORIGINAL names of the pairs up in a table keyed by the SPLIT names
            # (:679-731): every join comes back null and no pair survives.  Refused rather than reproduced or "fixed".
            raise NotImplementedError("pairs + HCR BED + split contigs: the reference's path filters every pair out")
        if split_on and self.edge_length:
            # likewise the edge filter joins lengths on original names against the split contigsizes (:977-992): the CLI
            # default --edge-length would empty the table; run with edge_length=0 when splitting pairs
            logger.warning("pairs + split contigs: the edge filter is skipped (the reference's join of original names "
                           "against split contig sizes keeps nothing)")
            self.edge_length = 0
        self.prefilter, self.names = _prefilter(contig_idx, split_length, split_contig_boundarys, hcr_bed, hcr_invert)
        self.threads = threads
        tables = _tables if _tables is not None else [t for p in self.pairs_pathes for t in self._load(p)]
        self.edges = self.generate_edges(tables)

    @classmethod
    def from_arrays(cls, chrom1, chrom2, mapq, contig_idx, contigsizes, pos1=None, pos2=None, **kw):
        t = {"chrom1": chrom1, "chrom2": chrom2, "mapq": mapq, "pos1": pos1, "pos2": pos2}
        return cls(None, contig_idx, contigsizes, _tables=[t], **kw)

    def _load(self, path):
        """-> list of column dicts: one for a pairs file, one per parquet chunk for a PQS directory"""
        from . import ingest
        if ingest.is_pqs(path):
            # PQS store (pqs.py:112-131, 674-809): q1/*.parquet when MAPQ >= 1 is asked for, else q0/*.parquet.  The
            # reference's q0 path additionally keeps only MAPQ-0 pairs and samples them (random, seed 0, :1190-1197,
            # 708-722) on top of q1; that sampling is tied to polars' RNG and is not reproduced.
            if self.min_mapq < 1:
                raise NotImplementedError("PQS input with min_quality = 0 (q0 sampling against q1, pqs.py:708-722) is not "
                                          "implemented; use min_quality >= 1 or a pairs file")
            files, with_mapq = ingest.pqs_chunks(path, self.min_mapq)
            if not files:
                logger.error(f"The pqs `{path}` holds no parquet chunk, please check it.")
                sys.exit(-1)
            lut = lambda ser: _name_codes(ser, self.names)
            return [ingest.read_pqs_chunk(f, lut, self.min_mapq, with_mapq) for f in files]
        # 4DN pairs columns (hypergraph.py:904-909): chrom1(1) pos1(2) chrom2(3) pos2(4) mapq(7)
        cols, bad = ingest.read_table(path, ingest.PAIRS_FIELDS, self.names, threads=max(int(getattr(self, "threads", 4) or 1), 1))
        if bad:
            logger.error(f"The pairs `{path}` holds {bad} line(s) without the columns this step reads (chrom1, pos1, chrom2, "
                         "pos2 in columns 2-5 and the mapq in column 8: a 7-column pairs file has no mapq) or with "
                         "non-numeric values in them, please check it.")
            sys.exit(-1)
        if len(cols["mapq"]) == 0:
            logger.error(f"The pairs `{path}` is empty, can not load anything, please check it.")
            sys.exit(-1)
        return [cols]

    def generate_edges(self, tables):
        n = len(self.contig_idx)
        ln = _lengths(self.contig_idx, self.contigsizes) if self.edge_length else None
        if len(tables) != 1:
            # the reference concatenates multi-file pairs before numbering (:1160-1240); do the same on the host side
            keys = tables[0].keys()
            tables = [{k: (np.concatenate([np.asarray(t[k]) for t in tables]) if tables[0][k] is not None else None)
                       for k in keys}]
        t = tables[0]
        raw = D.pairs_to_edges(t["chrom1"], t["chrom2"], t["mapq"], n, pos1=t.get("pos1"), pos2=t.get("pos2"),
                               contig_len=ln, min_mapq=self.min_mapq, edge_length=self.edge_length,
                               prefilter=self.prefilter)
        logger.info(f"Result of {raw.nnz:,} raw hyperedges of {raw.n_cols:,} pairs.")
        return _edges_from_raw(raw, self.contig_idx, self.contigsizes)

    def save(self, output):
        _save(self.edges, output)
