"""Streaming host decode of the extractors' input files.

The reference reads a whole Pore-C table / pairs file with the multi-threaded polars CSV reader before anything else
happens (cphasing/hypergraph.py:1706-1717, 904-909; PQS directories through pqs.py:1173-1316).  Here a file is read in
blocks of whole lines: one thread inflates / reads, one thread runs the multi-threaded field parser of libcph
(csrc/ingest.cu, `cph_tsv_parse`: text -> u32 / u32 / i8 columns), and the caller groups block k on the GPU while block
k + 1 is being decoded -- bounded host memory and the decode hidden behind the transfer instead of in front of it.

Everything in this module is host work (no CUDA); the parser is part of the C-ABI library like the kernels.
"""
from __future__ import annotations

import ctypes as C
import gzip
import os
import queue
import threading
from pathlib import Path

import numpy as np

from ._lib import check, lib

POREC_FIELDS = (("read_idx", 0, 0), ("contig", 5, 1), ("start", 6, 0), ("end", 7, 0), ("mapq", 8, 2))     # hypergraph.py:1500-1504
PAIRS_FIELDS = (("chrom1", 1, 1), ("pos1", 2, 0), ("chrom2", 3, 1), ("pos2", 4, 0), ("mapq", 7, 2))       # hypergraph.py:904-909
BLOCK_BYTES = 64 << 20
_KIND_DTYPE = {0: np.uint32, 1: np.uint32, 2: np.int8}


def default_threads():
    try:
        return max(1, min(16, len(os.sched_getaffinity(0)) - 2))
    except AttributeError:
        return max(1, min(16, (os.cpu_count() or 2) - 2))


class NameTable:
    """contig name -> code table in the layout `cph_tsv_parse` probes (open addressing, FNV-1a, linear probing).
    `names`: dict name -> code with the codes 0..n-1 (the extractor's contig index or split-name space)."""

    def __init__(self, names):
        n = len(names)
        by_code = [None] * n
        for nm, i in names.items():
            by_code[int(i)] = str(nm)
        if any(x is None for x in by_code):
            raise ValueError("contig codes must be exactly 0..n-1")
        enc = [x.encode() for x in by_code]
        self.n = n
        self.blob = b"".join(enc) or b"\0"
        self.off = np.zeros(n + 1, dtype=np.uint64)
        np.cumsum([len(x) for x in enc], out=self.off[1:])
        size = 2
        while size < 2 * max(n, 1):
            size *= 2
        self.slots = np.zeros(size, dtype=np.uint32)
        fn = lib().cph_fnv1a
        for i, x in enumerate(enc):
            s = int(fn(x, len(x))) & 0xFFFFFFFF & (size - 1)
            while self.slots[s]:
                s = (s + 1) & (size - 1)
            self.slots[s] = i + 1
        self._blob_buf = C.create_string_buffer(self.blob, len(self.blob))


def parse_block(buf, fields, names: NameTable | None, threads=None):
    """one block of whole TSV lines (bytes-like) -> ({field name: numpy array}, malformed line count)"""
    threads = threads or default_threads()
    n = len(buf)
    if isinstance(buf, bytes):
        keep = C.c_char_p(buf)                      # read-only view of the bytes object, no copy
        addr = C.cast(keep, C.c_void_p).value
    else:
        keep = (C.c_char * n).from_buffer(buf)
        addr = C.addressof(keep)
    cap = C.c_uint64(0)
    check(lib().cph_tsv_count_rows(addr, n, threads, C.byref(cap)))
    rows = int(cap.value)
    outs = [np.empty(max(rows, 1), dtype=_KIND_DTYPE[k]) for _, _, k in fields]
    nf = len(fields)
    cols = (C.c_int * nf)(*[c for _, c, _ in fields])
    kinds = (C.c_int * nf)(*[k for _, _, k in fields])
    ptrs = (C.c_void_p * nf)(*[o.ctypes.data for o in outs])
    got, bad = C.c_uint64(0), C.c_uint64(0)
    if names is not None:
        nargs = (C.addressof(names._blob_buf), names.off.ctypes.data, names.n, names.slots.ctypes.data, len(names.slots))
    else:
        nargs = (None, None, 0, None, 0)
    check(lib().cph_tsv_parse(addr, n, threads, nf, cols, kinds, ptrs, max(rows, 1), *nargs, C.byref(got),
                              C.byref(bad)))
    k = int(got.value)
    return {nm: o[:k] for (nm, _, _), o in zip(fields, outs)}, int(bad.value)


def iter_blocks(path, block_bytes=None):
    """blocks of whole lines of a plain or gzip-compressed text file"""
    path = str(path)
    block_bytes = block_bytes or BLOCK_BYTES
    opener = gzip.open if path.endswith(".gz") else open
    tail = b""
    with opener(path, "rb") as fh:
        while True:
            b = fh.read(block_bytes)
            if not b:
                break
            cut = b.rfind(b"\n")
            if cut < 0:
                tail += b
                continue
            yield tail + b[:cut + 1]
            tail = b[cut + 1:]
    if tail:
        yield tail


def _pump(gen, q):
    try:
        for x in gen:
            q.put(x)
        q.put(StopIteration)
    except BaseException as e:                     # noqa: BLE001 -- handed to the consumer
        q.put(e)


def prefetched(gen, depth=2):
    """runs `gen` in a daemon thread, `depth` items ahead of the consumer"""
    q = queue.Queue(maxsize=depth)
    th = threading.Thread(target=_pump, args=(gen, q), daemon=True)
    th.start()
    while True:
        x = q.get()
        if x is StopIteration:
            return
        if isinstance(x, BaseException):
            raise x
        yield x


def read_table_chunks(path, fields, names, threads=None, block_bytes=None):
    """path -> iterator of ({field: array}, malformed) per block; reading / inflating runs two blocks ahead of the parser,
    the parser one block ahead of the consumer"""
    nt = names if isinstance(names, NameTable) or names is None else NameTable(names)

    def parsed():
        for blk in prefetched(iter_blocks(path, block_bytes), 2):
            yield parse_block(blk, fields, nt, threads)
    return prefetched(parsed(), 1)


def read_table(path, fields, names, threads=None, block_bytes=None):
    """whole file -> ({field: array}, malformed)"""
    parts, bad = [], 0
    for cols, b in read_table_chunks(path, fields, names, threads, block_bytes):
        parts.append(cols)
        bad += b
    if not parts:
        return {nm: np.empty(0, dtype=_KIND_DTYPE[k]) for nm, _, k in fields}, bad
    return {nm: np.concatenate([p[nm] for p in parts]) for nm, _, _ in fields}, bad


# ------------------------------------------------------------------------------------------------------------------
# PQS directories (pqs.py:112-131: `_metadata`, `_contigsizes`, q0/*.parquet = every pair, q1/*.parquet = MAPQ >= 1)
# ------------------------------------------------------------------------------------------------------------------
def is_pqs(path):
    p = Path(path)
    if not p.is_dir() or not (p / "_metadata").exists() or not (p / "_contigsizes").exists():
        return False
    try:
        import ast
        meta = ast.literal_eval((p / "_metadata").read_text())
        return bool(meta.get("is_pqs"))
    except Exception:                              # noqa: BLE001
        return False


def pqs_chunks(path, min_mapq):
    """parquet chunk files the pairs extractor reads for this MAPQ cut (pqs.py:684-707): q1 when the store was written with
    MAPQ and the cut is >= 1, else q0"""
    import ast
    p = Path(path)
    meta = ast.literal_eval((p / "_metadata").read_text())
    with_mapq = bool(meta.get("is_with_mapq", True))
    sub = "q1" if (with_mapq and min_mapq >= 1) else "q0"
    files = sorted((p / sub).glob("*.parquet"), key=lambda f: (len(f.stem), f.stem))
    return files, with_mapq


def read_pqs_chunk(path, names_lut_fn, min_mapq, with_mapq):
    """one PQS chunk -> pairs columns (pqs.py:1184-1194: mapq >= min_mapq when the store carries MAPQ and the cut is > 1)"""
    import pyarrow.parquet as pq
    cols = ["chrom1", "pos1", "chrom2", "pos2"] + (["mapq"] if with_mapq else [])
    t = pq.read_table(path, columns=cols).to_pandas()
    if with_mapq and min_mapq > 1:
        t = t[t["mapq"] >= min_mapq]
    n = len(t)
    mq = t["mapq"].to_numpy().astype(np.int8) if with_mapq else np.full(n, 60, dtype=np.int8)
    return {"chrom1": names_lut_fn(t["chrom1"]), "chrom2": names_lut_fn(t["chrom2"]), "mapq": mq,
            "pos1": t["pos1"].to_numpy().astype(np.uint32), "pos2": t["pos2"].to_numpy().astype(np.uint32)}
