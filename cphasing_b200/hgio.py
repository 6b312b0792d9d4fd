"""``.hg`` files: the reference's msgpack document plus a binary sidecar of the raw columns.

The reference stores a HyperEdges struct as msgpack with every array turned into a Python list of ints
(algorithms/hypergraph.py:66-79, hypergraph.py:1302-1306, 1907-1911; read back at cli.py:4796-4834): measured here
3.3 s to write and 8.0 s to read back per 8e7 ints (tens of GB of boxed ints in between), against 0.1 s each way for the
raw columns.  Here

* ``save(edges, path)`` writes the SAME msgpack document (a map ``idx, row, col, count, contigsizes, mapq`` whose arrays
  are msgpack int arrays -- the reference's own loader reads it; ``msgpack=False`` skips it) and next to it
  ``<path>.bin``: a small header and the four columns as raw little-endian arrays.
* ``load(path)`` prefers the sidecar when it is present and belongs to that msgpack file (size + mtime recorded in the
  header), and falls back to decoding the msgpack document (files written by the reference).

Sidecar layout (all little-endian):
    magic   8 B   b"CPHHG\\x01\\x00\\x00"
    u64     nnz
    u64     size of the msgpack file it accompanies (0 = unknown)
    u64     mtime_ns of that file (0 = unknown)
    u64     length L of the metadata block
    L B     msgpack map {"idx": {...}, "contigsizes": {...}}
    pad to 64 B, then row i32[nnz], col i32[nnz], count u32[nnz], mapq i8[nnz], each padded to 64 B
"""
from __future__ import annotations

import os
import struct

import msgspec
import numpy as np

MAGIC = b"CPHHG\x01\x00\x00"
FIELDS = (("row", np.int32), ("col", np.int32), ("count", np.uint32), ("mapq", np.int8))


def encode_msgpack(edges) -> bytes:
    """the reference's own document (hypergraph.py:1909): msgspec encodes the Struct as a map in field order and
    numpy_enc_hook turns every array into a list of ints.  Measured at 8e7 ints: 3.3 s (a numpy re-implementation of the int
    encoder was tried and was 3x slower than msgspec's C loop, so it is not used); the sidecar below costs 0.1 s."""
    from .algorithms.hypergraph import numpy_enc_hook
    return msgspec.msgpack.encode(edges, enc_hook=numpy_enc_hook)


def _pad(fh, align=64):
    r = fh.tell() % align
    if r:
        fh.write(b"\0" * (align - r))


def save(edges, path, sidecar=True, msgpack=True):
    """write `path` (msgpack, what the reference reads) and `path + '.bin'` (what `load` prefers)"""
    cols = {f: np.ascontiguousarray(np.asarray(getattr(edges, f)), dtype=dt) for f, dt in FIELDS}
    if len(cols["mapq"]) not in (0, len(cols["row"])):
        raise ValueError("mapq must be empty or as long as row")
    size = mtime = 0
    if msgpack:
        with open(path, "wb") as fh:
            fh.write(encode_msgpack(edges))
        st = os.stat(path)
        size, mtime = st.st_size, st.st_mtime_ns
    if sidecar:
        meta = msgspec.msgpack.encode({"idx": {str(k): int(v) for k, v in edges.idx.items()},
                                       "contigsizes": {str(k): int(v) for k, v in edges.contigsizes.items()},
                                       "n_mapq": int(len(cols["mapq"]))})
        with open(str(path) + ".bin", "wb") as fh:
            fh.write(MAGIC)
            fh.write(struct.pack("<QQQQ", len(cols["row"]), size, mtime, len(meta)))
            fh.write(meta)
            for f, _ in FIELDS:
                _pad(fh)
                cols[f].tofile(fh)


def load_sidecar(path, msgpack_path=None):
    """-> HyperEdges from `<path>` (a .bin sidecar) or None when it does not belong to msgpack_path"""
    from .algorithms.hypergraph import HyperEdges
    with open(path, "rb") as fh:
        if fh.read(8) != MAGIC:
            return None
        nnz, size, mtime, L = struct.unpack("<QQQQ", fh.read(32))
        if msgpack_path is not None and os.path.exists(msgpack_path) and size:
            st = os.stat(msgpack_path)
            if st.st_size != size or st.st_mtime_ns != mtime:
                return None                      # the .hg was rewritten by someone else: the sidecar is stale
        meta = msgspec.msgpack.decode(fh.read(L))
        out = {}
        for f, dt in FIELDS:
            r = fh.tell() % 64
            if r:
                fh.seek(64 - r, 1)
            k = int(meta.get("n_mapq", nnz)) if f == "mapq" else nnz
            out[f] = np.fromfile(fh, dtype=dt, count=k)
            if len(out[f]) != k:
                return None
    return HyperEdges(idx=meta["idx"], row=out["row"], col=out["col"], count=out["count"], contigsizes=meta["contigsizes"],
                      mapq=out["mapq"])


def load(path):
    """`.hg` -> HyperEdges: the binary sidecar when present and current, else the msgpack document (cli.py:4811-4834)"""
    from .algorithms.hypergraph import HyperEdges, numpy_dec_hook
    side = str(path) + ".bin"
    if os.path.exists(side):
        e = load_sidecar(side, path)
        if e is not None:
            return e
    with open(path, "rb") as fh:
        head = fh.read(8)
        fh.seek(0)
        if head == MAGIC:                        # a bare sidecar handed over as the hypergraph
            e = load_sidecar(path)
            if e is not None:
                return e
        e = msgspec.msgpack.decode(fh.read(), type=HyperEdges, dec_hook=numpy_dec_hook)
    e.to_numpy()
    return e
