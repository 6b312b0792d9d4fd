"""`.hg` files (cphasing_b200/hgio.py): the msgpack document must be what the reference's encoder writes -- checked by
decoding it with msgspec into the reference's layout and, when the reference install is present, with the reference's own
HyperEdges + numpy_dec_hook (algorithms/hypergraph.py:66-123) -- and the binary sidecar must round-trip and go stale safely."""
import os
import sys
import time

import msgspec
import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def make_edges(nnz=50_000, seed=0, with_mapq=True):
    from cphasing_b200.algorithms.hypergraph import HyperEdges
    rng = np.random.default_rng(seed)
    idx = {f"ctg{i}": i for i in range(300)}
    sizes = {f"ctg{i}": int(rng.integers(10_000, 5_000_000)) for i in range(300)}
    row = np.sort(rng.integers(0, 300, nnz)).astype(np.int32)
    col = rng.integers(0, 3_000_000, nnz).astype(np.int32)
    col[:5] = [0, 127, 128, 65535, 65536]                               # representation boundaries
    cnt = rng.geometric(0.7, nnz).astype(np.uint32)
    cnt[:3] = [255, 256, 70000]
    mq = rng.integers(0, 61, nnz).astype(np.int8) if with_mapq else np.zeros(0, np.int8)
    return HyperEdges(idx=idx, row=row, col=col, count=cnt, contigsizes=sizes, mapq=mq)


@pytest.mark.parametrize("with_mapq", [True, False])
def test_msgpack_document_is_the_reference_layout(tmp_path, with_mapq):
    from cphasing_b200 import hgio
    from cphasing_b200.algorithms.hypergraph import HyperEdges, numpy_enc_hook, numpy_dec_hook
    e = make_edges(with_mapq=with_mapq)
    p = tmp_path / "x.hg"
    hgio.save(e, str(p))
    # byte for byte what the reference's own call writes (hypergraph.py:1909: msgspec.msgpack.encode(edges, enc_hook=...))
    assert p.read_bytes() == msgspec.msgpack.encode(e, enc_hook=numpy_enc_hook)
    d = msgspec.msgpack.decode(p.read_bytes(), type=HyperEdges, dec_hook=numpy_dec_hook)
    for f in ("row", "col", "count", "mapq"):
        assert np.array_equal(np.asarray(getattr(d, f)), np.asarray(getattr(e, f)))
    assert d.idx == e.idx and d.contigsizes == e.contigsizes
    sys.path.insert(0, os.path.join(ROOT, "baseline"))
    import ref_loader
    if ref_loader.available():            # the reference's own loader (cli.py:4811-4812)
        rah = ref_loader.load()
        r = msgspec.msgpack.decode(p.read_bytes(), type=rah.HyperEdges, dec_hook=rah.numpy_dec_hook)
        r.to_numpy()
        assert np.array_equal(r.row, e.row) and np.array_equal(r.count, e.count) and r.idx == e.idx


def test_sidecar_round_trip_and_staleness(tmp_path):
    from cphasing_b200 import hgio
    e = make_edges(seed=3)
    p = str(tmp_path / "y.hg")
    hgio.save(e, p)
    assert os.path.exists(p + ".bin")
    side = hgio.load_sidecar(p + ".bin", p)
    assert side is not None
    got = hgio.load(p)
    for f, dt in hgio.FIELDS:
        a = np.asarray(getattr(got, f))
        assert a.dtype == dt and np.array_equal(a, np.asarray(getattr(e, f)))
    assert got.idx == e.idx and got.contigsizes == e.contigsizes
    # somebody (the reference) rewrites the .hg: the sidecar no longer describes it and must be ignored
    e2 = make_edges(nnz=1000, seed=4)
    time.sleep(0.01)
    hgio.save(e2, p, sidecar=False)
    assert hgio.load_sidecar(p + ".bin", p) is None
    got2 = hgio.load(p)
    assert len(got2.row) == 1000 and np.array_equal(got2.col, e2.col)
    # a bare sidecar handed over as the hypergraph file
    hgio.save(e, str(tmp_path / "z.hg"), msgpack=False)
    got3 = hgio.load(str(tmp_path / "z.hg.bin"))
    assert np.array_equal(got3.row, e.row)
