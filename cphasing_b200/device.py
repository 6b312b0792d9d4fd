"""Device-resident stages of the hyperpartition hot path: thin Python over the libcph C-ABI.

Every function takes / returns torch CUDA tensors; nothing here computes on the host beyond
O(n_contigs) bookkeeping (row maps, frac-bit choice).  The reference-facing classes in
``cphasing_b200.algorithms.hypergraph`` / ``cphasing_b200.hypergraph`` are built on these.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass

import numpy as np
import torch

from . import _lib
from ._lib import call_ws, check, lib, nvtx, ptr, stream


def _dev(device=None):
    if not torch.cuda.is_available():
        raise _lib.CphError("no CUDA device: cphasing_b200 has no CPU fallback")
    return torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")


def to_dev(x, dtype, device):
    """numpy / tensor -> contiguous device tensor of `dtype` (None passes through)."""
    if x is None:
        return None
    if isinstance(x, torch.Tensor):
        return x.to(device=device, dtype=dtype, non_blocking=True).contiguous()
    a = np.ascontiguousarray(x)
    want = {torch.uint32: np.uint32, torch.int32: np.int32, torch.int8: np.int8, torch.int64: np.int64,
            torch.float32: np.float32, torch.float64: np.float64, torch.uint8: np.uint8}[dtype]
    if a.dtype != want:
        a = a.astype(want)
    elif not a.flags.writeable:
        a = a.copy()       # torch.from_numpy on a read-only array (np.load mmap, pandas views) is undefined once written to
    return torch.from_numpy(a).to(device, non_blocking=True)


PINNED_D2H_MAX_BYTES = 4 << 30


def to_host(t: torch.Tensor) -> np.ndarray:
    """device tensor -> numpy array.  The bytes land in page-locked memory from torch's caching host allocator (full
    PCIe rate, and no first-touch page faults on a fresh malloc for every result); the array keeps the block alive."""
    if t.dtype == torch.uint32:
        t = t.view(torch.int32)
    if not t.is_cuda:
        return t.numpy()
    nb = t.numel() * t.element_size()
    if nb == 0 or nb > PINNED_D2H_MAX_BYTES:
        return t.cpu().numpy()
    h = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
    h.copy_(t.contiguous(), non_blocking=True)
    torch.cuda.current_stream().synchronize()
    return h.numpy()


@dataclass
class RawEdges:
    """RAW EDGE-MAJOR triplets (include/cph.h): the CSC view of H, hyperedge runs marked by a head bit."""
    row: torch.Tensor      # uint32 contig | head bit
    col: torch.Tensor      # uint32 hyperedge id
    count: torch.Tensor    # uint32
    mapq: torch.Tensor     # int8 (or None)
    nnz: int
    n_cols: int            # max col + 1
    n_contigs: int
    packed: bool = False   # count carries D_e of the entry's hyperedge in its top byte (D_e << 24 | count): what the warp
    #                        group-by kernel writes.  Together with `row` that is the packed incidence of the table itself.
    n_edges: int = 0       # hyperedges (runs) in the arrays when known, else 0
    min_order: int = 0     # smallest hyperedge order the extraction let through (0 = unknown)

    def trimmed(self):
        k = self.nnz
        return RawEdges(self.row[:k], self.col[:k], self.count[:k], None if self.mapq is None else self.mapq[:k],
                        k, self.n_cols, self.n_contigs, self.packed, self.n_edges, self.min_order)

    def plain_count(self):
        """the count column without the D_e byte"""
        if not self.packed:
            return self.count
        return (self.count.view(torch.int32) & 0x00FFFFFF).view(torch.uint32)


class DeviceH:
    """Filtered incidence matrix on the device.  The packed edge-major form (pem) is what the kernels
    consume; the CSR view (what scipy callers see) is built on first use (cph_incidence_csr)."""

    def __init__(self, pem_row, pem_pk, nnz, n_rows, n_cols, row_keep=None, rowsum=None, pem_col=None, value_bound=None):
        self.pem_row = pem_row      # uint32 [nnz]  new row | head bit
        self.pem_pk = pem_pk        # uint32 [nnz]  D_e<<24 | count
        self.nnz = int(nnz)
        self.n_rows = int(n_rows)
        self.n_cols = int(n_cols)
        self.row_keep = row_keep    # bool [n_contigs] which input contigs survived (None = all)
        self.rowsum = rowsum        # uint64-as-int64 [n_contigs] summed counts before the row filter
        self.pem_col = pem_col      # uint32 [nnz] hyperedge id per entry; only set when columns keep their ids
        self._bound = value_bound   # python int >= max_ij sum_e c_ie c_je  (for the fixed-point scale)
        self._csr = None

    @property
    def device(self):
        return self.pem_row.device

    def _build_csr(self):
        if self._csr is None:
            dev = self.device
            cap = max(self.nnz, 1)
            indptr = torch.empty(self.n_rows + 1, dtype=torch.int64, device=dev)
            indices = torch.empty(cap, dtype=torch.int32, device=dev)
            pk = torch.empty(cap, dtype=torch.uint32, device=dev)
            call_ws("cph_incidence_csr", dev, ptr(self.pem_row), ptr(self.pem_pk), ptr(self.pem_col), self.nnz,
                    max(self.n_rows, 1), ptr(indptr), ptr(indices), ptr(pk))
            self._csr = (indptr, indices[:self.nnz], pk[:self.nnz])
        return self._csr

    indptr = property(lambda s: s._build_csr()[0])      # int64 [n+1]
    indices = property(lambda s: s._build_csr()[1])     # int32 [nnz] (sorted inside each row)
    pk = property(lambda s: s._build_csr()[2])          # uint32 [nnz]

    def value_bound(self):
        if self._bound is None:
            rs = torch.empty(max(self.n_rows, 1), dtype=torch.int64, device=self.device)
            mx = torch.zeros(1, dtype=torch.int32, device=self.device)
            check(lib().cph_row_sums(ptr(self.pem_row), ptr(self.pem_pk), None, self.nnz, max(self.n_rows, 1), 0, None, 1,
                                     ptr(rs), ptr(mx), stream()))
            self._bound = int(rs.max().item()) * max(int(mx.item()), 1)
        return self._bound

    def counts_f32(self):
        out = torch.empty(max(self.nnz, 1), dtype=torch.float32, device=self.device)[:self.nnz]
        check(lib().cph_csr_counts_f32(ptr(self.pk), self.nnz, ptr(out), stream()))
        return out


@dataclass
class DeviceCSR:
    indptr: torch.Tensor    # int64 [n+1]
    indices: torch.Tensor   # int32 [nnz]
    data32: torch.Tensor    # float32 or None
    data64: torch.Tensor    # float64 or None
    n: int
    nnz: int


# ------------------------------------------------------------------------------------------------
# stage 1: tables -> raw edge-major
# ------------------------------------------------------------------------------------------------
@dataclass
class PreFilter:
    """extractor pre-filters (include/cph.h: cph_prefilter).  The contig column(s) then hold codes of the ORIGINAL contig
    names, 0 <= code < n_names; bed_* describe the HCR intervals per name (sorted by name, start), split_init the first
    split bin of every name."""
    n_names: int
    bed_ptr: object = None        # int64 [n_names + 1]
    bed_start: object = None      # uint32
    bed_end: object = None        # uint32
    hcr_invert: bool = False
    split_length: int = 0
    split_init: object = None     # uint32 [n_names], 0xffffffff = name not in the split map

    def struct(self, dev):
        """-> (ctypes struct, tensors to keep alive)"""
        keep = []

        def d(x, dt):
            if x is None:
                return None
            t = to_dev(x, dt, dev)
            keep.append(t)
            return t.data_ptr()
        hcr = self.bed_ptr is not None
        pf = _lib.Prefilter(d(self.bed_ptr, torch.int64) if hcr else None, d(self.bed_start, torch.uint32) if hcr else None,
                            d(self.bed_end, torch.uint32) if hcr else None, int(bool(self.hcr_invert)),
                            int(self.split_length or 0), d(self.split_init, torch.uint32) if self.split_length else None,
                            int(self.n_names))
        return pf, keep


@nvtx("cph:porec_to_edges")
def porec_to_edges(read_idx, contig_idx, mapq, n_contigs, start=None, end=None, contig_len=None, min_mapq=1,
                   min_order=2, max_order=50, edge_length=0, read_offset=0, device=None, prefilter: PreFilter = None) -> RawEdges:
    dev = _dev(device)
    if prefilter is not None:
        return _porec_to_edges_pf(read_idx, contig_idx, mapq, n_contigs, start, end, contig_len, min_mapq, min_order,
                                  max_order, edge_length, read_offset, dev, prefilter)
    if not edge_length:
        host = _host_columns(read_idx, contig_idx, mapq)
        if host is not None and host[0].numel() >= STREAM_MIN_RECORDS:
            raw, staged = _porec_to_edges_host(host, n_contigs, min_mapq, min_order, max_order, read_offset, dev)
            if raw is not None:
                return raw
            read_idx, contig_idx, mapq = staged      # table not grouped by read: the columns are already on the device
    r = to_dev(read_idx, torch.uint32, dev)
    c = to_dev(contig_idx, torch.uint32, dev)
    q = to_dev(mapq, torch.int8, dev)
    N = r.numel()
    if edge_length:
        st = to_dev(start, torch.uint32, dev); en = to_dev(end, torch.uint32, dev)
        ln = to_dev(contig_len, torch.int64, dev)
    else:
        st = en = ln = None
    cap = max(N, 1)
    o_row = torch.empty(cap, dtype=torch.uint32, device=dev)
    o_col = torch.empty(cap, dtype=torch.uint32, device=dev)
    o_cnt = torch.empty(cap, dtype=torch.uint32, device=dev)
    o_mq = torch.empty(cap, dtype=torch.int8, device=dev)
    h = (C.c_uint64 * 6)()
    call_ws("cph_porec_to_edges", dev, ptr(r), ptr(c), ptr(q), ptr(st), ptr(en), ptr(ln), N, int(n_contigs),
            int(min_mapq), int(min_order), int(max_order), int(edge_length or 0), int(read_offset), 0,
            ptr(o_row), ptr(o_col), ptr(o_cnt), ptr(o_mq), h)
    return RawEdges(o_row, o_col, o_cnt, o_mq, int(h[0]), int(h[1]), int(n_contigs), bool(h[5]), int(h[4]),
                    int(min_order)).trimmed()


def _porec_to_edges_pf(read_idx, contig_idx, mapq, n_contigs, start, end, contig_len, min_mapq, min_order, max_order,
                       edge_length, read_offset, dev, prefilter):
    r = to_dev(read_idx, torch.uint32, dev)
    c = to_dev(contig_idx, torch.uint32, dev)
    q = to_dev(mapq, torch.int8, dev)
    st = to_dev(start, torch.uint32, dev); en = to_dev(end, torch.uint32, dev)
    ln = to_dev(contig_len, torch.int64, dev) if (edge_length and not prefilter.split_length) else None
    if prefilter.split_length:
        edge_length = 0                      # hypergraph.py:1406-1424: the split branch has no edge filter
    N = r.numel()
    cap = max(N, 1)
    o_row = torch.empty(cap, dtype=torch.uint32, device=dev)
    o_col = torch.empty(cap, dtype=torch.uint32, device=dev)
    o_cnt = torch.empty(cap, dtype=torch.uint32, device=dev)
    o_mq = torch.empty(cap, dtype=torch.int8, device=dev)
    h = (C.c_uint64 * 6)()
    pf, keep = prefilter.struct(dev)
    call_ws("cph_porec_to_edges_pf", dev, ptr(r), ptr(c), ptr(q), ptr(st), ptr(en), ptr(ln), N, int(n_contigs),
            int(min_mapq), int(min_order), int(max_order), int(edge_length or 0), int(read_offset), 0, C.byref(pf),
            ptr(o_row), ptr(o_col), ptr(o_cnt), ptr(o_mq), h)
    del keep
    return RawEdges(o_row, o_col, o_cnt, o_mq, int(h[0]), int(h[1]), int(n_contigs), bool(h[5]), int(h[4]),
                    int(min_order)).trimmed()


STREAM_MIN_RECORDS = 1 << 22      # below this the chunked host path has nothing to overlap
STREAM_CHUNK_RECORDS = 8 << 20
STREAM_PACK_THREADS = -1          # host threads that re-encode the chunks to 4 B per record (0: copy the columns as they are)
STREAM_PACK_CHUNK_RECORDS = 0     # records per packed chunk (0: the library default, 1 Mi)
LAST_H2D_BYTES = 0                # bytes the last host-table extraction sent to the device


def _host_columns(read_idx, contig_idx, mapq):
    """(read u32, contig u32, mapq i8) as contiguous CPU tensors when all three are host arrays, else None."""
    want = (torch.uint32, torch.uint32, torch.int8)
    npw = (np.uint32, np.uint32, np.int8)
    out = []
    for x, td, nd in zip((read_idx, contig_idx, mapq), want, npw):
        if isinstance(x, torch.Tensor):
            if x.is_cuda:
                return None
            if x.dtype != td:
                same_width = x.element_size() == (1 if td == torch.int8 else 4)
                x = x.view(td) if same_width and not x.dtype.is_floating_point else x.to(td)
            out.append(x.contiguous())
        else:
            a = np.ascontiguousarray(x)
            if a.dtype != nd:
                a = a.astype(nd)
            out.append(torch.from_numpy(a))
    return out


def bind_host_to_gpu_node(device_index=None):
    """Pin this process (and, by first touch, the pinned staging buffers it allocates afterwards) to the CPUs of the NUMA
    node the GPU hangs off.  With one process per GPU on a two-socket host the chunked H2D copies otherwise cross the
    socket interconnect for half of the ranks.  Best effort: returns the node number, or None when the topology cannot be
    read or the affinity cannot be set (containers with a restricted cpuset)."""
    try:
        idx = torch.cuda.current_device() if device_index is None else int(device_index)
        p = torch.cuda.get_device_properties(idx)
        bus = f"{getattr(p, 'pci_domain_id', 0):04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
        with open(f"/sys/bus/pci/devices/{bus}/numa_node") as fh:
            node = int(fh.read().strip())
        if node < 0:
            return None
        with open(f"/sys/devices/system/node/node{node}/cpulist") as fh:
            cpus = set()
            for part in fh.read().strip().split(","):
                a, _, b = part.partition("-")
                cpus.update(range(int(a), int(b or a) + 1))
        allowed = cpus & set(os.sched_getaffinity(0))
        if not allowed:
            return None
        os.sched_setaffinity(0, allowed)
        return node
    except Exception:                                   # noqa: BLE001 -- topology files absent, no permission, ...
        return None


def _pack_threads():
    """host threads for the packed transfer: STREAM_PACK_THREADS, or (auto) the cores of this process less two.  The
    re-encoding is bound by host memory bandwidth (about 17 B of traffic per record against 9 B for the plain copy), so
    it only pays when one rank has the host to itself and enough cores: measured for 286 M records on a 16-core box,
    4 threads 82 ms, 8 threads 48 ms, 14 threads 33 ms against 47 ms unpacked; with two ranks on one host (11 threads
    each) 60.8 ms packed against 53.9 ms unpacked -> several ranks per node send the columns as they are."""
    if STREAM_PACK_THREADS >= 0:
        return int(STREAM_PACK_THREADS)
    if int(os.environ.get("LOCAL_WORLD_SIZE", "1")) > 1:
        return 0
    try:
        cores = len(os.sched_getaffinity(0))
    except AttributeError:
        cores = os.cpu_count() or 1
    t = min(16, cores - 2)
    return t if t >= 8 else 0


def _porec_to_edges_host(host, n_contigs, min_mapq, min_order, max_order, read_offset, dev):
    """chunked H2D copy overlapped with the group-by kernel (cph_porec_to_edges_host) -> (RawEdges, None), or
    (None, staged device columns) when the table is not grouped by read."""
    hr, hc, hq = host
    N = hr.numel()
    r = torch.empty(N, dtype=torch.uint32, device=dev)
    c = torch.empty(N, dtype=torch.uint32, device=dev)
    q = torch.empty(N, dtype=torch.int8, device=dev)
    o_row = torch.empty(N, dtype=torch.uint32, device=dev)
    o_col = torch.empty(N, dtype=torch.uint32, device=dev)
    o_cnt = torch.empty(N, dtype=torch.uint32, device=dev)
    o_mq = torch.empty(N, dtype=torch.int8, device=dev)
    h = (C.c_uint64 * 6)()
    nt = _pack_threads()
    call_ws("cph_porec_to_edges_host", dev, ptr(hr), ptr(hc), ptr(hq), ptr(r), ptr(c), ptr(q), N, int(n_contigs),
            int(min_mapq), int(min_order), int(max_order), int(read_offset),
            int(STREAM_PACK_CHUNK_RECORDS if nt else STREAM_CHUNK_RECORDS), nt,
            ptr(o_row), ptr(o_col), ptr(o_cnt), ptr(o_mq), h)
    global LAST_H2D_BYTES
    LAST_H2D_BYTES = int(h[3])
    if int(h[2]):
        return None, (r, c, q)
    return RawEdges(o_row, o_col, o_cnt, o_mq, int(h[0]), int(h[1]), int(n_contigs), bool(h[5]), int(h[4]),
                    int(min_order)).trimmed(), None


@nvtx("cph:pairs_to_edges")
def pairs_to_edges(chrom1, chrom2, mapq, n_contigs, pos1=None, pos2=None, contig_len=None, min_mapq=1,
                   edge_length=0, device=None, prefilter: PreFilter = None) -> RawEdges:
    dev = _dev(device)
    c1 = to_dev(chrom1, torch.uint32, dev)
    c2 = to_dev(chrom2, torch.uint32, dev)
    q = to_dev(mapq, torch.int8, dev)
    P = c1.numel()
    if prefilter is not None:
        p1 = to_dev(pos1, torch.uint32, dev); p2 = to_dev(pos2, torch.uint32, dev)
        ln = to_dev(contig_len, torch.int64, dev) if edge_length else None
    elif edge_length:
        p1 = to_dev(pos1, torch.uint32, dev); p2 = to_dev(pos2, torch.uint32, dev)
        ln = to_dev(contig_len, torch.int64, dev)
    else:
        p1 = p2 = ln = None
    cap = max(2 * P, 2)
    o_row = torch.empty(cap, dtype=torch.uint32, device=dev)
    o_col = torch.empty(cap, dtype=torch.uint32, device=dev)
    o_cnt = torch.empty(cap, dtype=torch.uint32, device=dev)
    o_mq = torch.empty(cap, dtype=torch.int8, device=dev)
    h = (C.c_uint64 * 2)()
    if prefilter is not None:
        pf, keep = prefilter.struct(dev)
        call_ws("cph_pairs_to_edges_pf", dev, ptr(c1), ptr(c2), ptr(q), ptr(p1), ptr(p2), ptr(ln), P, int(n_contigs),
                int(min_mapq), int(edge_length or 0), C.byref(pf), ptr(o_row), ptr(o_col), ptr(o_cnt), ptr(o_mq), h)
        del keep
    else:
        call_ws("cph_pairs_to_edges", dev, ptr(c1), ptr(c2), ptr(q), ptr(p1), ptr(p2), ptr(ln), P, int(n_contigs),
                int(min_mapq), int(edge_length or 0), ptr(o_row), ptr(o_col), ptr(o_cnt), ptr(o_mq), h)
    # every surviving pair is a hyperedge of two entries with count 1: the kernel writes D_e = 1 into the count's top byte
    return RawEdges(o_row, o_col, o_cnt, o_mq, int(h[0]), int(h[1]), int(n_contigs), True, int(h[0]) // 2, 2).trimmed()


def edges_to_triplets(e: RawEdges):
    """-> (row i32, col i32, count u32, mapq i8) device tensors in HyperEdges wire order (row, col)."""
    dev = e.row.device
    k = e.nnz
    row = torch.empty(k, dtype=torch.int32, device=dev)
    col = torch.empty(k, dtype=torch.int32, device=dev)
    cnt = torch.empty(k, dtype=torch.uint32, device=dev)
    mq = torch.empty(k, dtype=torch.int8, device=dev) if e.mapq is not None else None
    call_ws("cph_edges_to_triplets", dev, ptr(e.row), ptr(e.col), ptr(e.plain_count()), ptr(e.mapq), k, int(e.n_contigs),
            ptr(row), ptr(col), ptr(cnt), ptr(mq))
    return row, col, cnt, mq


def triplets_to_edges(row, col, count, mapq, n_contigs, n_cols=None, device=None) -> RawEdges:
    """HyperEdges wire triplets (any order) -> raw edge-major."""
    dev = _dev(device)
    r = to_dev(row, torch.int32, dev)
    c = to_dev(col, torch.int32, dev)
    cnt = to_dev(count, torch.uint32, dev)
    mq = to_dev(mapq, torch.int8, dev) if mapq is not None and len(mapq) == len(row) else None
    k = r.numel()
    if n_cols is None:
        n_cols = int(c.max().item()) + 1 if k else 1
    o_row = torch.empty(max(k, 1), dtype=torch.uint32, device=dev)
    o_col = torch.empty(max(k, 1), dtype=torch.uint32, device=dev)
    o_cnt = torch.empty(max(k, 1), dtype=torch.uint32, device=dev)
    o_mq = torch.empty(max(k, 1), dtype=torch.int8, device=dev) if mq is not None else None
    call_ws("cph_triplets_to_edges", dev, ptr(r), ptr(c), ptr(cnt), ptr(mq), k, int(max(n_cols, 1)),
            ptr(o_row), ptr(o_col), ptr(o_cnt), ptr(o_mq))
    return RawEdges(o_row, o_col, o_cnt, o_mq, k, int(n_cols), int(n_contigs)).trimmed()


# ------------------------------------------------------------------------------------------------
# stage 1c: filters + incidence matrix
# ------------------------------------------------------------------------------------------------
def row_sums(e, min_mapq=0, row_drop=None, use_mapq=True, packed=False):
    """per-contig summed count (uint64 bits in an int64 tensor): matrix.sum(axis=1) of
    algorithms/hypergraph.py:287 over the entries get_data (:175-184) / remove_rows (:207-223) keep.
    Returns (rowsum int64[n], max_count int32[1])."""
    dev = e.row.device
    out = torch.empty(e.n_contigs, dtype=torch.int64, device=dev)
    mx = torch.zeros(1, dtype=torch.int32, device=dev)
    mq = e.mapq if (use_mapq and min_mapq > 0 and e.mapq is not None) else None
    check(lib().cph_row_sums(ptr(e.row), ptr(e.count), ptr(mq), e.nnz, e.n_contigs, int(min_mapq), ptr(row_drop),
                             int(packed or getattr(e, "packed", False)), ptr(out), ptr(mx), stream()))
    return out, mx


@nvtx("cph:incidence")
def incidence(e: RawEdges, min_contacts=1, min_mapq=0, row_drop=None, mapq_filter=True, rowsum_hook=None) -> DeviceH:
    """HyperGraph.get_data + remove_rows + incidence_matrix (algorithms/hypergraph.py:162-314).

    row_drop: uint8[n_contigs] flags of contigs removed by remove_rows (or None).
    rowsum_hook(rowsum, max_count) -> (rowsum, max_count): applied before the min_contacts test (the multi-GPU
    path all-reduces both across concatemer shards here).
    """
    dev = e.row.device
    n = e.n_contigs
    use_mq = bool(mapq_filter and min_mapq > 0 and e.mapq is not None and e.nnz > 1)   # :175
    rs, mx = row_sums(e, min_mapq, row_drop, use_mq)
    if rowsum_hook is not None:
        rs, mx = rowsum_hook(rs, mx)
    if min_contacts:
        keep = rs >= int(np.ceil(min_contacts))                                          # :287
        col_min = 2                                                                      # :305
    else:
        keep = torch.ones(n, dtype=torch.bool, device=dev)
        col_min = 0
    if row_drop is not None:
        # removed rows stay in the matrix as empty rows unless min_contacts drops them (:207-223 keeps nodes)
        live = row_drop == 0
    else:
        live = None
    new_id = torch.cumsum(keep.to(torch.int32), 0, dtype=torch.int32) - 1
    row_map = torch.where(keep if live is None else (keep & live), new_id, torch.full_like(new_id, -1))
    # one host round trip for the three scalars the host needs
    sc = torch.stack((keep.sum(), rs.max() if n else rs.sum(), mx[0].to(torch.int64))).cpu()
    n_rows, bound = int(sc[0]), int(sc[1]) * max(int(sc[2]), 1)
    if int(sc[2]) > 0x00FFFFFF:
        raise _lib.CphError("an alignment count >= 2^24 cannot be represented exactly (the reference stores counts "
                            "as float32)")
    if (e.packed and e.n_edges and not use_mq and row_drop is None and n_rows == n and e.nnz
            and col_min > 0 and e.min_order >= 2 and os.environ.get("CPH_INCIDENCE_FAST", "1") != "0"):
        # Nothing to drop: no entry filter, every contig passes min_contacts, every hyperedge holds >= 2 entries of count
        # >= 1 (so its summed count passes the column filter) -- the group-by kernel's own output already IS the packed
        # incidence (row | head, D_e << 24 | count, hyperedges numbered in order): no second pass over the entries.
        H = DeviceH(e.row, e.count, e.nnz, n, e.n_edges, keep, rs)
    else:
        H = _incidence_build(e.row, e.col, e.count, e.mapq if use_mq else None, e.nnz, n, min_mapq if use_mq else 0,
                             row_map, n_rows, col_min, bool(e.packed), keep, rs, e.n_cols)
    H._bound = bound
    return H


def _incidence_build(row, col, count, mapq, nnz, n_contigs, min_mapq, row_map, n_rows, col_min, packed, keep, rs,
                     n_cols_in):
    dev = row.device
    cap = max(nnz, 1)
    pem_row = torch.empty(cap, dtype=torch.uint32, device=dev)
    pem_pk = torch.empty(cap, dtype=torch.uint32, device=dev)
    h = (C.c_uint64 * 2)()
    call_ws("cph_incidence_build", dev, ptr(row), ptr(col), ptr(count), ptr(mapq), nnz, int(n_contigs), int(min_mapq),
            ptr(row_map), int(n_rows), int(col_min), int(packed), ptr(pem_row), ptr(pem_pk), None, None, None, h)
    k = int(h[0])
    n_cols = int(h[1]) if col_min > 0 else int(n_cols_in)
    return DeviceH(pem_row[:k], pem_pk[:k], k, int(n_rows), n_cols, keep, rs)


@nvtx("cph:sub_incidence")
def sub_incidence(H: DeviceH, idx, col_min=2) -> DeviceH:
    """extract_incidence_matrix2 (algorithms/hypergraph.py:938-954): rows `idx` (in that order) of H,
    hyperedges whose remaining summed count <= 1 dropped and renumbered."""
    dev = H.device
    idx_t = to_dev(np.asarray(idx, dtype=np.int64), torch.int64, dev)
    row_map = torch.full((H.n_rows,), -1, dtype=torch.int32, device=dev)
    row_map[idx_t] = torch.arange(idx_t.numel(), dtype=torch.int32, device=dev)
    sub = _incidence_build(H.pem_row, None, H.pem_pk, None, H.nnz, H.n_rows, 0, row_map, int(idx_t.numel()), col_min,
                           True, None, None, H.n_cols)
    sub.column_shard = getattr(H, "column_shard", False)      # rows of a shard are a shard (the column filter is column-local)
    return sub


# ------------------------------------------------------------------------------------------------
# stage 2: clique expansion
# ------------------------------------------------------------------------------------------------
def frac_bits_for(bound):
    """largest fixed-point scale such that sums bounded by `bound` stay below 2^61."""
    return max(0, min(52, 61 - int(bound + 1).bit_length()))


def choose_frac_bits(H: DeviceH):
    b = H.value_bound()
    return frac_bits_for(b), b


@dataclass
class Accumulator:
    acc: torch.Tensor       # int64 (uint64 bits) packed upper triangle
    diag: torch.Tensor      # int64 (uint64 bits) [n]
    n: int
    frac_bits: int
    use_high_order: bool


BINNED_MIN_N = 5600          # n (n + 1) / 2 x 8 B beyond ~126 MB: the accumulator no longer stays in L2
BINNED_MAX_N = 46 * 2508     # 46 block rows of the binned kernels
BINNED_MAX_RECORD_BYTES = 16 << 30   # scratch of 8 B per pair; beyond this the direct kernel runs instead
SWEEP_MAX_N = 1 << 20        # row field of the 4-byte packed entry (expand.cu sw_prep_kernel)
BINNED_MIN_NNZ = 1 << 22     # small inputs: the extra passes cost more than the DRAM read-modify-writes they avoid


def sweep_bands(n, band_bytes=None, mean_order=4.0):
    """row bands [r0[k], r0[k+1]) of the packed upper triangle with at most ~band_bytes of u64 cells each.
    Measured on B200 (scripts/diag_swept.py, profiles/r2f_swept_*.txt): with few pairs per pass (workload D, mean order 4)
    128 MB bands are fastest -- fewer streaming passes, the 126 MB L2 still absorbs the REDs; with many pairs per pass
    (workload E, mean order 8.5) 96 MB bands are, 128 MB thrashes."""
    if band_bytes is None:
        mb = os.environ.get("CPH_BAND_MB", "")
        band_bytes = int((float(mb) if mb else (128.0 if mean_order <= 6.0 else 96.0)) * (1 << 20))
    cells = max(band_bytes // 8, n)          # at least one row
    # cells before row i: i (2n - i + 1) / 2; walk the rows with numpy (n <= 2^20)
    i = np.arange(n + 1, dtype=np.int64)
    before = i * (2 * n - i + 1) // 2
    r0 = [0]
    while r0[-1] < n:
        nxt = int(np.searchsorted(before, before[r0[-1]] + cells, side="right")) - 1
        r0.append(min(n, max(nxt, r0[-1] + 1)))
    return r0


class LocalConsensus:
    """Choices of the accumulate stage that every rank of a column-shard job must make alike, because they decide which
    collectives follow and how large they are: direct or swept form, the band size, and whether the packed 4-byte form
    fits.  One process agrees with itself; parallel.RankConsensus is the several-rank form (all-reduce SUM / MAX)."""
    world_size = 1

    def sums(self, *vals):
        return tuple(int(v) for v in vals)

    def any(self, flag):
        return bool(flag)


_LOCAL = LocalConsensus()


def expand_accumulate_swept(H: DeviceH, use_high_order, frac_bits, acc, diag, zero, band_hook=None, consensus=None,
                            mean_order=None):
    """sorted 4-byte entries + one pass per L2-sized row band (expand.cu, swept form).  `band_hook(k, lo, hi, cell_lo,
    cell_hi)` runs after band k was launched (multi-GPU: start that band's exchange).  -> False when the packed form
    does not fit (the caller falls back).  The prep kernel's flag is read (one 4-byte copy) BEFORE any band is launched:
    a flagged run leaves pe4 slots unwritten, and the band kernel must never walk those.  With several ranks the flag is
    MAX-reduced first, so either every rank sweeps or none does."""
    dev = H.device
    n = H.n_rows
    cons = consensus or _LOCAL
    pe4 = torch.empty(max(H.nnz, 1), dtype=torch.int32, device=dev)
    flag = torch.empty(1, dtype=torch.int32, device=dev)
    rc = lib().cph_expand_sweep_prep(ptr(H.pem_row), ptr(H.pem_pk), H.nnz, n, int(use_high_order), ptr(pe4), ptr(flag), stream())
    if rc == _lib.ERR_TOO_LARGE:
        return False
    check(rc)
    if cons.any(int(flag.item()) != 0):
        return False
    if mean_order is None:
        mean_order = H.nnz / max(H.n_cols, 1)
    r0 = sweep_bands(n, mean_order=mean_order)
    for k in range(len(r0) - 1):
        lo, hi = r0[k], r0[k + 1]
        check(lib().cph_expand_sweep_band(ptr(pe4), H.nnz, n, lo, hi, int(frac_bits), int(zero), ptr(acc), ptr(diag), stream()))
        if band_hook is not None:
            band_hook(k, lo, hi, lo * (2 * n - lo + 1) // 2, hi * (2 * n - hi + 1) // 2)
    return True


@nvtx("cph:expand_accumulate")
def expand_accumulate(H: DeviceH, use_high_order=True, entry_w=None, csr_entry_w=None, frac_bits=None,
                      acc=None, band_hook=None, acc_out=None, consensus=None) -> Accumulator:
    """fixed-point accumulation of A = H diag(d) H^T (algorithms/hypergraph.py:385-446).  `acc`: add onto an existing
    triangle; `acc_out`: caller-owned storage for a fresh one; `band_hook`: see expand_accumulate_swept; `consensus`:
    how the shard-dependent choices (form, band size, packed-form flag) are made the same on every rank (LocalConsensus)."""
    dev = H.device
    n = H.n_rows
    if frac_bits is None:
        b = H.value_bound()
        if entry_w is not None and H.nnz:
            b = int(np.ceil(b * max(float(entry_w.max().item()), 1.0)))
        frac_bits = frac_bits_for(b)
        if not use_high_order and entry_w is None:
            frac_bits = 0            # pure integer contact counts: bit-exact
    tri = n * (n + 1) // 2
    zero = acc is None
    if acc is None:
        acc = acc_out if acc_out is not None else torch.empty(max(tri, 1), dtype=torch.int64, device=dev)
    diag = torch.empty(max(n, 1), dtype=torch.int64, device=dev)
    form = os.environ.get("CPH_EXPAND_FORM", "")          # "", "swept", "binned", "direct": A/B switch for measurements
    cons = consensus or _LOCAL
    big, mean_order = False, None
    if entry_w is None and (n >= BINNED_MIN_N or form == "swept"):     # a condition every rank evaluates alike
        g_nnz, g_cols = cons.sums(H.nnz, H.n_cols)                     # shard sizes differ: decide on the job's totals
        mean_order = g_nnz / max(g_cols, 1)
        big = form == "swept" or g_nnz // cons.world_size >= BINNED_MIN_NNZ
    if big and form != "direct":
        # the packed triangle does not fit in L2: update it region by region (expand.cu: "swept" = sorted 4-byte entries,
        # one streaming pass per row band; "binned" = the round-1 pair records grouped by 2-D block)
        if form != "binned" and n <= SWEEP_MAX_N:
            if not zero:
                diag.zero_()
            if expand_accumulate_swept(H, use_high_order, frac_bits, acc, diag, zero, band_hook, cons, mean_order):
                return Accumulator(acc, diag, n, int(frac_bits), bool(use_high_order))
            if band_hook is not None:
                band_hook(-1, 0, 0, 0, 0)        # abandoned: the direct kernel below starts over
            if not zero:
                raise _lib.CphError("swept expansion on top of an existing accumulator: entries do not fit the packed form")
        elif n <= BINNED_MAX_N:
            args = (ptr(H.pem_row), ptr(H.pem_pk), H.nnz, n, int(use_high_order), int(frac_bits), int(zero), ptr(acc), ptr(diag))
            fn = lib().cph_expand_accumulate_binned
            need = C.c_size_t(0)
            hp = C.c_uint64(0)
            check(fn(1, *args, None, C.byref(hp), None, 0, C.byref(need), None))
            ws = _lib.workspace(need.value, dev)
            check(fn(1, *args, None, C.byref(hp), ws.data_ptr(), ws.numel(), C.byref(need), stream()))
            if int(hp.value) * 8 <= BINNED_MAX_RECORD_BYTES:
                rec = torch.empty(max(int(hp.value), 1), dtype=torch.int64, device=dev)
                check(fn(2, *args, ptr(rec), C.byref(hp), ws.data_ptr(), ws.numel(), C.byref(need), stream()))
                return Accumulator(acc, diag, n, int(frac_bits), bool(use_high_order))
        # the packed / record form does not fit: the direct kernel below redoes the (cheap) zeroing and the diagonal
    check(lib().cph_expand_accumulate_diag(ptr(H.pem_row), ptr(H.pem_pk), ptr(entry_w), H.nnz, max(n, 1),
                                           int(use_high_order), int(frac_bits), int(zero), ptr(acc), ptr(diag), stream()))
    return Accumulator(acc, diag, n, int(frac_bits), bool(use_high_order))


@nvtx("cph:expand_emit")
def expand_emit(a: Accumulator, min_weight=0.1, node_w=None, want32=True, want64=True) -> DeviceCSR:
    """threshold (:453-455), optional w_i w_j normalisation (:590-602), CSR emit."""
    dev = a.acc.device
    n = a.n
    need = C.c_size_t(0)
    hn = C.c_uint64(0)
    fn = lib().cph_expand_count
    check(fn(ptr(a.acc), ptr(a.diag), n, a.frac_bits, float(min_weight), C.byref(hn), None, 0, C.byref(need), None))
    ws = torch.empty(need.value, dtype=torch.uint8, device=dev)     # private: emit reads it back
    check(fn(ptr(a.acc), ptr(a.diag), n, a.frac_bits, float(min_weight), C.byref(hn), ptr(ws), ws.numel(),
             C.byref(need), stream()))
    nnz = int(hn.value)
    indptr = torch.empty(n + 1, dtype=torch.int64, device=dev)
    indices = torch.empty(max(nnz, 1), dtype=torch.int32, device=dev)
    d32 = torch.empty(max(nnz, 1), dtype=torch.float32, device=dev) if want32 else None
    d64 = torch.empty(max(nnz, 1), dtype=torch.float64, device=dev) if want64 else None
    nw = to_dev(node_w, torch.float64, dev) if node_w is not None else None
    check(lib().cph_expand_emit(ptr(a.acc), ptr(a.diag), n, a.frac_bits, float(min_weight), ptr(nw), ptr(ws),
                                ptr(indptr), ptr(indices), ptr(d32), ptr(d64), stream()))
    return DeviceCSR(indptr, indices[:nnz], None if d32 is None else d32[:nnz], None if d64 is None else d64[:nnz],
                     n, nnz)


def clique_expansion(H: DeviceH, use_high_order=True, min_weight=0.1, node_w=None, entry_w=None, csr_entry_w=None,
                     want32=True, want64=True):
    acc = expand_accumulate(H, use_high_order, entry_w, csr_entry_w)
    return expand_emit(acc, min_weight, node_w, want32, want64), acc
