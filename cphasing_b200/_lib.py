"""ctypes binding of libcph.so (include/cph.h) -- the only door from Python to the CUDA kernels.

There is NO CPU fallback: if the library is missing or a call fails, the caller gets an exception.
PyTorch owns every device buffer; this module only passes raw pointers, sizes and the current
stream through the C-ABI.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libcph.so")

CPH_OK = 0
ERR_TOO_LARGE = 4      # CPH_ERR_TOO_LARGE (include/cph.h)
HEAD_BIT = 0x80000000
ROW_MASK = 0x7FFFFFFF
COUNT_MASK = 0x00FFFFFF


class CphError(RuntimeError):
    pass


_lib = None

_P = C.c_void_p
_U64 = C.c_uint64
_U32 = C.c_uint32
_I = C.c_int
_F = C.c_float
_D = C.c_double
_SZ = C.c_size_t
_WS = [_P, _SZ, C.POINTER(_SZ), _P]          # ws, ws_bytes, h_ws_need, stream

class Prefilter(C.Structure):
    """cph_prefilter of include/cph.h (device pointers; 0 / NULL = off)."""
    _fields_ = [("bed_ptr", _P), ("bed_start", _P), ("bed_end", _P), ("hcr_invert", _I), ("split_length", _U64),
                ("split_init", _P), ("n_names", _U32)]


_PF = C.POINTER(Prefilter)

_SIGS = {
    "cph_version": (C.c_int, []),
    "cph_last_error": (C.c_char_p, []),
    "cph_launch_count": (_U64, []),
    "cph_pairs_to_edges": (_I, [_P, _P, _P, _P, _P, _P, _U64, _U32, _I, _U64, _P, _P, _P, _P, C.POINTER(_U64)] + _WS),
    "cph_porec_to_edges": (_I, [_P, _P, _P, _P, _P, _P, _U64, _U32, _I, _U32, _U32, _U64, _U32, _I,
                                _P, _P, _P, _P, C.POINTER(_U64)] + _WS),
    "cph_pairs_to_edges_pf": (_I, [_P, _P, _P, _P, _P, _P, _U64, _U32, _I, _U64, _PF, _P, _P, _P, _P, C.POINTER(_U64)] + _WS),
    "cph_porec_to_edges_pf": (_I, [_P, _P, _P, _P, _P, _P, _U64, _U32, _I, _U32, _U32, _U64, _U32, _I, _PF,
                                   _P, _P, _P, _P, C.POINTER(_U64)] + _WS),
    "cph_porec_to_edges_host": (_I, [_P, _P, _P, _P, _P, _P, _U64, _U32, _I, _U32, _U32, _U32, _U64, _I,
                                     _P, _P, _P, _P, C.POINTER(_U64)] + _WS),
    "cph_release_host_staging": (_I, []),
    "cph_edges_to_triplets": (_I, [_P, _P, _P, _P, _U64, _U32, _P, _P, _P, _P] + _WS),
    "cph_triplets_to_edges": (_I, [_P, _P, _P, _P, _U64, _U32, _P, _P, _P, _P] + _WS),
    "cph_row_sums": (_I, [_P, _P, _P, _U64, _U32, _I, _P, _I, _P, _P, _P]),
    "cph_incidence_csr": (_I, [_P, _P, _P, _U64, _U32, _P, _P, _P] + _WS),
    "cph_incidence_csr_fixed": (_I, [_P, _P, _P, _U64, _U32, _P, _P, _P, _P]),
    "cph_diag_hist": (_I, [_P, _P, _P, _U64, _U32, _I, _I, _P, _P]),
    "cph_incidence_build": (_I, [_P, _P, _P, _P, _U64, _U32, _I, _P, _U32, _U32, _I, _P, _P, _P, _P, _P,
                                 C.POINTER(_U64)] + _WS),
    "cph_csr_counts_f32": (_I, [_P, _U64, _P, _P]),
    "cph_row_sq_sums": (_I, [_P, _P, _U32, _P, _P]),
    "cph_diag_terms": (_I, [_P, _P, _P, _U32, _I, _I, _P, _P]),
    "cph_expand_accumulate": (_I, [_P, _P, _P, _U64, _U32, _I, _I, _I, _P, _P]),
    "cph_expand_accumulate_diag": (_I, [_P, _P, _P, _U64, _U32, _I, _I, _I, _P, _P, _P]),
    "cph_expand_sweep_prep": (_I, [_P, _P, _U64, _U32, _I, _P, _P, _P]),
    "cph_expand_sweep_band": (_I, [_P, _U64, _U32, _U32, _U32, _I, _I, _P, _P, _P]),
    "cph_expand_accumulate_binned": (_I, [_I, _P, _P, _U64, _U32, _I, _I, _I, _P, _P, _P, C.POINTER(_U64)] + _WS),
    "cph_expand_count": (_I, [_P, _P, _U32, _I, _F, C.POINTER(_U64)] + _WS),
    "cph_expand_emit": (_I, [_P, _P, _U32, _I, _F, _P, _P, _P, _P, _P, _P, _P]),
    "cph_csr_find_pairs": (_I, [_P, _P, _U32, _P, _P, _U64, _P, _P]),
    "cph_csr_reweight_pairs": (_I, [_P, _P, _U32, _U64, _P, _P, _U64, _F, _P, _P, _U64, _F, _P, _P, _P, _P]),
    "cph_csr_scale_nodes": (_I, [_P, _P, _U32, _P, _P, _P, _P]),
    "cph_csr_abs_degree": (_I, [_P, _P, _U32, _P, _P, _I, _P, _P, _P]),
    "cph_csr_row_ids": (_I, [_P, _U32, _P, _P]),
    "cph_kprune": (_I, [_P, _P, _P, _P, _U32, _U64, _P, _P, _P, _I, _P, _P, _U64, _P, _P, _P, _P]),
    "cph_community": (_I, [_P, _P, _P, _P, _U32, _U64, _D, _I, _I, _I, _I, _P, _P, C.POINTER(_D), C.POINTER(_U32)] + _WS),
    "cph_modularity": (_I, [_P, _P, _P, _P, _U32, _U64, _D, _P, C.POINTER(_D)] + _WS),
    "cph_group_sums": (_I, [_P, _P, _P, _P, _U32, _U64, _P, _U32, _P, _P] + _WS),
    "cph_irmm_edge_weights": (_I, [_P, _P, _U64, _P, _U32, _P, _P, _I, _P, _P, _P]),
    "cph_expand_full": (_I, [_P, _P, _P, _P, _P, _U64, _U32, _I, _P, _P]),
    "cph_dense_from_csr": (_I, [_P, _P, _P, _P, _U32, _P, _P]),
    "cph_irmm_blend": (_I, [_P, _I, _D, _P, _U32, _D, _P, _P]),
    "cph_dense_scale_pairs": (_I, [_P, _U32, _P, _P, _U64, _D, _P]),
    "cph_dense_to_csr_count": (_I, [_P, _U32, _I, C.POINTER(_U64)] + _WS),
    "cph_dense_to_csr_emit": (_I, [_P, _U32, _I, _P, _P, _P, _P, _P, _P]),
    "cph_fnv1a": (_U64, [C.c_char_p, _U64]),
    "cph_tsv_count_rows": (_I, [_P, _U64, _I, C.POINTER(_U64)]),
    "cph_tsv_parse": (_I, [_P, _U64, _I, _I, C.POINTER(_I), C.POINTER(_I), C.POINTER(_P), _U64, _P, _P, _U32, _P, _U32,
                           C.POINTER(_U64), C.POINTER(_U64)]),
    "cph_synth_porec_count": (_I, [_U64, _U64, _U64, _I, _P, C.POINTER(_U64)] + _WS),
    "cph_synth_porec_fill": (_I, [_U64, _U64, _U64, _I, _U32, _U32, _U32, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "cph_synth_pairs": (_I, [_U64, _U64, _U64, _D, _U32, _U32, _U32, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
}


def declared_symbols():
    """Every symbol include/cph.h declares (tests check the header and this table agree)."""
    return sorted(_SIGS)


def lib():
    """Load libcph.so once.  Fails loudly when it has not been built (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise CphError(f"{LIB_PATH} is missing: build it with `make -C cphasing_b200/csrc` "
                           "(or __graft_entry__.build()); there is no CPU fallback")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGS.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def register(sigs):
    """Later modules (kprune, community, multi-GPU) add their entry points here."""
    _SIGS.update(sigs)
    if _lib is not None:
        for name, (res, args) in sigs.items():
            fn = getattr(_lib, name)
            fn.restype = res
            fn.argtypes = args


def ptr(t):
    """device (or host) pointer of a tensor, NULL for None."""
    if t is None:
        return None
    assert t.is_contiguous(), "libcph needs contiguous buffers"
    return t.data_ptr()


def stream():
    return torch.cuda.current_stream().cuda_stream


def check(status):
    if status != CPH_OK:
        raise CphError(f"libcph status {status}: {lib().cph_last_error().decode()}")


def nvtx(name):
    """decorator: an NVTX range per stage of the path (tracing) -- shows up in nsys / ncu --nvtx timelines;
    a no-op cost of two calls when no tool listens."""
    import functools

    def deco(fn):
        @functools.wraps(fn)
        def wrapped(*a, **k):
            torch.cuda.nvtx.range_push(name)
            try:
                return fn(*a, **k)
            finally:
                torch.cuda.nvtx.range_pop()
        return wrapped
    return deco


def launches():
    return int(lib().cph_launch_count())


_ws_cache = {}


def workspace(nbytes, device):
    """A grow-only scratch buffer per device (the C-ABI never allocates)."""
    key = str(device)
    t = _ws_cache.get(key)
    if t is None or t.numel() < nbytes:
        _ws_cache[key] = None
        t = torch.empty(max(int(nbytes), 1 << 20), dtype=torch.uint8, device=device)
        _ws_cache[key] = t
    return t


def release_workspace():
    _ws_cache.clear()


def call_ws(name, device, *args, ws_tail=()):
    """Two-phase call of an entry point that ends in (ws, ws_bytes, h_ws_need, stream)."""
    fn = getattr(lib(), name)
    need = _SZ(0)
    check(fn(*args, None, 0, C.byref(need), None))
    ws = workspace(need.value, device)
    check(fn(*args, ws.data_ptr(), ws.numel(), C.byref(need), stream()))
    return ws
