"""ctypes binding of libdavis_b200.so (include/davis_b200.h).

There is no CPU fallback: if the library is missing or cannot be loaded the
import of the product modules fails with the reason.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import build as _build

OK, E_INVAL, E_SIZE, E_CUDA, E_NOMEM, E_OVERFLOW, E_STATE = 0, -1, -2, -3, -4, -5, -6


class DavisCfg(C.Structure):
    _fields_ = [("bit_rate", C.c_int32), ("symbol_length", C.c_int32),
                ("preamble_symbols", C.c_int32), ("packet_symbols", C.c_int32),
                ("preamble", C.c_uint32), ("block_size", C.c_int32),
                ("max_blocks", C.c_int32), ("max_hits", C.c_int32),
                ("host_slots", C.c_int32), ("host_group", C.c_int32),
                ("flavor", C.c_int32)]


class DavisStats(C.Structure):
    _fields_ = [("samples", C.c_uint64), ("launches", C.c_uint64), ("candidates", C.c_uint64),
                ("confirmed", C.c_uint64), ("records", C.c_uint64), ("demod_ms", C.c_float),
                ("search_ms", C.c_float), ("finish_ms", C.c_float), ("total_ms", C.c_float),
                ("h2d_bytes", C.c_uint64), ("d2h_bytes", C.c_uint64),
                ("host_submit_ms", C.c_float), ("host_wait_ms", C.c_float),
                ("host_copy_ms", C.c_float), ("host_sort_ms", C.c_float), ("reruns", C.c_uint64),
                ("graph_replays", C.c_uint64)]


HIT_DTYPE = np.dtype([("position", "<i8"), ("stream", "<u4"), ("call", "<u4"), ("index", "<i4"),
                      ("data", "u1", (10,)), ("crc_ok", "u1"), ("flags", "u1"),
                      ("rssi", "<f8"), ("snr", "<f8"), ("disc_mean", "<f8")])
assert HIT_DTYPE.itemsize == 56

EXPORTS = ["davis_abi_version", "davis_create", "davis_destroy", "davis_reset", "davis_demod_u8",
           "davis_demod_c64", "davis_demod_c128", "davis_fetch_hits", "davis_fetch_hits_view", "davis_get_stats", "davis_debug_bits",
           "davis_debug_disc", "davis_debug_disc32", "davis_host_alloc", "davis_host_free",
           "davis_crc16_batch", "davis_last_error", "davis_strerror",
           "davis_synth_fill", "davis_synth_truth", "davis_disc32_static"]

_lib = None


class DavisError(RuntimeError):
    def __init__(self, code: int, text: str):
        super().__init__(f"libdavis_b200: {text} (code {code})")
        self.code = code


def library_path() -> str:
    return _build.LIB_PATH


def load() -> C.CDLL:
    """Load (building first if the sources are newer and nvcc is present)."""
    global _lib
    if _lib is not None:
        return _lib
    path = _build.LIB_PATH
    if _build.needs_build():
        try:
            _build.build_library()
        except Exception as e:  # no nvcc on this box and no prebuilt library
            if not os.path.exists(path):
                raise ImportError(f"libdavis_b200.so is not built and cannot be built here: {e}")
    L = C.CDLL(path)
    vp, sz, i32 = C.c_void_p, C.c_size_t, C.c_int
    L.davis_abi_version.restype = i32
    L.davis_create.argtypes = [C.POINTER(DavisCfg), i32, i32, C.POINTER(vp)]
    L.davis_destroy.argtypes = [vp]
    L.davis_reset.argtypes = [vp, i32]
    L.davis_demod_u8.argtypes = [vp, vp, sz, sz, i32, vp]
    L.davis_demod_c64.argtypes = [vp, vp, sz, sz, i32, vp]
    L.davis_demod_c128.argtypes = [vp, vp, sz, sz, i32, vp]
    L.davis_fetch_hits.argtypes = [vp, vp, sz, C.POINTER(sz)]
    L.davis_fetch_hits_view.argtypes = [vp, C.POINTER(vp), C.POINTER(sz)]
    L.davis_get_stats.argtypes = [vp, C.POINTER(DavisStats)]
    L.davis_debug_bits.argtypes = [vp, i32, vp, sz, C.POINTER(C.c_int64), C.POINTER(sz)]
    L.davis_debug_disc.argtypes = [vp, i32, vp, sz]
    L.davis_debug_disc32.argtypes = [vp, i32, vp, sz]
    L.davis_disc32_static.argtypes = [vp, sz, vp, i32, vp, C.c_char_p, sz]
    L.davis_host_alloc.argtypes = [C.POINTER(vp), sz]
    L.davis_host_free.argtypes = [vp]
    L.davis_crc16_batch.argtypes = [vp, sz, sz, C.c_uint16, C.c_uint16, i32, i32, vp]
    L.davis_last_error.argtypes = [vp]
    L.davis_last_error.restype = C.c_char_p
    L.davis_strerror.argtypes = [i32]
    L.davis_strerror.restype = C.c_char_p
    L.davis_synth_fill.argtypes = [vp, sz, sz, sz, C.c_uint64, i32, C.c_float, C.c_float,
                                   C.c_float, i32, vp]
    L.davis_synth_truth.argtypes = [sz, sz, C.c_uint64, i32, C.c_float, vp, vp]
    for name in EXPORTS:
        fn = getattr(L, name)
        if fn.restype is C.c_int or fn.restype is None:
            fn.restype = i32
    if L.davis_abi_version() != 3:
        raise ImportError("libdavis_b200.so ABI version mismatch")
    _lib = L
    return L


def check(code: int, handle=None) -> None:
    if code == OK:
        return
    L = load()
    text = L.davis_strerror(code).decode()
    if handle:
        detail = L.davis_last_error(handle).decode()
        if detail:
            text = f"{text}: {detail}"
    if code == E_SIZE:
        raise ValueError("Incompatible array sizes")        # the reference's own error
    raise DavisError(code, text)
