"""ctypes front end of oracle/davis_oracle.c (TEST INFRASTRUCTURE)."""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import List, Optional

import numpy as np

from .davis_oracle import OraclePacket

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libdavis_oracle.so")


class _Pkt(C.Structure):
    _fields_ = [("stream", C.c_uint32), ("call", C.c_uint32), ("index", C.c_int32),
                ("data", C.c_uint8 * 10), ("crc_ok", C.c_uint8), ("pad", C.c_uint8)]


PKT_DTYPE = np.dtype([("stream", "<u4"), ("call", "<u4"), ("index", "<i4"),
                      ("data", "u1", (10,)), ("crc_ok", "u1"), ("pad", "u1")])
assert PKT_DTYPE.itemsize == C.sizeof(_Pkt) == 24

_lib: Optional[C.CDLL] = None


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "davis_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B"])
    return _SO


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_SO)
        L.dor_crc16.restype = C.c_uint16
        L.dor_crc16.argtypes = [C.c_void_p, C.c_size_t, C.c_uint16]
        L.dor_swap_bits.restype = C.c_uint8
        L.dor_swap_bits.argtypes = [C.c_uint8]
        L.dor_crc_gate.restype = C.c_int
        L.dor_crc_gate.argtypes = [C.c_void_p]
        L.dor_stream_bits.restype = None
        L.dor_stream_bits.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p]
        L.dor_packets_from_bits.restype = C.c_long
        L.dor_packets_from_bits.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.c_uint32,
                                            C.c_uint32, C.c_void_p, C.c_size_t]
        L.dor_batch.restype = None
        L.dor_batch.argtypes = [C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_int,
                                C.c_uint32, C.c_void_p, C.c_size_t, C.c_void_p, C.c_int]
        _lib = L
    return _lib


def crc16(data: bytes, init: int = 0) -> int:
    buf = (C.c_uint8 * len(data)).from_buffer_copy(bytes(data)) if data else None
    return int(lib().dor_crc16(buf, len(data), init))


def stream_bits(u8: np.ndarray, want_disc: bool = False, want_filt: bool = False):
    u8 = np.ascontiguousarray(u8, dtype=np.uint8)
    n = u8.size // 2
    bits = np.empty(n, dtype=np.uint8)
    disc = np.empty(n, dtype=np.float64) if want_disc else None
    filt = np.empty(2 * n, dtype=np.float64) if want_filt else None
    lib().dor_stream_bits(u8.ctypes.data, n, bits.ctypes.data,
                          disc.ctypes.data if want_disc else None,
                          filt.ctypes.data if want_filt else None)
    res = [bits]
    if want_disc:
        res.append(disc)
    if want_filt:
        res.append(filt.view(np.complex128))
    return res[0] if len(res) == 1 else tuple(res)


def batch_records(iq: np.ndarray, block: int = 8192, preamble: int = 0xCB89,
                  per_stream_cap: int = 4096, n_threads: int = 0) -> np.ndarray:
    """Packets of every call of every stream of ``iq[n_streams, 2*n_samples]`` as
    a structured array (stream-major, then call, then the reference's order)."""
    iq = np.ascontiguousarray(iq, dtype=np.uint8)
    if iq.ndim == 1:
        iq = iq[None, :]
    n_streams, nbytes = iq.shape
    out = np.zeros((n_streams, per_stream_cap), dtype=PKT_DTYPE)
    counts = np.zeros(n_streams, dtype=np.int64)
    if n_threads <= 0:
        n_threads = os.cpu_count() or 1
    lib().dor_batch(iq.ctypes.data, n_streams, iq.strides[0], nbytes // 2, block, preamble,
                    out.ctypes.data, per_stream_cap, counts.ctypes.data, n_threads)
    if (counts < 0).any():
        raise RuntimeError("oracle packet buffer too small")
    return np.concatenate([out[s, :counts[s]] for s in range(n_streams)]) if n_streams else out[0, :0]


def batch_packets(iq: np.ndarray, **kw) -> List[List[OraclePacket]]:
    rec = batch_records(iq, **kw)
    n_streams = 1 if np.ndim(iq) == 1 else np.shape(iq)[0]
    per: List[List[OraclePacket]] = [[] for _ in range(n_streams)]
    for r in rec:
        per[int(r["stream"])].append(OraclePacket(int(r["call"]), int(r["index"]),
                                                  r["data"].tobytes(), bool(r["crc_ok"])))
    return per
