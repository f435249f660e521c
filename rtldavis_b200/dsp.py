"""Drop-in for ``rtldavis.dsp`` backed by the sm_100a kernels in libdavis_b200.so.

Same names and behaviour as the reference module for the demodulation path:
``PacketConfig`` (src/rtldavis/dsp.py:101-125), ``Packet`` (dsp.py:12-17) and
``Demodulator`` with ``demodulate(block) -> List[Packet]`` / ``reset()`` /
``.cfg`` / ``.discriminated`` (dsp.py:127-253).  ``BatchDemodulator`` is the
batched surface the GPU is built for: many independent streams, many blocks per
call, identical results to running one reference ``Demodulator`` per stream.

Everything numeric happens on the GPU through the C ABI (include/davis_b200.h);
this module only marshals buffers.  There is no CPU implementation to fall back
to: without the CUDA library or a CUDA device construction fails.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import List, Optional, Sequence

import numpy as np

from . import _lib
from ._lib import HIT_DTYPE, DavisCfg, DavisStats, check


@dataclass
class Packet:
    """dsp.Packet, src/rtldavis/dsp.py:12-17."""
    index: int
    data: np.ndarray
    rssi: float
    snr: float


class PacketConfig:
    """dsp.PacketConfig, src/rtldavis/dsp.py:101-125 (same derived attributes).

    ``flavor`` is not in the reference: ``"python"`` (default) follows src/rtldavis,
    ``"go"`` follows the Go mirror where its numbers differ -- no epsilon in the
    discriminator (dsp/dsp.go:93) and the parser's mean taken over symbols 8..24
    (protocol/protocol.go:156-164); pair it with ``block_size=512`` (dsp/dsp.go:212)."""

    FLAVORS = {"python": 0, "go": 1}

    def __init__(self, bit_rate: int, symbol_length: int, preamble_symbols: int,
                 packet_symbols: int, preamble: str, block_size: int = 512,
                 flavor: str = "python") -> None:
        if flavor not in self.FLAVORS:
            raise ValueError("flavor must be 'python' or 'go'")
        self.flavor = flavor
        self.bit_rate = bit_rate
        self.symbol_length = symbol_length
        self.preamble_symbols = preamble_symbols
        self.packet_symbols = packet_symbols
        self.preamble = preamble
        self.preamble_bytes = np.array([int(ch) for ch in preamble], dtype=np.uint8)
        self.preamble_str = self.preamble_bytes.tobytes()
        self.sample_rate = bit_rate * symbol_length
        self.block_size = block_size
        self.block_size2 = 2 * block_size
        self.preamble_length = preamble_symbols * symbol_length
        self.packet_length = packet_symbols * symbol_length
        self.buffer_length = (self.packet_length // block_size + 2) * block_size

    def _to_c(self, max_blocks: int, max_hits: int = 0, host_slots: int = 0,
              host_group: int = 0) -> DavisCfg:
        if len(self.preamble) != self.preamble_symbols or set(self.preamble) - {"0", "1"}:
            raise ValueError("preamble must be a string of preamble_symbols 0/1 characters")
        return DavisCfg(self.bit_rate, self.symbol_length, self.preamble_symbols,
                        self.packet_symbols, int(self.preamble, 2), self.block_size,
                        max_blocks, max_hits, host_slots, host_group,
                        self.FLAVORS[self.flavor])


def _device_pointer(x):
    """(pointer, nbytes-per-row info) for torch CUDA tensors / __cuda_array_interface__."""
    if hasattr(x, "data_ptr") and hasattr(x, "is_cuda"):
        if not x.is_cuda:
            return None
        return int(x.data_ptr())
    cai = getattr(x, "__cuda_array_interface__", None)
    if cai is not None:
        return int(cai["data"][0])
    return None


_UNSUPPORTED = ("unsupported PacketConfig for the B200 demodulator "
                "(needs symbol_length 14, 16-symbol preamble, 80-symbol packet, "
                "block_size a multiple of 8 and >= 256)")


def _check_supported(cfg: "PacketConfig") -> None:
    """davis_create's configuration checks (davis_capi.cu), made without touching CUDA."""
    cfg._to_c(1)                                              # preamble string
    if (cfg.symbol_length != 14 or cfg.preamble_symbols != 16 or cfg.packet_symbols != 80
            or cfg.block_size < 256 or cfg.block_size % 8):
        raise ValueError(_UNSUPPORTED)


class BatchDemodulator:
    """``n_streams`` independent reference Demodulators advanced together.

    ``submit(iq)`` takes ``uint8[n_streams, n_blocks * 2 * block_size]`` (numpy /
    pinned host memory, or a CUDA tensor used in place) and enqueues the work;
    ``fetch()`` returns a structured array (``_lib.HIT_DTYPE``) ordered by stream,
    call and the reference's own per-call order.  ``demodulate(iq)`` does both and
    returns ``result[stream][call] -> List[Packet]``.
    """

    def __init__(self, cfg: PacketConfig, n_streams: int = 1, device: int = 0,
                 max_blocks: int = 1, max_hits: int = 0, host_slots: int = 0,
                 host_group: int = 0) -> None:
        self._L = _lib.load()
        self.cfg = cfg
        self.n_streams = int(n_streams)
        self.device = int(device)
        self.max_blocks = int(max_blocks)
        self._h = C.c_void_p()
        ccfg = cfg._to_c(self.max_blocks, max_hits, host_slots, host_group)
        rc = self._L.davis_create(C.byref(ccfg), self.n_streams, self.device, C.byref(self._h))
        if rc != 0:
            self._h = C.c_void_p()
            if rc == _lib.E_INVAL:
                raise ValueError(_UNSUPPORTED)
            check(rc)
        self._keepalive = None
        self._last_blocks = 0
        # blocks each stream has consumed since its last reset: the library numbers the calls
        # of a stream from there (davis_hit.call), and reset(stream) restarts one stream only
        self.calls_done = np.zeros(self.n_streams, dtype=np.int64)

    # -- lifetime ---------------------------------------------------------
    def close(self) -> None:
        if getattr(self, "_h", None) and self._h.value:
            self._L.davis_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # -- core -------------------------------------------------------------
    def reset(self, stream: int = -1) -> None:
        """Demodulator.reset (dsp.py:248-253) for one stream or all (-1)."""
        check(self._L.davis_reset(self._h, int(stream)), self._h)
        if stream < 0:
            self.calls_done[:] = 0
        else:
            self.calls_done[int(stream)] = 0

    def submit(self, iq, cuda_stream: Optional[int] = None) -> None:
        B = self.cfg.block_size
        ptr = _device_pointer(iq)
        if ptr is not None:
            shape = tuple(iq.shape)
            if getattr(iq, "dtype", None) is not None and "uint8" not in str(iq.dtype):
                raise TypeError("device input must be uint8")
            if len(shape) == 1:
                shape = (1, shape[0])
            if len(shape) != 2 or shape[0] != self.n_streams:
                raise ValueError("Incompatible array sizes")
            stride = int(iq.stride(0)) if hasattr(iq, "stride") and len(iq.shape) == 2 else shape[1]
            if hasattr(iq, "stride") and int(iq.stride(-1)) != 1:
                raise ValueError("device input rows must be contiguous")
            nbytes = shape[1]
        else:
            iq = np.asarray(iq)
            if np.iscomplexobj(iq):                        # dsp.py:144-151: dest[:] = input_data
                return self._submit_complex(iq, cuda_stream)
            if iq.dtype != np.uint8:
                raise TypeError("input must be uint8 interleaved I/Q or complex")
            if iq.ndim == 1:
                iq = iq[None, :]
            if iq.ndim != 2 or iq.shape[0] != self.n_streams:
                raise ValueError("Incompatible array sizes")
            if iq.strides[1] != 1:
                iq = np.ascontiguousarray(iq)
            ptr, stride, nbytes = iq.ctypes.data, iq.strides[0], iq.shape[1]
        if nbytes == 0 or nbytes % (2 * B):
            raise ValueError("Incompatible array sizes")
        n_blocks = nbytes // (2 * B)
        if n_blocks > self.max_blocks:
            raise ValueError(f"{n_blocks} blocks per call exceed max_blocks={self.max_blocks}")
        self._keepalive = iq
        rc = self._L.davis_demod_u8(self._h, ptr, nbytes, stride, n_blocks,
                                    C.c_void_p(cuda_stream) if cuda_stream else None)
        if rc != 0:
            self._keepalive = None
        check(rc, self._h)
        self._last_blocks = n_blocks
        self.calls_done += n_blocks

    def _submit_complex(self, iq: np.ndarray, cuda_stream: Optional[int]) -> None:
        """Complex host input: complex64 stays complex64, everything else becomes
        complex128 (the type the reference's raw_samples buffer has)."""
        B = self.cfg.block_size
        if iq.dtype != np.complex64:
            iq = iq.astype(np.complex128, copy=False)
        if iq.ndim == 1:
            iq = iq[None, :]
        if iq.ndim != 2 or iq.shape[0] != self.n_streams:
            raise ValueError("Incompatible array sizes")
        if iq.strides[1] != iq.itemsize:
            iq = np.ascontiguousarray(iq)
        n = iq.shape[1]
        if n == 0 or n % B:
            raise ValueError("Incompatible array sizes")
        n_blocks = n // B
        if n_blocks > self.max_blocks:
            raise ValueError(f"{n_blocks} blocks per call exceed max_blocks={self.max_blocks}")
        fn = self._L.davis_demod_c64 if iq.dtype == np.complex64 else self._L.davis_demod_c128
        self._keepalive = iq
        rc = fn(self._h, iq.ctypes.data, n, iq.strides[0], n_blocks,
                C.c_void_p(cuda_stream) if cuda_stream else None)
        if rc != 0:
            self._keepalive = None
        check(rc, self._h)
        self._last_blocks = n_blocks
        self.calls_done += n_blocks

    def fetch(self, copy: bool = True) -> np.ndarray:
        """Packets of the submitted call.  With ``copy=False`` the array is a view of
        the library's pinned result buffer, valid until the next submit()/reset()."""
        n = C.c_size_t(0)
        ptr = C.c_void_p()
        rc = self._L.davis_fetch_hits_view(self._h, C.byref(ptr), C.byref(n))
        self._keepalive = None
        self._last_blocks = 0
        check(rc, self._h)
        if n.value == 0:
            return np.empty(0, dtype=HIT_DTYPE)
        buf = (C.c_char * (n.value * HIT_DTYPE.itemsize)).from_address(ptr.value)
        view = np.frombuffer(buf, dtype=HIT_DTYPE, count=n.value)
        return view.copy() if copy else view

    def stats(self) -> DavisStats:
        st = DavisStats()
        check(self._L.davis_get_stats(self._h, C.byref(st)), self._h)
        return st

    def demodulate(self, iq) -> List[List[List[Packet]]]:
        first_call = self.calls_done.copy()
        self.submit(iq)
        n_blocks = self._last_blocks
        hits = self.fetch()
        out: List[List[List[Packet]]] = [[[] for _ in range(n_blocks)] for _ in range(self.n_streams)]
        for h in hits:
            s = int(h["stream"])
            out[s][int(h["call"]) - int(first_call[s])].append(
                Packet(int(h["index"]), h["data"].copy(), float(h["rssi"]), float(h["snr"])))
        return out

    # -- verification taps --------------------------------------------------
    def debug_bits(self, stream: int = 0):
        """(first_global_sample, uint8 0/1 bits) decided by the fp32 kernel in the last call."""
        cap = (self.cfg.buffer_length + self.max_blocks * self.cfg.block_size + 64)
        out = np.empty(cap, dtype=np.uint8)
        first, n = C.c_int64(0), C.c_size_t(0)
        check(self._L.davis_debug_bits(self._h, stream, out.ctypes.data, cap, C.byref(first),
                                       C.byref(n)), self._h)
        return int(first.value), out[: n.value]

    def discriminated(self, stream: int = 0) -> np.ndarray:
        out = np.empty(2 * self.cfg.block_size, dtype=np.float64)
        check(self._L.davis_debug_disc(self._h, stream, out.ctypes.data, out.size), self._h)
        return out

    def discriminated_fp32(self, stream: int = 0) -> np.ndarray:
        out = np.empty(2 * self.cfg.block_size, dtype=np.float32)
        check(self._L.davis_debug_disc32(self._h, stream, out.ctypes.data, out.size), self._h)
        return out


class Demodulator:
    """dsp.Demodulator, src/rtldavis/dsp.py:127-253: one stream, one block per call."""

    def __init__(self, cfg: PacketConfig, device: int = 0) -> None:
        # Nothing here touches CUDA.  rtldavis builds a Parser (and with it a Demodulator) in the
        # parent process just to read ``p.cfg`` and the hop tables (runners/rtlsdr.py:30), then
        # forks the DSP worker (rtlsdr.py:61-65, start method "fork": __main__.py:277), which
        # builds its own (worker.py:28).  A CUDA context created before the fork would make every
        # CUDA call in the child fail, so the device handle is created by the first call that
        # needs it -- in whichever process makes that call.  The library itself is loaded now
        # (no CPU fallback: a missing libdavis_b200.so fails here, not at the first block), and an
        # unsupported configuration is refused now as well.
        self.cfg = cfg
        self.device = int(device)
        _lib.load()
        _check_supported(cfg)
        self._lazy: Optional[BatchDemodulator] = None
        self._last_hits = np.empty(0, dtype=HIT_DTYPE)

    @property
    def _batch(self) -> BatchDemodulator:
        if self._lazy is None:
            self._lazy = BatchDemodulator(self.cfg, n_streams=1, device=self.device, max_blocks=1)
        return self._lazy

    def demodulate(self, input_data: np.ndarray) -> List[Packet]:
        arr = np.asarray(input_data)
        if np.iscomplexobj(arr):
            if arr.size != self.cfg.block_size:                  # dsp.py:145-149
                raise ValueError("Incompatible array sizes")
            self._batch.submit(arr.reshape(1, -1))
        else:
            if arr.size != 2 * self.cfg.block_size:              # dsp.py:32-36
                raise ValueError("Incompatible array sizes")
            self._batch.submit(arr.reshape(1, -1).astype(np.uint8, copy=False))
        self._last_hits = self._batch.fetch()
        return [Packet(int(h["index"]), h["data"].copy(), float(h["rssi"]), float(h["snr"]))
                for h in self._last_hits]

    @property
    def discriminated(self) -> np.ndarray:
        """float64[2*block_size], the buffer Parser.parse reads (protocol.py:307)."""
        return self._batch.discriminated(0)

    @property
    def last_hits(self) -> np.ndarray:
        """Structured records of the last call (adds crc_ok, disc_mean, position)."""
        return self._last_hits

    def reset(self) -> None:
        if self._lazy is not None:                           # a stream that never ran has no history
            self._lazy.reset(-1)

    def close(self) -> None:
        if self._lazy is not None:
            self._lazy.close()
            self._lazy = None
