"""``rtldavis.crc`` surface (src/rtldavis/crc.py:4-42) on the GPU library.

``CRC(name, init, poly, residue).checksum(data)`` keeps the reference signature;
the arithmetic runs in the library's CRC kernel (one launch per call -- this
entry exists for API parity and parity tests; on the hot path the CRC is fused
into the slicing kernel and arrives as ``davis_hit.crc_ok``).
``checksum_many`` is the batched form.
"""
from __future__ import annotations

import numpy as np

from . import _lib


class CRC:
    def __init__(self, name: str, init: int, poly: int, residue: int, device: int = 0) -> None:
        self.name = name
        self.init = np.uint16(init)
        self.poly = np.uint16(poly)
        self.residue = np.uint16(residue)
        self.device = device
        self._L = _lib.load()

    def __str__(self) -> str:
        return (f"{{Name:{self.name} Init:0x{int(self.init):04X} Poly:0x{int(self.poly):04X} "
                f"Residue:0x{int(self.residue):04X}}}")

    def checksum_many(self, rows: np.ndarray, swap_bits: bool = False) -> np.ndarray:
        rows = np.ascontiguousarray(rows, dtype=np.uint8)
        if rows.ndim != 2:
            raise ValueError("rows must be uint8[n, len]")
        out = np.zeros(rows.shape[0], dtype=np.uint16)
        if rows.shape[0] == 0:
            return out
        if rows.shape[1] == 0:
            out[:] = self.init
            return out
        _lib.check(self._L.davis_crc16_batch(rows.ctypes.data, rows.shape[0], rows.shape[1],
                                             int(self.init), int(self.poly), int(swap_bits),
                                             self.device, out.ctypes.data))
        return out

    def checksum(self, data: bytes) -> np.uint16:
        row = np.frombuffer(bytes(data), dtype=np.uint8).reshape(1, -1)
        return np.uint16(self.checksum_many(row)[0])
