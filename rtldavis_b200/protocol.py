"""The demodulation-path part of ``rtldavis.protocol``.

``new_packet_config`` (src/rtldavis/protocol.py:68-76), ``swap_bit_order``
(protocol.py:79-83) and the CRC gate of ``Parser.parse`` (protocol.py:290-300).
Hop tables, frequency correction bookkeeping and sensor decoding stay in the
reference (out of scope).
"""
from __future__ import annotations

from typing import List, Tuple

import numpy as np

from . import dsp


def new_packet_config(symbol_length: int) -> dsp.PacketConfig:
    return dsp.PacketConfig(bit_rate=19200, symbol_length=symbol_length, preamble_symbols=16,
                            packet_symbols=80, preamble="1100101110001001", block_size=8192)


def swap_bit_order(b: int) -> int:
    b = int(b) & 0xFF
    return int(f"{b:08b}"[::-1], 2)


def crc_gate(hits: np.ndarray) -> List[Tuple[bytes, bool]]:
    """For records returned by the library: (swapped 10 bytes, CRC verdict).

    The verdict is the one computed on the device; per-call dedupe on the swapped
    payload (protocol.py:293-295) is implied by the library's dedupe on the
    unswapped payload (a bijection)."""
    return [(bytes(swap_bit_order(x) for x in h["data"]), bool(h["crc_ok"])) for h in hits]
