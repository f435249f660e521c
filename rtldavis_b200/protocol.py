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


def new_go_packet_config(symbol_length: int = 14) -> dsp.PacketConfig:
    """protocol.NewPacketConfig, protocol/protocol.go:30-38 over dsp.NewPacketConfig,
    dsp/dsp.go:187-218 (block size 512), with the Go flavor of the numbers."""
    return dsp.PacketConfig(bit_rate=19200, symbol_length=symbol_length, preamble_symbols=16,
                            packet_symbols=80, preamble="1100101110001001", block_size=512,
                            flavor="go")


def swap_bit_order(b: int) -> int:
    b = int(b) & 0xFF
    return int(f"{b:08b}"[::-1], 2)


def crc_gate(hits: np.ndarray) -> List[Tuple[bytes, bool]]:
    """For records returned by the library: (swapped 10 bytes, CRC verdict).

    The verdict is the one computed on the device; per-call dedupe on the swapped
    payload (protocol.py:293-295) is implied by the library's dedupe on the
    unswapped payload (a bijection)."""
    return [(bytes(swap_bit_order(x) for x in h["data"]), bool(h["crc_ok"])) for h in hits]


def parse_hits(hits: np.ndarray, cfg: dsp.PacketConfig, station_id=None) -> List[dict]:
    """The part of ``Parser.parse`` (src/rtldavis/protocol.py:282-332) that sits on the
    demodulation path, for a whole batch of records: bit-order swap, CRC gate, the
    frequency error ``-int(mean(discriminated[index:index+224]) * Fs / 2 pi)``
    (protocol.py:304-311; the mean comes from the device, ``davis_hit.disc_mean``),
    message id and the station filter (protocol.py:325-329).  Hop bookkeeping and sensor
    decoding (protocol.py:316-323,331) stay with the reference's Parser.  With
    ``cfg.flavor == "go"`` the device mean is the one over symbols 8..24 and the error is
    ``-int(9600 + mean * Fs / 2 pi)`` (protocol/protocol.go:156-168; Go's int() truncates
    toward zero like Python's).  Each returned
    dict carries what ``_parse_sensor_data`` needs (``data`` = the 8 message bytes,
    ``rssi``, ``snr``) plus ``stream`` and ``call``."""
    import math
    out: List[dict] = []
    for h in hits:
        if not h["crc_ok"]:
            continue
        data = bytes(swap_bit_order(x) for x in h["data"])
        msg = data[2:]
        msg_id = msg[0] & 0x7
        if station_id is not None and msg_id != station_id:
            continue
        hz = (float(h["disc_mean"]) * float(cfg.sample_rate)) / (2 * math.pi)
        if getattr(cfg, "flavor", "python") == "go":
            # 0/0 in Go's discriminator (zero history under the window) gives NaN, whose
            # int conversion is platform-defined in Go: reported as None here
            freq_err = None if math.isnan(hz) else -int(9600 + hz)
        else:
            freq_err = -int(hz)
        out.append({"stream": int(h["stream"]), "call": int(h["call"]), "index": int(h["index"]),
                    "data": msg, "id": msg_id, "freq_err": freq_err,
                    "rssi": float(h["rssi"]), "snr": float(h["snr"])})
    return out
