"""Seeded synthetic Davis ISS captures (uint8 interleaved I/Q, RTL-SDR format).

The recipe is SURVEY.md section 8(d): each packet is 6 payload bytes + a
big-endian CRC-16 (poly 0x1021, init 0), sent LSB-first per byte after the
``AA AA AA AA CB 89`` training/sync sequence, 2-FSK with Gaussian shaping
(BT = 0.5), 14 samples per symbol at 268.8 kS/s, with the carrier sitting at
-Fs/4 in the raw stream (the reference demodulator rotates by +Fs/4 before its
low-pass, src/rtldavis/dsp.py:42-49, so that is the only place it can decode).

This is host-side input generation for tests and for the numpy-sized bench
configurations; the full-size bench inputs are produced on the device by the
generator kernel in ``csrc/davis_synth.cu`` with the same signal model.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Optional, Sequence, Tuple

import numpy as np

FS = 268_800            # 19200 * 14, src/rtldavis/dsp.py:118
SPS = 14
SYNC = bytes([0xAA, 0xAA, 0xAA, 0xAA, 0xCB, 0x89])
BURST_BITS = (len(SYNC) + 8) * 8          # 112 bits on air
DEVIATION_HZ = 4800.0
AMPLITUDE = 0.6


def _rev8(b: int) -> int:
    return int(f"{b:08b}"[::-1], 2)


def crc16_xmodem(data: Sequence[int]) -> int:
    """Bitwise CRC-16, poly 0x1021, init 0 (the generator's own copy so the
    product package does not import the oracle)."""
    reg = 0
    for byte in data:
        reg ^= (int(byte) & 0xFF) << 8
        for _ in range(8):
            reg = ((reg << 1) ^ 0x1021) & 0xFFFF if reg & 0x8000 else (reg << 1) & 0xFFFF
    return reg


def make_message(station: int, sensor: int, body: Sequence[int]) -> bytes:
    """8-byte Davis message: header nibble pair, 5 body bytes, CRC hi, CRC lo."""
    head = ((sensor & 0xF) << 4) | (station & 0x7)
    six = bytes([head, *[int(b) & 0xFF for b in body][:5]])
    c = crc16_xmodem(six)
    return six + bytes([c >> 8, c & 0xFF])


def air_bits(message: bytes) -> np.ndarray:
    """Bits in transmit order: sync sequence MSB-first, message bytes bit-reversed
    (the receiver undoes that with swap_bit_order, src/rtldavis/protocol.py:290)."""
    on_air = SYNC + bytes(_rev8(b) for b in message)
    return np.unpackbits(np.frombuffer(on_air, dtype=np.uint8))


def _gauss_taps(bt: float = 0.5, span: int = 4, sps: int = SPS) -> np.ndarray:
    t = (np.arange(span * sps + 1) - span * sps / 2) / sps
    sigma = np.sqrt(np.log(2.0)) / (2 * np.pi * bt)
    h = np.exp(-0.5 * (t / sigma) ** 2)
    return h / h.sum()


_GAUSS = _gauss_taps()


@dataclass
class Embedded:
    start: int            # first sample of the burst (start of the AA training)
    message: bytes        # the 8 message bytes (host bit order)


@dataclass
class Capture:
    iq: np.ndarray                        # uint8, 2 * n_samples
    packets: List[Embedded] = field(default_factory=list)


def burst_waveform(message: bytes, df_hz: float = 0.0, clock_ppm: float = 0.0,
                   phase0: float = 0.0) -> np.ndarray:
    """Complex baseband burst at unit amplitude, carrier at -Fs/4 + df_hz."""
    bits = air_bits(message)
    sps = SPS * (1.0 + clock_ppm * 1e-6)
    n = int(np.ceil(len(bits) * sps))
    which = np.minimum((np.arange(n) / sps).astype(np.int64), len(bits) - 1)
    nrz = 2.0 * bits[which] - 1.0
    pad = len(_GAUSS) // 2
    shaped = np.convolve(np.pad(nrz, pad, mode="edge"), _GAUSS, mode="valid")
    freq = DEVIATION_HZ * shaped + df_hz - FS / 4.0
    phase = phase0 + 2 * np.pi * np.cumsum(freq) / FS
    return np.exp(1j * phase)


def to_u8(x: np.ndarray) -> np.ndarray:
    """Inverse of the reference's LUT (src/rtldavis/dsp.py:26) with rounding and
    clipping, interleaved I, Q."""
    out = np.empty(2 * x.size, dtype=np.uint8)
    out[0::2] = np.clip(np.rint(x.real * 127.6 + 127.4), 0, 255).astype(np.uint8)
    out[1::2] = np.clip(np.rint(x.imag * 127.6 + 127.4), 0, 255).astype(np.uint8)
    return out


def make_capture(seed: int, n_samples: int, n_packets: int, snr_db: float = 30.0,
                 station: Optional[int] = None, df_hz: float = 0.0,
                 clock_ppm: float = 0.0, starts: Optional[Sequence[int]] = None,
                 noise_only: bool = False) -> Capture:
    """One stream.  ``seed`` fully determines payloads, jitter and noise.

    SNR is burst power over noise power in the full 268.8 kHz band.  Packets are
    spread evenly with 0..13 samples of jitter unless ``starts`` is given."""
    rng = np.random.default_rng(seed)
    burst_len = int(np.ceil(BURST_BITS * SPS * (1 + abs(clock_ppm) * 1e-6))) + 2
    x = np.zeros(n_samples, dtype=np.complex128)
    packets: List[Embedded] = []
    if not noise_only and n_packets > 0:
        if starts is None:
            gap = n_samples // n_packets
            if gap < burst_len + 64:
                raise ValueError("capture too short for that many packets")
            starts = [k * gap + gap // 4 + int(rng.integers(0, SPS)) for k in range(n_packets)]
        for s in starts:
            st = int(rng.integers(0, 8)) if station is None else station
            sensor = int(rng.integers(2, 11))
            msg = make_message(st, sensor, rng.integers(0, 256, size=5))
            w = burst_waveform(msg, df_hz, clock_ppm, phase0=float(rng.uniform(0, 2 * np.pi)))
            e = min(n_samples, s + w.size)
            if s < 0 or e <= s:
                continue
            x[s:e] += AMPLITUDE * w[: e - s]
            packets.append(Embedded(int(s), msg))
    sigma = AMPLITUDE / np.sqrt(2.0) * 10 ** (-snr_db / 20.0)
    x += sigma * (rng.standard_normal(n_samples) + 1j * rng.standard_normal(n_samples))
    return Capture(to_u8(x), packets)


def make_sync_flood(n_samples: int, seed: int, snr_db: float = 25.0, df_hz: float = 800.0) -> np.ndarray:
    """A transmitter that sends the sync word ``CB 89 CB 89 ...`` for the whole capture: every 16th
    symbol starts a preamble match, at several sample phases.  The reference answers such input
    (src/rtldavis/dsp.py:171-205); lists sized for the matches of noise overflow on it."""
    rng = np.random.default_rng(seed)
    n_bits = n_samples // SPS + 2
    bits = np.resize(np.unpackbits(np.array([0xCB, 0x89], dtype=np.uint8)), n_bits)
    nrz = np.repeat(2.0 * bits - 1.0, SPS)[:n_samples]
    pad = len(_GAUSS) // 2
    shaped = np.convolve(np.pad(nrz, pad, mode="edge"), _GAUSS, mode="valid")
    phase = 2 * np.pi * np.cumsum(DEVIATION_HZ * shaped + df_hz - FS / 4.0) / FS
    sigma = AMPLITUDE / np.sqrt(2.0) * 10 ** (-snr_db / 20.0)
    x = AMPLITUDE * np.exp(1j * phase) + sigma * (rng.standard_normal(n_samples) + 1j * rng.standard_normal(n_samples))
    return to_u8(x)


def make_batch(n_streams: int, n_samples: int, n_packets: int, snr_db: float = 30.0,
               seed0: int = 0, df_range_hz: float = 0.0, ppm_range: float = 0.0
               ) -> Tuple[np.ndarray, List[List[Embedded]]]:
    """``uint8[n_streams, 2*n_samples]``; stream s uses seed ``seed0 + s``,
    station ``s mod 8``, and offsets drawn uniformly from the given ranges."""
    out = np.empty((n_streams, 2 * n_samples), dtype=np.uint8)
    truth: List[List[Embedded]] = []
    for s in range(n_streams):
        r = np.random.default_rng(10_000_019 * (seed0 + s) + 7)
        df = float(r.uniform(-df_range_hz, df_range_hz)) if df_range_hz else 0.0
        ppm = float(r.uniform(-ppm_range, ppm_range)) if ppm_range else 0.0
        cap = make_capture(seed0 + s, n_samples, n_packets, snr_db, station=s % 8,
                           df_hz=df, clock_ppm=ppm)
        out[s] = cap.iq
        truth.append(cap.packets)
    return out, truth


# The US-band hop sequence of the Davis protocol as rtldavis tabulates it (the 51 values of
# Parser.hop_pattern, src/rtldavis/protocol.py:173-225): hop k of a transmitter is on channel
# HOP_PATTERN_US[(hop0 + k) % 51] of Parser.channels (protocol.py:119-171).
HOP_PATTERN_US = (0, 19, 41, 25, 8, 47, 32, 13, 36, 22, 3, 29, 44, 16, 5, 27, 38, 10, 49, 21, 2, 30, 42,
                  14, 48, 7, 24, 34, 45, 1, 17, 39, 26, 9, 31, 50, 37, 12, 20, 33, 4, 43, 28, 15, 35, 6,
                  40, 11, 23, 46, 18)
assert sorted(HOP_PATTERN_US) == list(range(51))


def make_hop_set(n_channels: int, n_samples: int, n_transmitters: int, hop_samples: int,
                 snr_db: float = 30.0, seed: int = 0, df_range_hz: float = 3000.0
                 ) -> Tuple[np.ndarray, List[List[Embedded]]]:
    """BASELINE config 3: ``n_channels`` parallel narrowband streams (one per channel
    of the hop set, src/rtldavis/protocol.py:119-171).  Transmitter t sends one packet
    every ``hop_samples`` samples and moves to the next channel of the hop pattern after
    each one, so every packet appears on exactly one stream.  With the full set of 51
    channels the pattern is the reference's own table (``HOP_PATTERN_US``, indexing
    Parser.channels); a smaller set hops a seeded permutation.  Each channel has a fixed
    residual carrier offset (the tuner's error on that channel) inside ``df_range_hz``."""
    rng = np.random.default_rng(seed)
    df = rng.uniform(-df_range_hz, df_range_hz, size=n_channels)
    pattern = rng.permutation(n_channels)              # drawn either way: keeps the seed's other draws
    if n_channels == len(HOP_PATTERN_US):
        pattern = np.asarray(HOP_PATTERN_US)
    x = np.zeros((n_channels, n_samples), dtype=np.complex128)
    truth: List[List[Embedded]] = [[] for _ in range(n_channels)]
    burst_len = BURST_BITS * SPS + 2
    for t in range(n_transmitters):
        offset = int(rng.integers(0, hop_samples))
        hop0 = int(rng.integers(0, n_channels))
        k = 0
        while True:
            s = offset + k * hop_samples + int(rng.integers(0, SPS))
            if s + burst_len >= n_samples:
                break
            ch = int(pattern[(hop0 + k) % n_channels])
            msg = make_message(t % 8, int(rng.integers(2, 11)), rng.integers(0, 256, size=5))
            w = burst_waveform(msg, float(df[ch]), 0.0, phase0=float(rng.uniform(0, 2 * np.pi)))
            x[ch, s:s + w.size] += AMPLITUDE * w
            truth[ch].append(Embedded(s, msg))
            k += 1
    sigma = AMPLITUDE / np.sqrt(2.0) * 10 ** (-snr_db / 20.0)
    out = np.empty((n_channels, 2 * n_samples), dtype=np.uint8)
    for ch in range(n_channels):
        noise = sigma * (rng.standard_normal(n_samples) + 1j * rng.standard_normal(n_samples))
        out[ch] = to_u8(x[ch] + noise)
    return out, truth
