"""numpy restatement of the reference demodulation path (TEST INFRASTRUCTURE).

Parity status: PINNED.  ``tests/golden/*.npz`` were produced by importing the
unmodified reference (``/root/reference/src/rtldavis``) in the build container
with ``oracle/gen_golden.py``; ``tests/test_oracle_golden.py`` replays them
against every function in this file on the CPU.

Two formulations of the same arithmetic live here:

* :class:`CallOracle` follows the reference call by call (a sliding window of
  ``buffer_length`` samples, one ``demodulate(block)`` per ``block_size``
  samples).  It is what ``bench.py`` times as the CPU baseline ("port").
* :func:`stream_oracle` is the *global* formulation the CUDA design uses: the
  whole capture is demodulated once into a decision-bit stream ``Q[g]`` and
  every preamble position ``g`` is mapped back onto the reference's calls.
  The tests prove both produce identical packet lists.

The Go mirror (:class:`GoCallOracle`, :func:`go_freq_error`) is restated from
``dsp/dsp.go`` and ``protocol/protocol.go``.  PARITY UNPINNED for that part: there is
no Go toolchain in the build container, so no output of the Go code itself exists to
check it against; it shares every stage but FIR summation order and the discriminator
epsilon with the pinned Python restatement.

Reference citations are ``file:line`` under ``/root/reference``.
"""
from __future__ import annotations

import struct
from dataclasses import dataclass
from typing import Iterable, List, Sequence, Tuple

import numpy as np

# --------------------------------------------------------------------------
# constants of the path
# --------------------------------------------------------------------------
#: fir9 taps, src/rtldavis/dsp.py:56-69 (Go: dsp/dsp.go:65-83)
FIR_TAPS = np.array(
    [0.017682261285, 0.048171339939, 0.122424706672, 0.197408519126,
     0.228626345955,
     0.197408519126, 0.122424706672, 0.048171339939, 0.017682261285],
    dtype=np.float64,
)
#: discriminator epsilon, src/rtldavis/dsp.py:88 (the Go mirror has none)
DISC_EPS = 1e-10
#: byte -> float map, src/rtldavis/dsp.py:26
LUT = (np.arange(256, dtype=np.float64) - 127.4) / 127.6
#: multipliers of rotate_fs4, src/rtldavis/dsp.py:46-49
ROT = np.array([1, 1j, -1, -1j], dtype=np.complex128)


@dataclass(frozen=True)
class OracleConfig:
    """Derived sizes of ``PacketConfig`` (src/rtldavis/dsp.py:101-125) with the
    values ``protocol.new_packet_config`` passes (src/rtldavis/protocol.py:68-76)."""

    symbol_length: int = 14
    block_size: int = 8192
    bit_rate: int = 19200
    preamble_symbols: int = 16
    packet_symbols: int = 80
    preamble: str = "1100101110001001"

    @property
    def sample_rate(self) -> int:
        return self.bit_rate * self.symbol_length

    @property
    def preamble_length(self) -> int:
        return self.preamble_symbols * self.symbol_length

    @property
    def packet_length(self) -> int:
        return self.packet_symbols * self.symbol_length

    @property
    def buffer_length(self) -> int:
        return (self.packet_length // self.block_size + 2) * self.block_size

    @property
    def window_blocks(self) -> int:
        return self.buffer_length // self.block_size

    @property
    def preamble_bits(self) -> np.ndarray:
        return np.array([int(c) for c in self.preamble], dtype=np.uint8)

    @property
    def packet_bytes(self) -> int:
        return (self.packet_symbols + 7) // 8


# --------------------------------------------------------------------------
# stage functions (one per reference function)
# --------------------------------------------------------------------------
def bytes_to_complex(u8: np.ndarray) -> np.ndarray:
    """ByteToCmplxLUT.execute, src/rtldavis/dsp.py:28-39."""
    u8 = np.asarray(u8)
    if u8.size % 2:
        raise ValueError("Incompatible array sizes")
    out = np.empty(u8.size // 2, dtype=np.complex128)
    out.real = LUT[u8[0::2]]
    out.imag = LUT[u8[1::2]]
    return out


def rotate_quarter(x: np.ndarray) -> np.ndarray:
    """rotate_fs4, src/rtldavis/dsp.py:42-49: sample i times j**(i mod 4).

    The products are exact (components swap and change sign), so multiplying by
    the tiled table is bit-identical to the reference's four strided assignments.
    """
    out = np.array(x, dtype=np.complex128, copy=True)
    out[1::4] = x[1::4] * 1j
    out[2::4] = x[2::4] * -1
    out[3::4] = x[3::4] * -1j
    return out


def fir9_valid(r: np.ndarray, n_out: int) -> np.ndarray:
    """fir9, src/rtldavis/dsp.py:52-73: ``np.convolve(.., 'valid')[:n_out]``.

    ``out[k] = sum_m taps[m] * r[k+m]`` (the taps are symmetric, so convolution
    and correlation coincide)."""
    return np.convolve(r, FIR_TAPS, mode="valid")[:n_out]


def discriminate(f: np.ndarray) -> np.ndarray:
    """discriminate, src/rtldavis/dsp.py:76-90 on ``f`` of length n+1 -> n."""
    a, b = f[:-1], f[1:]
    return (a.imag * b.real - a.real * b.imag) / (a.real ** 2 + a.imag ** 2 + DISC_EPS)


def quantize_literal(d: Sequence[float]) -> np.ndarray:
    """quantize, src/rtldavis/dsp.py:93-98: IEEE-754 bit 63 of every value,
    taken one sample at a time exactly as the reference does (this loop is
    ~90 % of the reference's run time and is kept for the timed baseline)."""
    out = np.empty(len(d), dtype=np.uint8)
    for i, v in enumerate(d):
        out[i] = struct.unpack("<Q", struct.pack("<d", v))[0] >> 63
    return out


def quantize_fast(d: np.ndarray) -> np.ndarray:
    """Same bit as :func:`quantize_literal` (``np.signbit`` *is* bit 63)."""
    return np.signbit(d).astype(np.uint8)


def swap_bit_order(b: int) -> int:
    """protocol.swap_bit_order, src/rtldavis/protocol.py:79-83."""
    r = 0
    for i in range(8):
        r |= ((b >> i) & 1) << (7 - i)
    return r


_SWAP = np.array([swap_bit_order(i) for i in range(256)], dtype=np.uint8)


def crc_table(poly: int) -> np.ndarray:
    """CRC._new_table, src/rtldavis/crc.py:28-42 (MSB-first, 16 bit)."""
    tbl = np.zeros(256, dtype=np.uint16)
    for i in range(256):
        reg = i << 8
        for _ in range(8):
            reg = ((reg << 1) ^ poly) & 0xFFFF if reg & 0x8000 else (reg << 1) & 0xFFFF
        tbl[i] = reg
    return tbl


_CCITT = crc_table(0x1021)


def crc16(data: Iterable[int], init: int = 0, table: np.ndarray = _CCITT) -> int:
    """CRC.checksum, src/rtldavis/crc.py:19-26.  The path's instance is
    ``CRC("CCITT-16", 0, 0x1021, 0)`` (src/rtldavis/protocol.py:117)."""
    reg = init & 0xFFFF
    for byte in data:
        reg = ((reg << 8) & 0xFFFF) ^ int(table[(reg >> 8) ^ int(byte)])
    return reg


def crc_gate(air_bytes: Sequence[int]) -> Tuple[bytes, bool]:
    """Parser.parse lines 290-297 (src/rtldavis/protocol.py): bit-reverse every
    byte, then the packet is valid iff the CRC over bytes 2..9 is zero."""
    data = bytes(int(_SWAP[int(b)]) for b in air_bytes)
    return data, crc16(data[2:]) == 0


# --------------------------------------------------------------------------
# result record shared by both formulations
# --------------------------------------------------------------------------
@dataclass(frozen=True)
class OraclePacket:
    call: int          # which demodulate() call returned it (0-based)
    index: int         # Packet.index, src/rtldavis/dsp.py:14
    data: bytes        # Packet.data (air order, 10 bytes)
    crc_ok: bool       # outcome of crc_gate on it
    rssi: float = 0.0
    snr: float = 0.0

    @property
    def key(self) -> Tuple[int, int, bytes]:
        return (self.call, self.index, self.data)


# --------------------------------------------------------------------------
# formulation A: call by call, as the reference runs
# --------------------------------------------------------------------------
class CallOracle:
    """Demodulator, src/rtldavis/dsp.py:127-253, restated.

    State is kept as the reference keeps it (sliding buffers moved by
    ``block_size`` per call) so that windowing quirks -- duplicates at
    ``index == block_size``, RSSI windows that index the one-block ``filtered``
    buffer with two-block positions -- come out the same.
    """

    def __init__(self, cfg: OracleConfig = OracleConfig(), literal_quantize: bool = False):
        self.cfg = cfg
        self._quant = quantize_literal if literal_quantize else quantize_fast
        self.reset()
        self.calls = 0

    def reset(self) -> None:
        c = self.cfg
        self.tail9 = np.zeros(9, dtype=np.complex128)        # iq[0:9], dsp.py:130
        self.filtered = np.zeros(c.block_size + 1, dtype=np.complex128)
        self.discriminated = np.zeros(2 * c.block_size, dtype=np.float64)
        self.quantized = np.zeros(c.buffer_length, dtype=np.uint8)

    def demodulate(self, block: np.ndarray) -> List[OraclePacket]:
        c = self.cfg
        B = c.block_size
        block = np.asarray(block)
        if np.iscomplexobj(block):
            if block.size != B:
                raise ValueError("Incompatible array sizes")          # dsp.py:145-149
            x = block.astype(np.complex128)
        else:
            if block.size != 2 * B:
                raise ValueError("Incompatible array sizes")          # dsp.py:32-36
            x = bytes_to_complex(block)

        r = rotate_quarter(x)                                          # dsp.py:160
        iq = np.concatenate([self.tail9, r])                           # dsp.py:154,159
        f_new = self._fir(iq, B)                                       # dsp.py:161
        f = np.concatenate([self.filtered[-1:], f_new])                # dsp.py:155
        d_new = self._disc(f)                                          # dsp.py:162
        self.tail9 = iq[-9:].copy()
        self.filtered = f
        self.discriminated = np.concatenate([self.discriminated[B:], d_new])
        self.quantized = np.concatenate([self.quantized[B:], self._quant(d_new)])

        out = self._slice(self._search())
        self.calls += 1
        return out

    _fir = staticmethod(fir9_valid)
    _disc = staticmethod(discriminate)

    # dsp.py:171-188 -- every occurrence, offset-major then ascending
    def _search(self) -> List[int]:
        c = self.cfg
        pat = c.preamble_bits.tobytes()
        found: List[int] = []
        for phase in range(c.symbol_length):
            lane = self.quantized[phase::c.symbol_length].tobytes()
            at = lane.find(pat)
            while at >= 0:
                found.append(at * c.symbol_length + phase)
                at = lane.find(pat, at + 1)
        return found

    # dsp.py:190-246
    def _slice(self, found: List[int]) -> List[OraclePacket]:
        c = self.cfg
        taken = set()
        out: List[OraclePacket] = []
        for q in found:
            if q > c.block_size:                                       # dsp.py:194
                continue
            bits = self.quantized[q:q + c.packet_length:c.symbol_length][:c.packet_symbols]
            data = np.packbits(bits).tobytes()                         # MSB first, dsp.py:197-200
            if data in taken:
                continue
            taken.add(data)
            rssi, snr = self._power(q)
            out.append(OraclePacket(self.calls, q, data, crc_gate(data)[1], rssi, snr))
        return out

    # dsp.py:207-236 (Python only)
    def _power(self, q: int) -> Tuple[float, float]:
        c = self.cfg
        sig = self.filtered[q:q + c.preamble_length]
        lo = max(0, q - c.preamble_length)
        if q > lo:
            noise = float(np.mean(np.abs(self.filtered[lo:q]) ** 2))
        else:
            noise = 1e-9
        with np.errstate(invalid="ignore", divide="ignore"):
            power = float(np.mean(np.abs(sig) ** 2)) if sig.size else float("nan")
        import math
        rssi = 10 * math.log10(power) if power > 0 else -120
        snr = 10 * math.log10(power / noise) if noise > 0 else 50
        return rssi, snr


# --------------------------------------------------------------------------
# the Go mirror (parity unpinned, see the module header)
# --------------------------------------------------------------------------
def go_fir9(iq: np.ndarray, n_out: int) -> np.ndarray:
    """FIR9, dsp/dsp.go:65-83: symmetric taps folded, pairs added before the multiply."""
    c = FIR_TAPS
    w = [iq[m:m + n_out] for m in range(9)]
    acc = (w[0] + w[8]) * c[0]
    acc = acc + (w[1] + w[7]) * c[1]
    acc = acc + (w[2] + w[6]) * c[2]
    acc = acc + (w[3] + w[5]) * c[3]
    acc = acc + w[4] * c[4]
    return acc


def go_discriminate(f: np.ndarray) -> np.ndarray:
    """Discriminate, dsp/dsp.go:85-95: no epsilon; 0/0 (zero history) is NaN."""
    a, b = f[:-1], f[1:]
    with np.errstate(invalid="ignore", divide="ignore"):
        return (a.imag * b.real - a.real * b.imag) / (a.real * a.real + a.imag * a.imag)


def go_quantize(d: np.ndarray) -> np.ndarray:
    """Quantize, dsp/dsp.go:97-103: bit 63.  The NaN of 0/0 has a platform-defined sign
    (set on amd64, clear on arm64); it is taken as clear here (the samples this
    touches are global samples 0 and 1 of a fresh stream)."""
    return np.where(np.isnan(d), 0, np.signbit(d)).astype(np.uint8)


class GoCallOracle(CallOracle):
    """Demodulator.Demodulate, dsp/dsp.go:271-291 (block size 512, dsp.go:212).  Pack +
    Search (dsp.go:105-131,258-265) look only at the first block_size + preamble_length
    samples, which is every position Slice keeps (dsp.go:148: qIdx <= BlockSize); RSSI/SNR
    do not exist in Go and are whatever the Python formula gives."""

    def __init__(self, cfg: OracleConfig = OracleConfig(block_size=512)):
        super().__init__(cfg)
        self._quant = go_quantize

    _fir = staticmethod(go_fir9)
    _disc = staticmethod(go_discriminate)


def go_freq_error(discriminated: np.ndarray, index: int, cfg: OracleConfig):
    """Parser.Parse, protocol/protocol.go:154-168: mean over symbols 8..24 after the
    preamble index, then -int(9600 + mean * Fs / 2 pi).  Returns (mean, error); error is
    None when the mean is NaN (Go's int(NaN) is platform-defined)."""
    import math
    lower = index + 8 * cfg.symbol_length
    upper = index + 24 * cfg.symbol_length
    tail = discriminated[lower:upper]
    mean = float(np.sum(tail) / len(tail)) if not np.isnan(tail).any() else float("nan")
    if math.isnan(mean):
        return mean, None
    return mean, -int(9600 + (mean * float(cfg.sample_rate)) / (2 * math.pi))


def run_calls(u8: np.ndarray, cfg: OracleConfig = OracleConfig(),
              literal_quantize: bool = False) -> List[OraclePacket]:
    """Feed a whole uint8 capture block by block (the trailing partial block is
    never consumed, as a caller of the reference could not pass it)."""
    o = CallOracle(cfg, literal_quantize)
    step = 2 * cfg.block_size
    out: List[OraclePacket] = []
    for k in range(len(u8) // step):
        out.extend(o.demodulate(u8[k * step:(k + 1) * step]))
    return out


# --------------------------------------------------------------------------
# formulation B: whole stream at once + mapping onto calls
# --------------------------------------------------------------------------
def stream_filtered(u8: np.ndarray) -> np.ndarray:
    """F[g] for g = 0..n-1 with F[g] = sum_m taps[m] * r[g-9+m], r[<0] = 0
    (the reference's block pipeline is this, because block_size % 4 == 0 keeps
    the rotation phase continuous: dsp.py:159-161)."""
    r = rotate_quarter(bytes_to_complex(u8))
    padded = np.concatenate([np.zeros(9, dtype=np.complex128), r])
    return fir9_valid(padded, r.size)


def stream_discriminated(u8: np.ndarray) -> np.ndarray:
    """D[g] = disc(F[g-1], F[g]) with F[-1] = 0 (dsp.py:155,162)."""
    f = stream_filtered(u8)
    return discriminate(np.concatenate([np.zeros(1, dtype=np.complex128), f]))


def stream_bits(u8: np.ndarray) -> np.ndarray:
    return quantize_fast(stream_discriminated(u8))


def find_preambles(bits: np.ndarray, cfg: OracleConfig = OracleConfig()) -> np.ndarray:
    """All g with bits[g + 14*i] == preamble[i] for i < 16, ascending."""
    sl, pat = cfg.symbol_length, cfg.preamble_bits
    n = bits.size - (len(pat) - 1) * sl
    if n <= 0:
        return np.zeros(0, dtype=np.int64)
    ok = np.ones(n, dtype=bool)
    for i, p in enumerate(pat):
        ok &= bits[i * sl:i * sl + n] == p
    return np.flatnonzero(ok).astype(np.int64)


def map_to_calls(positions: Iterable[int], bits: np.ndarray, n_calls: int,
                 cfg: OracleConfig = OracleConfig(), first_call: int = 0) -> List[OraclePacket]:
    """Turn global preamble positions into the reference's per-call packets.

    Call k sees the window of global samples [(k-W+1)B, (k+1)B) (W =
    buffer_length / block_size) and returns positions whose window index is
    <= B (dsp.py:194), ordered by (index mod 14, index) (dsp.py:175-186) with
    duplicate payloads inside one call dropped (dsp.py:203-205)."""
    B, W, sl = cfg.block_size, cfg.window_blocks, cfg.symbol_length
    per_call: dict = {}
    for g in positions:
        g = int(g)
        # window index i = g - (k-W+1)B must satisfy 0 <= i <= B
        k_hi = g // B + W - 1
        for k in (k_hi, k_hi - 1):
            i = g - (k - W + 1) * B
            if 0 <= i <= B and first_call <= k < n_calls:
                per_call.setdefault(k, []).append((i % sl, i, g))
    out: List[OraclePacket] = []
    for k in sorted(per_call):
        taken = set()
        for _, i, g in sorted(per_call[k]):
            sym = bits[g:g + cfg.packet_length:sl][:cfg.packet_symbols]
            data = np.packbits(sym).tobytes()
            if data in taken:
                continue
            taken.add(data)
            out.append(OraclePacket(k, i, data, crc_gate(data)[1]))
    return out


def stream_oracle(u8: np.ndarray, cfg: OracleConfig = OracleConfig()) -> List[OraclePacket]:
    """Whole-capture equivalent of :func:`run_calls` (without RSSI/SNR)."""
    n_calls = len(u8) // (2 * cfg.block_size)
    u8 = u8[: n_calls * 2 * cfg.block_size]
    bits = stream_bits(u8)
    return map_to_calls(find_preambles(bits, cfg), bits, n_calls, cfg)
