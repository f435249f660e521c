#!/usr/bin/env python3
"""Generate tests/golden/*.npz by running the UNMODIFIED reference.

Run in the build container only (``/root/reference`` does not exist on the GPU
box):  ``python oracle/gen_golden.py``.  Nothing from the reference is copied;
it is imported from where it lies and only its inputs/outputs are stored.

Per fixture:
  iq          uint8 interleaved I/Q fed to the reference, block by block
  block       block_size passed to PacketConfig
  pk_call / pk_index / pk_data / pk_crc / pk_rssi / pk_snr
              every dsp.Packet the reference's Demodulator.demodulate returned
              (src/rtldavis/dsp.py:139-169), in order, with the verdict of the
              CRC gate of Parser.parse (src/rtldavis/protocol.py:290-297)
  bits        np.packbits of the quantized stream (new half of ``quantized``
              after every call, concatenated)
  disc_lo/hi, disc   a slice of the float64 ``discriminated`` stream
  truth_start / truth_msg   where the generator put packets (not reference output)

The 10 s fixtures (``*_10s``) do not store the IQ (5.4 MB each): they store the generator's
recipe (``synth.make_capture`` keyword arguments as JSON) and the SHA-256 of the bytes it must
reproduce, next to the reference's rows and ``np.packbits(quantized)`` for the whole capture.
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, "/root/reference/src")
sys.path.insert(0, ROOT)

from rtldavis import crc as ref_crc            # noqa: E402
from rtldavis import dsp as ref_dsp            # noqa: E402
from rtldavis import protocol as ref_protocol  # noqa: E402

from rtldavis_b200 import synth                # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def run_reference(iq: np.ndarray, block: int, as_complex: bool = False):
    cfg = ref_protocol.new_packet_config(14)
    if block != cfg.block_size:
        cfg = ref_dsp.PacketConfig(cfg.bit_rate, cfg.symbol_length, cfg.preamble_symbols,
                                   cfg.packet_symbols, cfg.preamble, block_size=block)
    dem = ref_dsp.Demodulator(cfg)
    gate = ref_crc.CRC("CCITT-16", 0, 0x1021, 0)
    rows, bits, disc = [], [], []
    for k in range(iq.size // (2 * block)):
        blk = iq[k * 2 * block:(k + 1) * 2 * block]
        if as_complex:
            lut = ref_dsp.ByteToCmplxLUT()
            z = np.zeros(block, dtype=np.complex128)
            lut.execute(blk, z)
            blk = z.astype(np.complex64)
        for p in dem.demodulate(blk):
            host = bytes(ref_protocol.swap_bit_order(int(b)) for b in p.data)
            rows.append((k, p.index, bytes(p.data), int(gate.checksum(host[2:])) == 0,
                         p.rssi, p.snr))
        bits.append(dem.quantized[cfg.buffer_length - block:].copy())
        disc.append(dem.discriminated[block:].copy())
    return rows, np.concatenate(bits), np.concatenate(disc)


def save(name, cap, block, disc_lo, disc_hi, as_complex=False):
    rows, bits, disc = run_reference(cap.iq, block, as_complex)
    np.savez_compressed(
        os.path.join(OUT, name + ".npz"),
        iq=cap.iq, block=np.int64(block), as_complex=np.bool_(as_complex),
        pk_call=np.array([r[0] for r in rows], dtype=np.int64),
        pk_index=np.array([r[1] for r in rows], dtype=np.int64),
        pk_data=np.array([list(r[2]) for r in rows], dtype=np.uint8).reshape(-1, 10),
        pk_crc=np.array([r[3] for r in rows], dtype=np.bool_),
        pk_rssi=np.array([r[4] for r in rows], dtype=np.float64),
        pk_snr=np.array([r[5] for r in rows], dtype=np.float64),
        bits=np.packbits(bits), n_bits=np.int64(bits.size),
        disc_lo=np.int64(disc_lo), disc_hi=np.int64(disc_hi), disc=disc[disc_lo:disc_hi],
        truth_start=np.array([p.start for p in cap.packets], dtype=np.int64),
        truth_msg=np.array([list(p.message) for p in cap.packets], dtype=np.uint8).reshape(-1, 8),
    )
    ok = sum(r[3] for r in rows)
    print(f"{name}: {len(rows)} packets from the reference, {ok} CRC-valid, "
          f"{len(cap.packets)} embedded")


def save_seeded(name, block, n_samples, **recipe):
    """A whole 10 s capture: recipe + hash instead of the IQ; every full block is fed."""
    cap = synth.make_capture(n_samples=n_samples, **recipe)
    n_blocks = n_samples // block
    iq = cap.iq[: 2 * block * n_blocks]
    rows, bits, _ = run_reference(iq, block)
    np.savez_compressed(
        os.path.join(OUT, name + ".npz"),
        recipe=np.array(json.dumps(dict(recipe, n_samples=n_samples))),
        sha256=np.array(hashlib.sha256(cap.iq.tobytes()).hexdigest()),
        block=np.int64(block), n_blocks=np.int64(n_blocks),
        pk_call=np.array([r[0] for r in rows], dtype=np.int64),
        pk_index=np.array([r[1] for r in rows], dtype=np.int64),
        pk_data=np.array([list(r[2]) for r in rows], dtype=np.uint8).reshape(-1, 10),
        pk_crc=np.array([r[3] for r in rows], dtype=np.bool_),
        pk_rssi=np.array([r[4] for r in rows], dtype=np.float64),
        pk_snr=np.array([r[5] for r in rows], dtype=np.float64),
        bits=np.packbits(bits), n_bits=np.int64(bits.size),
        truth_start=np.array([p.start for p in cap.packets], dtype=np.int64),
        truth_msg=np.array([list(p.message) for p in cap.packets], dtype=np.uint8).reshape(-1, 8),
    )
    print(f"{name}: {len(rows)} packets from the reference, {sum(r[3] for r in rows)} CRC-valid, "
          f"{len(cap.packets)} embedded, {n_blocks} blocks")


def main():
    os.makedirs(OUT, exist_ok=True)
    if "--only-10s" not in sys.argv:
        short_fixtures()
    ten_second_fixtures()


def ten_second_fixtures():
    """What the short fixtures do not pin: BASELINE config 1 as it is
    stated (10 s, 40 packets, station 0, high SNR), a 6 dB capture (marginal decisions, many
    false candidates), a noise-only capture and config 2's literal +-50 ppm RF offset."""
    B, n = 8192, 2_688_000                  # 10 s at 268.8 kS/s: 328 full blocks + a 1024-sample tail
    save_seeded("config1_10s", B, n, seed=1001, n_packets=40, snr_db=30.0, station=0)
    save_seeded("snr6_10s", B, n, seed=1002, n_packets=40, snr_db=6.0, station=3, df_hz=1500.0,
                clock_ppm=-30.0)
    save_seeded("noise_10s", B, n, seed=1003, n_packets=0, noise_only=True)
    save_seeded("rf50ppm_10s", B, n, seed=1004, n_packets=40, snr_db=30.0, station=6,
                df_hz=45750.0, clock_ppm=50.0)


def short_fixtures():
    B = 8192
    # (1) high SNR; one burst placed so that its sync word starts within a few
    #     samples of a block boundary (duplicate reports in consecutive calls).
    sync_off = 32 * 14                      # AA AA AA AA precede the sync word
    starts = [3000, 3 * B - sync_off - 9, 6 * B + 1234, 8 * B + 5000]
    save("hi_snr_8192", synth.make_capture(11, 12 * B, 4, 30.0, station=0, starts=starts),
         B, 3 * B - 512, 3 * B + 1536)
    # (2) low SNR, residual carrier offset and sample-clock error
    save("lo_snr_8192", synth.make_capture(12, 8 * B, 3, 12.0, station=5, df_hz=2100.0,
                                           clock_ppm=-50.0), B, 0, 2048)
    # (3) the dataclass default / Go block size (window of 4 blocks)
    save("go_block_512", synth.make_capture(13, 96 * 512, 3, 25.0, station=2), 512, 0, 2048)
    # (4) complex64 input, the live pyrtlsdr format (src/rtldavis/dsp.py:144-150)
    save("complex_in_8192", synth.make_capture(14, 4 * B, 2, 30.0, station=1), B, 0, 1024,
         as_complex=True)


if __name__ == "__main__":
    main()
