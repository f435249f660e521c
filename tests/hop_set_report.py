#!/usr/bin/env python3
"""BASELINE config 3 at full size (a script, not collected by pytest): the 51 channels of the US
hop set (src/rtldavis/protocol.py:119-171) as 51 parallel streams of 10 s, eight transmitters each
sending one packet per 2.5625 s dwell (protocol.py:118) and hopping to the next channel of the
pattern, a fixed carrier offset per channel in +-3 kHz.  Fed block by block (328 calls, the way
worker.py:49 feeds the reference) and in ragged multi-block chunks: the carried per-stream state
must give the oracle's records both ways.
Usage: python tests/hop_set_report.py [blocks=328]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from oracle import c_oracle as CO
from rtldavis_b200 import dsp, protocol, synth


def rows(rec):
    return [(int(r["stream"]), int(r["call"]), int(r["index"]), r["data"].tobytes(), bool(r["crc_ok"]))
            for r in rec]


def main():
    n_blocks = int(sys.argv[1]) if len(sys.argv) > 1 else 328
    B, hop = 8192, int(round(2.5625 * 268800))
    iq, truth = synth.make_hop_set(51, n_blocks * B, n_transmitters=8, hop_samples=hop, snr_db=25.0, seed=51)
    want = rows(CO.batch_records(iq))
    sent = [(ch, e.message) for ch, pk in enumerate(truth) for e in pk
            if e.start + synth.BURST_BITS * synth.SPS < (n_blocks - 1) * B]      # the last block is never flushed
    valid = {(r[0], bytes(protocol.swap_bit_order(b) for b in r[3][2:])) for r in want if r[4]}
    print(f"51 streams x {n_blocks} blocks; {len(sent)} packets sent on {len({c for c, _ in sent})} channels; "
          f"oracle: {len(want)} records, {sum(r[4] for r in want)} CRC-valid, {sum(p in valid for p in sent)} of the sent recovered")
    cfg = protocol.new_packet_config(14)
    rng = np.random.default_rng(3)
    ragged, k = [], 0
    while k < n_blocks:
        c = int(min(n_blocks - k, rng.integers(1, 24)))
        ragged.append(c); k += c
    bad = 0
    for name, chunks in (("block by block", [1] * n_blocks), (f"ragged ({len(ragged)} chunks)", ragged)):
        bd = dsp.BatchDemodulator(cfg, 51, max_blocks=max(chunks))
        got, k, t0 = [], 0, time.time()
        for c in chunks:
            bd.submit(np.ascontiguousarray(iq[:, k * 2 * B:(k + c) * 2 * B]))
            got += rows(bd.fetch())
            k += c
        dt = time.time() - t0
        bd.close()
        same = sorted(got) == sorted(want)
        bad += not same
        print(f"  {name:28s} {len(got):5d} records, {'equal to the oracle' if same else 'MISMATCH'}  ({dt:.2f} s, host input)")
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
