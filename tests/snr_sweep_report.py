#!/usr/bin/env python3
"""BASELINE config 4 at full size (a script, not collected by pytest): 4 096 streams x 10 s,
AWGN 0..20 dB in 2 dB steps, residual carrier +-3 kHz, sample clock +-50 ppm, 40 packets per
stream.  Per SNR: every record of the CUDA path (CRC-valid or not: stream, call, index, 10 bytes,
verdict) against the C oracle on a host copy of the same bytes, the share of embedded packets that
come back CRC-valid, and how many records needed a decision bit corrected by the fp64 re-check.
Usage: python tests/snr_sweep_report.py [streams_per_snr=372] [blocks=328] [snr,snr,...]
With an explicit SNR list the stream count is not topped up to 4 096 (config 2 at full size is
`snr_sweep_report.py 1024 328 20`: bench.py's workload, every record against the oracle)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from oracle import c_oracle as CO
from rtldavis_b200 import _lib, dsp, protocol


def rows(rec):
    return [(int(r["stream"]), int(r["call"]), int(r["index"]), r["data"].tobytes(), bool(r["crc_ok"]))
            for r in rec]


def main():
    per = int(sys.argv[1]) if len(sys.argv) > 1 else 372
    n_blocks = int(sys.argv[2]) if len(sys.argv) > 2 else 328
    n_packets, B = 40, 8192
    n_samples = n_blocks * B
    snrs = [int(x) for x in sys.argv[3].split(",")] if len(sys.argv) > 3 else list(range(0, 21, 2))
    top_up = 0 if len(sys.argv) > 3 else max(0, 4096 - per * len(snrs))
    L = _lib.load()
    cfg = protocol.new_packet_config(14)
    print(f"{len(snrs)} SNRs x {per} streams (+{top_up} at the last) x {n_blocks} blocks of {B} samples, "
          f"{n_packets} packets per stream")
    print(f"{'SNR dB':>6} {'streams':>7} {'records':>8} {'mismatch':>8} {'CRC-valid':>9} {'embedded':>8} "
          f"{'recovered':>9} {'fp64-fixed':>10} {'oracle s':>8}")
    tot_rec = tot_bad = tot_streams = 0
    for i, snr in enumerate(snrs):
        ns = per + (top_up if i == len(snrs) - 1 else 0)
        seed = 7000 + i
        buf = torch.empty((ns, 2 * n_samples), dtype=torch.uint8, device="cuda")
        _lib.check(L.davis_synth_fill(buf.data_ptr(), ns, 2 * n_samples, n_samples, seed, n_packets,
                                      float(snr), 3000.0, 50.0, 0, None))
        torch.cuda.synchronize()
        msgs = np.zeros((ns, n_packets, 8), dtype=np.uint8)
        starts = np.zeros((ns, n_packets), dtype=np.int64)
        _lib.check(L.davis_synth_truth(ns, n_samples, seed, n_packets, float(snr),
                                       msgs.ctypes.data, starts.ctypes.data))
        bd = dsp.BatchDemodulator(cfg, ns, max_blocks=n_blocks)
        bd.submit(buf)
        hits = bd.fetch()
        bd.close()
        host = buf.cpu().numpy()
        del buf
        t0 = time.time()
        want = rows(CO.batch_records(host))
        t_or = time.time() - t0
        got = rows(hits)
        bad = len(set(got) ^ set(want)) + (0 if got == want else int(sorted(got) == sorted(want)))
        valid = {(int(h["stream"]), bytes(protocol.swap_bit_order(b) for b in h["data"][2:]))
                 for h in hits if h["crc_ok"]}
        rec = sum((s, msgs[s, k].tobytes()) in valid for s in range(ns) for k in range(n_packets))
        fixed = int((hits["flags"] & 1).sum())
        print(f"{snr:6d} {ns:7d} {len(want):8d} {bad:8d} {int(hits['crc_ok'].sum()):9d} {ns * n_packets:8d} "
              f"{100.0 * rec / (ns * n_packets):8.2f}% {fixed:10d} {t_or:8.1f}", flush=True)
        tot_rec += len(want); tot_bad += bad; tot_streams += ns
    print(f"{tot_streams} streams, {tot_rec} records, {tot_bad} differing from the oracle")
    return 1 if tot_bad else 0


if __name__ == "__main__":
    sys.exit(main())
