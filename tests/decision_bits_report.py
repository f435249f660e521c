#!/usr/bin/env python3
"""Gate (ii) of SURVEY 8(d) at scale (a script, not collected by pytest): the decision bits the
CUDA path holds for whole 10 s streams against the oracle's float64 sign bits, sample by sample.
Every disagreement is counted and must sit where the discriminator is within 1e-5 of zero on its
natural scale max(|D|, |F[g]|/|F[g-1]|) (tests/test_gpu_parity.py::test_decision_bits_and_fp32_discriminator).
Usage: python tests/decision_bits_report.py [streams=64] [blocks=328]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from oracle import c_oracle as CO
from rtldavis_b200 import _lib, dsp, protocol


def main():
    ns = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    n_blocks = int(sys.argv[2]) if len(sys.argv) > 2 else 328
    n_samples = n_blocks * 8192
    L = _lib.load()
    snrs = [4.0, 10.0, 20.0, 30.0]
    total = bad_total = outside = 0
    print(f"{'SNR dB':>6} {'streams':>7} {'samples':>12} {'disagree':>9} {'outside 1e-5 band':>18}")
    for i, snr in enumerate(snrs):
        per = ns // len(snrs)
        buf = torch.empty((per, 2 * n_samples), dtype=torch.uint8, device="cuda")
        _lib.check(L.davis_synth_fill(buf.data_ptr(), per, 2 * n_samples, n_samples, 9100 + i, 40,
                                      snr, 3000.0, 50.0, 0, None))
        torch.cuda.synchronize()
        bd = dsp.BatchDemodulator(protocol.new_packet_config(14), per, max_blocks=n_blocks)
        bd.submit(buf)
        bd.fetch()
        host = buf.cpu().numpy()
        n = nbad = nout = 0
        for s in range(per):
            first, bits = bd.debug_bits(s)
            ref_bits, ref_disc, ref_filt = CO.stream_bits(host[s], want_disc=True, want_filt=True)
            lo = max(first, 16)
            got = bits[lo - first:lo - first + ref_bits.size - lo]
            want = ref_bits[lo:lo + got.size]
            bad = np.flatnonzero(got != want)
            if bad.size:
                mag = np.abs(ref_filt)
                prev = np.concatenate([[0.0], mag[:-1]])
                scale = np.maximum(np.abs(ref_disc), mag * prev / (prev * prev + 1e-10))
                nout += int(np.sum(np.abs(ref_disc[lo + bad]) > 1e-5 * scale[lo + bad]))
            n += want.size
            nbad += bad.size
        bd.close()
        print(f"{snr:6.0f} {per:7d} {n:12d} {nbad:9d} {nout:18d}", flush=True)
        total += n; bad_total += nbad; outside += nout
    print(f"{total} samples, {bad_total} fp32 decisions differ from the float64 sign bit "
          f"({bad_total / max(total, 1):.2e}), {outside} of them outside the 1e-5 band")
    return 1 if outside else 0


if __name__ == "__main__":
    sys.exit(main())
