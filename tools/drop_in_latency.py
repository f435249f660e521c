#!/usr/bin/env python3
"""Latency of the single-stream drop-in call: dsp.Demodulator.demodulate(one 8192-sample block)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from rtldavis_b200 import dsp, protocol, synth

cap = synth.make_capture(5, 64 * 8192, 8, snr_db=20.0)
dem = dsp.Demodulator(protocol.new_packet_config(14))
blocks = [cap.iq[k * 16384:(k + 1) * 16384] for k in range(64)]
for b in blocks[:8]:
    dem.demodulate(b)
t = []
n_pk = 0
for rep in range(5):
    for b in blocks:
        t0 = time.perf_counter()
        n_pk += len(dem.demodulate(b))
        t.append(time.perf_counter() - t0)
t = np.array(t) * 1e6
print(f"drop-in demodulate(): median {np.median(t):.0f} us, p95 {np.percentile(t, 95):.0f} us per 8192-sample block "
      f"({8192 / 268800 * 1e3:.1f} ms of signal), {n_pk} packets")
