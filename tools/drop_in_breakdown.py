#!/usr/bin/env python3
"""Where the time of one single-block drop-in call goes (host submit / device chain / wait / copy)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from rtldavis_b200 import dsp, protocol, synth

cap = synth.make_capture(5, 64 * 8192, 8, snr_db=20.0)
bd = dsp.BatchDemodulator(protocol.new_packet_config(14), 1, max_blocks=1)
blocks = [cap.iq[k * 16384:(k + 1) * 16384].reshape(1, -1) for k in range(64)]
for b in blocks[:8]:
    bd.submit(b); bd.fetch()
rows = []
for rep in range(5):
    for b in blocks:
        t0 = time.perf_counter(); bd.submit(b); t1 = time.perf_counter(); bd.fetch(copy=False); t2 = time.perf_counter()
        st = bd.stats()
        rows.append(((t1 - t0) * 1e6, (t2 - t1) * 1e6, st.host_submit_ms * 1e3, st.total_ms * 1e3, st.host_wait_ms * 1e3, st.host_copy_ms * 1e3, st.graph_replays))
r = np.median(np.array(rows), axis=0)
print(f"submit() {r[0]:.0f} us (library {r[2]:.0f}), fetch() {r[1]:.0f} us (wait {r[4]:.0f}, copy {r[5]:.0f}); device chain {r[3]:.0f} us; graph replays {r[6]:.0f}")
