#!/usr/bin/env python3
"""Where one step of the hot path spends its time (device-resident input).
Usage: python tools/step_breakdown.py [streams] [blocks] [steps]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch

from rtldavis_b200 import _lib, dsp, protocol


def main():
    streams = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
    blocks = int(sys.argv[2]) if len(sys.argv) > 2 else 328
    steps = int(sys.argv[3]) if len(sys.argv) > 3 else 5
    L = _lib.load()
    n = blocks * 8192
    dev = torch.empty((streams, 2 * n), dtype=torch.uint8, device="cuda")
    _lib.check(L.davis_synth_fill(dev.data_ptr(), streams, 2 * n, n, 2024, 40, 20.0, 3000.0, 50.0, 0, None))
    torch.cuda.synchronize()
    bd = dsp.BatchDemodulator(protocol.new_packet_config(14), streams, max_blocks=blocks)
    for i in range(steps + 2):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        bd.submit(dev)
        t1 = time.perf_counter()
        hits = bd.fetch(copy=False)
        t2 = time.perf_counter()
        st = bd.stats()
        if i >= 2:
            print(f"step {i - 2}: wall {1e3 * (t2 - t0):7.2f} ms  submit {1e3 * (t1 - t0):6.2f}  fetch {1e3 * (t2 - t1):6.2f} | "
                  f"dev demod {st.demod_ms:5.2f} search {st.search_ms:5.2f} finish {st.finish_ms:5.2f} | "
                  f"host submit {st.host_submit_ms:5.2f} wait {st.host_wait_ms:5.2f} copy {st.host_copy_ms:5.2f} "
                  f"sort {st.host_sort_ms:5.2f} | cand {st.candidates} conf {st.confirmed} rec {st.records} "
                  f"hits {len(hits)} valid {int(hits['crc_ok'].sum())}")
    bd.close()


if __name__ == "__main__":
    main()
