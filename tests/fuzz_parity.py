#!/usr/bin/env python3
"""Randomised parity soak: random stream counts, block counts, SNRs, chunkings, host/device input,
every record against the C oracle.  Usage: python tests/fuzz_parity.py [cases] [seed]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from oracle import c_oracle as CO
from rtldavis_b200 import dsp, protocol, synth

def rows(h):
    return [(int(x["stream"]), int(x["call"]), int(x["index"]), x["data"].tobytes(), bool(x["crc_ok"])) for x in h]

def main():
    cases = int(sys.argv[1]) if len(sys.argv) > 1 else 30
    rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 1)
    t0 = time.time()
    total = 0
    for c in range(cases):
        ns = int(rng.choice([1, 2, 3, 7, 31, 63, 64, 65, 97, 130]))
        nb = int(rng.integers(2, 13))
        snr = float(rng.uniform(2.0, 30.0))
        iq, _ = synth.make_batch(ns, nb * 8192, int(rng.integers(1, 5)), snr, seed0=int(rng.integers(0, 10**6)),
                                 df_range_hz=float(rng.uniform(0, 3000)), ppm_range=float(rng.uniform(0, 50)))
        flood = nb >= 8 and rng.random() < 0.2       # one stream of nothing but sync words: list re-runs
        if flood:
            iq[int(rng.integers(0, ns))] = synth.make_sync_flood(nb * 8192, int(rng.integers(0, 10**6)),
                                                                 df_hz=float(rng.uniform(-2000, 2000)))
        want = [(int(r["stream"]), int(r["call"]), int(r["index"]), r["data"].tobytes(), bool(r["crc_ok"]))
                for r in CO.batch_records(iq, per_stream_cap=1 << 16)]
        # random chunking
        cuts, k = [], 0
        while k < nb:
            step = int(rng.integers(1, nb - k + 1))
            cuts.append(step); k += step
        on_dev = bool(rng.integers(0, 2))
        # the flavor (davis_cfg.flavor) must not change a record: one case in three runs as Go
        go = c % 3 == 2
        if go:
            cfg = dsp.PacketConfig(19200, 14, 16, 80, "1100101110001001", block_size=8192, flavor="go")
        else:
            cfg = protocol.new_packet_config(14)
        kw = {} if on_dev or rng.random() < 0.5 else {"host_slots": int(rng.integers(1, 4)), "host_group": int(rng.integers(1, ns + 1))}
        bd = dsp.BatchDemodulator(cfg, ns, max_blocks=max(cuts), **kw)
        got, k, reruns = [], 0, 0
        for step in cuts:
            part = np.ascontiguousarray(iq[:, k * 16384:(k + step) * 16384])
            bd.submit(torch.from_numpy(part).cuda() if on_dev else part)
            got += rows(bd.fetch())
            reruns += int(bd.stats().reruns)
            k += step
        bd.close()
        ok = sorted(got) == sorted(want)
        total += len(want)
        print(f"case {c:3d}: streams {ns:4d} blocks {nb:2d} snr {snr:5.1f} cuts {cuts} {'dev' if on_dev else 'host'} {'go' if go else 'py'} "
              f"records {len(want):5d}{' flood' if flood else ''}{f' reruns {reruns}' if reruns else ''} {'ok' if ok else 'MISMATCH'}", flush=True)
        if not ok:
            sw, sg = set(want), set(got)
            print("  missing", sorted(sw - sg)[:5]); print("  extra", sorted(sg - sw)[:5])
            sys.exit(1)
    print(f"{cases} cases, {total} records, all equal to the oracle ({time.time() - t0:.1f} s)")

if __name__ == "__main__":
    main()
