#!/usr/bin/env python3
"""bench.py -- IQ Msample/s demodulated on B200, beside the reference's CPU path.

One "step" = one pass of the whole hot path (demodulate -> sync search -> confirm / slice /
CRC -> carry) over one batch of synthetic captures.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
                    [--streams S] [--blocks NB] [--snr DB] [--e2e-waves V]

N = 1 workload: BASELINE.json configs[1] -- 1,024 independent 10 s captures (328 blocks of
8192 samples each), mixed station IDs, +-50 ppm clock error, +-3 kHz residual carrier.
N > 1 (torchrun): N x 1,024 streams sharded by stream through rtldavis_b200.shard, no
data-path collective ("weak" scaling: 1,024 streams per GPU in the device-resident leg; in
the end-to-end leg the same global batch is split in proportion to the host->device copy
rate every GPU gets, probed at start-up).

value        device-resident throughput (inputs already in HBM), CUDA-event timed
e2e          the same metric through the public API with PINNED HOST input: H2D of every
             byte and D2H of the packet records inside the timed region
roofline     the demodulation kernel alone (frac) and the whole step (path_frac), 2 algorithmic
             bytes per complex sample
sustained    the device-resident loop again for >= 1.5 s with its own clock / power record
parity_check records of the first 8 streams of every rank against the C oracle, gathered
             with shard.gather_packets (outside the timed regions)
cpu_baseline the reference's own CPU code (oracle/_ref: the unmodified rtldavis modules,
             kind "ref") or, when that is not staged, the oracle's restatement (kind "port")
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "iq_msamples_per_s_demodulated"
UNIT = "Msample/s"
BLOCK = 8192
SAMPLE_STREAMS, SAMPLE_BLOCKS = 8, 41          # the bounded sample the CPU legs are timed on


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None,
                    help="timed steps (default: 400 = a timed region of >= 1 s; reference arm: 20)")
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="graft", choices=["graft", "reference"])
    ap.add_argument("--streams", type=int, default=1024, help="streams per GPU")
    ap.add_argument("--blocks", type=int, default=328, help="8192-sample blocks per stream (328 = 10 s)")
    ap.add_argument("--snr", type=float, default=20.0)
    ap.add_argument("--packets", type=int, default=40)
    ap.add_argument("--seed", type=int, default=2024)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-sustained", action="store_true")
    ap.add_argument("--uniform-e2e", action="store_true", help="end-to-end leg with the uniform split")
    ap.add_argument("--e2e-waves", type=int, default=1,
                    help="end-to-end: feed each GPU's streams in this many waves through a pinned ring of two "
                         "wave buffers (BASELINE config 5: --streams 8192 --e2e-waves 8)")
    ap.add_argument("--e2e-steps", type=int, default=0, help="steps of the end-to-end leg (default: --steps)")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    a = ap.parse_args()
    if a.steps is None:
        a.steps = 400 if a.impl == "graft" else 20
    return a


def workload_name(a):
    return (f"{a.streams} streams x {a.blocks} blocks x {BLOCK} samples "
            f"({a.blocks * BLOCK / 268800.0:.2f} s @268.8 kS/s) per GPU, {a.packets} packets/stream, "
            f"stations 0-7, SNR {a.snr:g} dB, df +-3 kHz, clock +-50 ppm")


def config_dict(a):
    """Identical in both arms (the driver compares them)."""
    return {"workload": workload_name(a), "streams_per_gpu": a.streams, "blocks_per_stream": a.blocks,
            "block_size": BLOCK,
            "l2": "inputs (%.2f GB per GPU) far exceed the 126 MB L2; no flush needed" % (a.streams * 2 * a.blocks * BLOCK / 1e9)}


# ---------------------------------------------------------------------------
# CPU legs: the reference's own Demodulator + CRC gate, one process per core, block by
# block (worker.py:49-50), on a bounded sample of the workload.
# ---------------------------------------------------------------------------
def _cpu_worker(job):
    iq, seconds, use_ref = job
    import logging
    logging.disable(logging.CRITICAL)              # the parser logs every packet
    step = 2 * BLOCK
    n_blocks = iq.size // step
    if use_ref:
        from oracle import ref_loader
        _, _, ref_protocol = ref_loader.load()
        parser = ref_protocol.Parser(symbol_length=14)

        def one(blk):
            return len(parser.parse(parser.demodulator.demodulate(blk)))      # worker.py:49-50
        reset = parser.demodulator.reset
    else:
        from oracle import davis_oracle as O
        o = O.CallOracle(O.OracleConfig(), literal_quantize=True)

        def one(blk):
            return sum(p.crc_ok for p in o.demodulate(blk))
        reset = o.reset
    done, packets = 0, 0
    t0 = time.perf_counter()
    while True:
        k = done % n_blocks
        if k == 0 and done:
            reset()
        packets += one(iq[k * step:(k + 1) * step])
        done += 1
        if time.perf_counter() - t0 >= seconds:
            break
    return done * BLOCK, time.perf_counter() - t0, packets


def cpu_kind():
    from oracle import ref_loader
    return ref_loader.kind()


def cpu_reference_rate(sample_iq: np.ndarray, seconds: float, cores: int):
    """Aggregate Msample/s of `cores` processes each demodulating its own stream
    of the sample for ~`seconds` (block by block, as the reference is driven)."""
    import multiprocessing as mp
    use_ref = cpu_kind() == "ref"
    ctx = mp.get_context("fork")
    jobs = [(sample_iq[i % sample_iq.shape[0]], seconds, use_ref) for i in range(cores)]
    t0 = time.perf_counter()
    with ctx.Pool(cores) as pool:
        res = pool.map(_cpu_worker, jobs)
    wall = time.perf_counter() - t0
    samples = sum(r[0] for r in res)
    busy = max(r[1] for r in res)
    return samples / busy / 1e6, samples, wall, sum(r[2] for r in res)


def cpu_sample_text(cores, seconds, kind, source="bytes of the GPU arm's own device generator, same seed"):
    what = ("the unmodified rtldavis Parser: demodulator.demodulate + parse (oracle/_ref)" if kind == "ref"
            else "the oracle's literal restatement of dsp.Demodulator (per-sample quantize kept)")
    return (f"first {SAMPLE_STREAMS} streams x {SAMPLE_BLOCKS} blocks of this workload ({source}); "
            f"{cores} processes x ~{seconds:g} s, block by block through {what}")


def c_port_rate(sample_iq: np.ndarray, cores: int):
    from oracle import c_oracle as CO
    CO.lib()
    t0 = time.perf_counter()
    CO.batch_records(sample_iq, block=BLOCK, n_threads=cores)
    dt = time.perf_counter() - t0
    n = sample_iq.shape[0] * (sample_iq.shape[1] // (2 * BLOCK)) * BLOCK
    return n / dt / 1e6


def usable_cores() -> int:
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return os.cpu_count() or 1


# ---------------------------------------------------------------------------
# clocks
# ---------------------------------------------------------------------------
class ClockSampler:
    """Polls NVML (nvidia_ml_py) from a thread while the timed region runs: SM clock,
    power and throttle reasons every ~0.5 ms."""
    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown",
               0x4: "sw_power_cap"}

    def __init__(self, index: int):
        self.index, self.rows, self.h, self.nv = index, [], None, None
        self.stop_flag = threading.Event()
        self.thread = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            try:
                import torch
                uuid = str(torch.cuda.get_device_properties(index).uuid)
                if not uuid.startswith("GPU-"):
                    uuid = "GPU-" + uuid
                self.h = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode())
            except Exception:
                self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_sm = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.h = None

    def _poll(self):
        nv = self.nv
        i, pw = 0, 0.0
        while not self.stop_flag.is_set():
            try:
                sm = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                try:
                    rs = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    rs = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                if i % 8 == 0:                      # the power query is the slow one
                    pw = nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0
                self.rows.append((float(sm), pw, int(rs)))
            except Exception:
                pass
            i += 1
            time.sleep(0.0002)

    def start(self):
        if self.h is None:
            return
        self.thread = threading.Thread(target=self._poll, daemon=True)
        self.thread.start()

    def stop(self):
        if self.h is None or self.thread is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvml unavailable"], "samples": 0}
        self.stop_flag.set()
        self.thread.join(timeout=2)
        sm = [r[0] for r in self.rows]
        reasons = set()
        for r in self.rows:
            for bit, nm in self.REASONS.items():
                if r[2] & bit:
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": self.max_sm,
                "sm_mhz_min": min(sm) if sm else None,
                "power_w_max": max((r[1] for r in self.rows), default=None),
                "samples": len(sm), "reasons": sorted(reasons)}


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ---------------------------------------------------------------------------
# reference arm: the reference's own CPU implementation on all host cores
# ---------------------------------------------------------------------------
def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    cores = usable_cores()
    kind = cpu_kind()
    # Nothing of the product runs in this arm: the sample comes from the numpy generator (same
    # signal model, packet density, SNR and offsets as the GPU arm's device generator), so this
    # process never loads libdavis_b200.so and its workers fork from a process without CUDA.
    from rtldavis_b200 import synth
    sample_blocks = min(a.blocks, SAMPLE_BLOCKS)
    iq, _ = synth.make_batch(SAMPLE_STREAMS, sample_blocks * BLOCK,
                             max(1, a.packets * sample_blocks // max(a.blocks, 1)), a.snr, seed0=a.seed,
                             df_range_hz=3000.0, ppm_range=50.0)
    input_note = ("numpy generator rtldavis_b200.synth.make_batch (the signal model of the GPU arm's device "
                  "generator: same packet density, SNR, +-3 kHz, +-50 ppm); no CUDA library in this process")
    per_step = max(1.0, min(a.cpu_seconds, 150.0 / max(1, a.steps + a.warmup)))  # whole run <= ~2.5 min
    rates = []
    for i in range(a.warmup + a.steps):
        rate, samples, wall, _ = cpu_reference_rate(iq, per_step, cores)
        if i >= a.warmup:
            rates.append((rate, samples, wall))
    value = float(np.mean([r[0] for r in rates]))
    samples = int(np.mean([r[1] for r in rates]))
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": a.gpus,
        "steps": a.steps, "warmup": a.warmup,
        "ms_per_step": float(np.mean([r[2] for r in rates])) * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": config_dict(a),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind,
                         "sample": cpu_sample_text(cores, per_step, kind, "numpy generator, same signal model and seed"),
                         "samples_per_step": samples,
                         "input": input_note},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)
    return 0


# ---------------------------------------------------------------------------
# graft arm
# ---------------------------------------------------------------------------
def _rows(rec):
    return [(int(r["stream"]), int(r["call"]), int(r["index"]), r["data"].tobytes(), bool(r["crc_ok"]))
            for r in rec]


def run_graft(a):
    import torch
    import torch.distributed as dist

    from rtldavis_b200 import _lib, protocol, shard
    from rtldavis_b200._lib import HIT_DTYPE

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def all_max(x):
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def all_sum(x):
        t = torch.tensor([x], dtype=torch.int64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return int(t.item())

    L = _lib.load()
    n_samples = a.blocks * BLOCK
    row = 2 * n_samples
    seed = a.seed + 7919 * rank                         # every rank generates its own streams

    def synth_fill(buf, n, seed_):
        _lib.check(L.davis_synth_fill(buf.data_ptr(), n, row, n_samples, seed_, a.packets,
                                      a.snr, 3000.0, 50.0, local, None))
        torch.cuda.synchronize()

    dev = torch.empty((a.streams, row), dtype=torch.uint8, device="cuda")
    synth_fill(dev, a.streams, seed)

    cfg = protocol.new_packet_config(14)
    n_global = a.streams * world
    # The global batch goes through the sharding layer (uniform split: a.streams per rank).
    # Two sharded handles alternate (each carries its own copy of the per-stream state): batch i
    # is submitted before batch i-1 is fetched, so the device->host copy of one batch's packet
    # records and the host bookkeeping overlap the next batch's kernels.  Every kernel of both
    # handles runs on ONE stream, in order, so k_demod's event timer is that of a kernel running alone.
    sds = (shard.ShardedDemodulator(cfg, n_global, max_blocks=a.blocks, device=local),
           shard.ShardedDemodulator(cfg, n_global, max_blocks=a.blocks, device=local))
    assert sds[0].count == a.streams
    bds = (sds[0].local, sds[1].local)
    stream = torch.cuda.Stream(device=local)
    sptr = stream.cuda_stream
    assert sptr != 0

    def run_steps(n, collect, on_stream=sptr):
        pending, hits = None, None
        for i in range(n):
            cur = bds[i & 1]
            cur.submit(dev, cuda_stream=on_stream)
            if pending is not None:
                hits = pending.fetch(copy=False)
                collect(pending.stats())
            pending = cur
        hits = pending.fetch(copy=False)
        collect(pending.stats())
        return hits

    # ---- device-resident: `value` ----
    hits = run_steps(max(a.warmup, 2), lambda st: None)
    sampler = ClockSampler(local)
    barrier()
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    demod_ms, search_ms, finish_ms, launches = [], [], [], [0]

    def collect(st):
        demod_ms.append(st.demod_ms); search_ms.append(st.search_ms); finish_ms.append(st.finish_ms)
        launches[0] += st.launches

    e0.record(stream)
    hits = run_steps(a.steps, collect)
    e1.record(stream)
    barrier()
    clocks = sampler.stop()
    launches = launches[0]
    ms_step = all_max(e0.elapsed_time(e1)) / a.steps
    n_valid = all_sum(int(hits["crc_ok"].sum()))
    samples_step = a.streams * n_samples * world
    value = samples_step / (ms_step * 1e-3) / 1e6

    # ---- the same loop for >= 1.5 s: what clocks and power do under a sustained FP32-bound load ----
    sustained = None
    if not a.no_sustained:
        s_steps = max(a.steps, int(np.ceil(1500.0 / ms_step)))
        sampler2 = ClockSampler(local)
        barrier()
        sampler2.start()
        s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s0.record(stream)
        run_steps(s_steps, lambda st: None)
        s1.record(stream)
        barrier()
        sc = sampler2.stop()
        s_ms = all_max(s0.elapsed_time(s1)) / s_steps
        sustained = {"steps": s_steps, "ms_per_step": s_ms, "value": samples_step / (s_ms * 1e-3) / 1e6,
                     "unit": UNIT, "timed_region_s": s_ms * s_steps * 1e-3, "clocks": sc}

    # Extra, not the headline: the same loop with each handle on its own stream, so the finish
    # kernels of one batch run beside the FP32-bound k_demod of the other.
    run_steps(2, lambda st: None, on_stream=None)
    barrier()
    t0 = time.perf_counter()
    run_steps(a.steps, lambda st: None, on_stream=None)
    torch.cuda.synchronize()
    two_streams_ms = all_max(time.perf_counter() - t0) / a.steps * 1e3

    # ---- roofline: the demodulation kernel (CUDA events inside the library, on the launching
    #      stream, around that kernel only) and the whole step ----
    peak, peak_src = measured_peak()
    demod_avg = float(np.mean(demod_ms))
    alg_bytes = 2 * a.streams * n_samples
    ach = alg_bytes / (demod_avg * 1e-3) / 1e9
    path_ach = alg_bytes / (ms_step * 1e-3) / 1e9
    roofline = {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                "traffic": None, "peak_source": peak_src, "kernel": "k_demod",
                "limited_by": "FP32 issue rate (FMA pipe ~80 % busy, DRAM ~33 %): the HBM figure is the "
                              "north_star's roofline, not this kernel's binding resource",
                "kernel_ms": demod_avg,
                "path_achieved": path_ach, "path_frac": path_ach / peak,
                "behind_demod_ms": ms_step - demod_avg,
                "chain_search_ms_sum": float(np.mean(search_ms)),
                "chain_finish_ms_sum": float(np.mean(finish_ms)),
                "chain_note": "sums over the two search/finish chains, which overlap each other behind "
                              "k_demod; behind_demod_ms (ms_per_step - kernel_ms) is their critical path",
                "frac_of_nominal_8TBs": ach / 8000.0,
                "algorithmic_bytes_per_launch": alg_bytes}
    prof = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(prof):
        try:
            with open(prof) as f:
                cap = json.load(f)
            # the ncu capture is of the default workload; any other size has no capture
            if cap.get("algorithmic_bytes_per_launch") == alg_bytes:
                roofline["traffic"] = cap.get("k_demod_dram_bytes_per_launch")
        except Exception:
            pass

    # ---- parity: the first 8 streams of every rank against the C oracle, gathered through
    #      the sharding layer (fresh state, one batch; outside every timed region) ----
    from oracle import c_oracle as CO

    def oracle_records(rows_u8, local_ids, call0=0):
        """Records of the C oracle for the given rows, numbered as the handle numbers them."""
        rec = CO.batch_records(rows_u8, block=BLOCK)
        if call0:
            rec = rec[rec["call"] >= call0]
        orc = np.zeros(len(rec), dtype=HIT_DTYPE)
        for k in ("call", "index", "data", "crc_ok"):
            orc[k] = rec[k]
        orc["call"] -= call0
        orc["stream"] = np.asarray(local_ids, dtype=np.uint32)[rec["stream"]]
        return orc

    # 63 streams from the front of the rank's shard and its very last one (the shard edge)
    chk = sorted(set(range(min(63, a.streams))) | {a.streams - 1})
    n_chk = len(chk)
    sds[0].reset()
    sds[0].submit(dev)
    mine = sds[0].fetch()
    got_all = shard.gather_packets(mine[np.isin(mine["stream"], chk)], sds[0].first_stream)
    orc = oracle_records(dev[torch.tensor(chk, device=dev.device)].cpu().numpy(), chk)
    want_all = shard.gather_packets(orc, sds[0].first_stream)
    parity = None
    if rank == 0:
        g_rows, w_rows = _rows(got_all), _rows(want_all)
        parity = {"streams": n_chk * world, "records": len(w_rows),
                  "mismatch": 0 if g_rows == w_rows else max(1, len(set(g_rows) ^ set(w_rows))),
                  "crc_valid": sum(r[4] for r in w_rows),
                  "how": "fresh state, one batch through shard.ShardedDemodulator; records of the first "
                         f"{n_chk - 1} streams and the last stream of every rank gathered with shard.gather_packets and compared "
                         "(stream, call, index, 10 bytes, CRC verdict, order) with oracle/davis_oracle.c"}

    # ---- end to end: pinned host input, H2D + records D2H inside the timed region ----
    e2e = None
    if not a.no_e2e:
        weights, probe = None, None
        if world > 1 and not a.uniform_e2e:
            probe = shard.probe_h2d_weights(local)
            weights = probe
        waves = max(1, a.e2e_waves)
        e_steps = a.e2e_steps or min(a.steps, 20)        # ~0.1-0.2 s per step: 20 steps are >= 2 s
        mine_n = shard.partition(n_global, world, weights)[rank][1]
        per_wave = (mine_n + waves - 1) // waves
        wave_n = [min(per_wave, mine_n - w * per_wave) for w in range(waves) if mine_n - w * per_wave > 0]
        # pinned ring of (at most) two wave buffers, filled once from the device generator; a wave
        # reads ring[w % 2] -- every byte of every wave crosses the host link in every step
        ring = []
        for w in range(min(2, len(wave_n))):
            tmp = torch.empty((per_wave, row), dtype=torch.uint8, device="cuda")
            synth_fill(tmp, per_wave, seed + 101 * (w + 1))
            hbuf = torch.empty((per_wave, row), dtype=torch.uint8, pin_memory=True)
            hbuf.copy_(tmp)
            torch.cuda.synchronize()
            ring.append(hbuf.numpy())
            del tmp
        from rtldavis_b200 import dsp
        handles = [dsp.BatchDemodulator(cfg, n, device=local, max_blocks=a.blocks) for n in wave_n]

        last_views = [None] * len(wave_n)               # packet records of the last step, per wave

        def e2e_step():
            h2d = d2h = valid = 0
            for w, (bd2, n) in enumerate(zip(handles, wave_n)):
                bd2.submit(ring[w % 2][:n])
                h2 = bd2.fetch(copy=False)
                last_views[w] = h2
                st2 = bd2.stats()
                h2d += int(st2.h2d_bytes); d2h += int(st2.d2h_bytes); valid += int(h2["crc_ok"].sum())
            return h2d, d2h, valid

        e_warm = max(1, min(a.warmup, 2))
        for _ in range(e_warm):
            e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(e_steps):
            h2d, d2h, valid = e2e_step()
        torch.cuda.synchronize()
        mine_dt = time.perf_counter() - t0
        dt = all_max(mine_dt)
        # Parity of what was just timed: the records the LAST timed step returned for eight
        # streams of every wave (front, middle and end of the rank's share: the shard and
        # pipeline-group edges) against the C oracle on the same pinned bytes.  The handles carry
        # their state from step to step and every step feeds the same bytes, so the oracle is run
        # over two passes of a row and its second pass is compared (the carried state is a halo of
        # less than two blocks); call numbers are taken relative to the pass.
        passes_before = e_warm + e_steps - 1
        e_mis = e_rec = e_streams = 0
        for w, n in enumerate(wave_n):
            ids = sorted({i for i in (0, 1, 2, n // 2, n // 2 + 1, n - 3, n - 2, n - 1) if 0 <= i < n})
            rows = ring[w % 2][ids]
            two = np.concatenate([rows, rows], axis=1) if passes_before else rows
            want_w = oracle_records(two, ids, call0=a.blocks if passes_before else 0)
            got_w = last_views[w][np.isin(last_views[w]["stream"], ids)].copy()
            got_w["call"] -= passes_before * a.blocks
            g_rows, w_rows = _rows(got_w), _rows(want_w)
            e_mis += 0 if g_rows == w_rows else max(1, len(set(g_rows) ^ set(w_rows)))
            e_rec += len(w_rows)
            e_streams += len(ids)
        e2e_parity = {"streams": all_sum(e_streams), "records": all_sum(e_rec), "mismatch": all_sum(e_mis),
                      "how": "records of the last timed end-to-end step (pinned host input through the host "
                             "pipeline) for 8 streams per wave per rank -- first, middle, last of the rank's share "
                             "-- against oracle/davis_oracle.c on the same bytes"}
        e2e_samples = n_global * n_samples
        e2e = {"value": e2e_samples * e_steps / dt / 1e6, "unit": UNIT,
               "h2d_bytes_per_step": all_sum(h2d), "d2h_bytes_per_step": all_sum(d2h),
               "ms_per_step": dt / e_steps * 1e3, "steps": e_steps,
               "valid_packets_last_step": all_sum(valid),
               "streams_this_rank": mine_n, "waves": len(wave_n),
               "split": "uniform" if weights is None else "weighted by the probed host->device rate of every rank",
               "h2d_probe_gbs": probe, "parity_check": e2e_parity}
        if world > 1:
            counts = [torch.zeros(1, dtype=torch.int64, device="cuda") for _ in range(world)]
            dist.all_gather(counts, torch.tensor([mine_n], dtype=torch.int64, device="cuda"))
            e2e["streams_per_rank"] = [int(c.item()) for c in counts]
        for bd2 in handles:
            bd2.close()

    # ---- CPU baseline on rank 0, N = 1 only ----
    cpu = None
    if rank == 0 and world == 1 and not a.no_cpu:
        cores = usable_cores()
        kind = cpu_kind()
        n_distinct = min(a.streams, SAMPLE_STREAMS)
        sample_blocks = min(a.blocks, SAMPLE_BLOCKS)
        sample = dev[:n_distinct, : 2 * BLOCK * sample_blocks].cpu().numpy()
        rate, samples, wall, _ = cpu_reference_rate(sample, a.cpu_seconds, cores)
        c_rate = c_port_rate(dev[: min(a.streams, 2 * cores), :].cpu().numpy(), cores)
        cpu = {"value": rate, "unit": UNIT, "cores": cores, "kind": kind,
               "sample": cpu_sample_text(cores, a.cpu_seconds, kind),
               "samples_timed": samples, "wall_s": wall,
               "c_port_value": c_rate,
               "c_port_note": f"oracle/davis_oracle.c, {cores} threads, whole captures (not the reference's speed; a stronger CPU bar)"}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_dict(a),
            "notes": {"pipelining": "two sharded handles alternate: batch i is submitted before batch i-1 is fetched "
                                    "(its record copy overlaps the next batch's kernels); k_demod of both runs on one "
                                    "stream, the second search/finish chain and the record copies on the handles' own streams",
                      "decision_arithmetic": "fp32 streaming kernel; every bit of every reported packet re-derived with an "
                                             "error bound (fp32 outside the guard band, fp64 inside), RSSI/SNR/discriminator mean in fp64",
                      "sharding": f"{n_global} streams over {world} rank(s) through rtldavis_b200.shard, no data-path collective"},
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "sustained": sustained,
            "parity_check": parity,
            "two_streams": {"value": samples_step / (two_streams_ms * 1e-3) / 1e6, "unit": UNIT,
                            "ms_per_step": two_streams_ms,
                            "note": "same loop, one stream per handle (kernels of consecutive batches overlap); "
                                    "host wall clock; not used for value/roofline"},
            "gpu_launches": int(launches), "clocks": clocks,
            "valid_packets_per_step": n_valid,
            "embedded_packets_per_step": a.streams * a.packets * world,
        }
        emit(line)
    for sd in sds:
        sd.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


_JSON_FD = None


def emit(line: dict) -> None:
    """The ONE stdout line of the contract."""
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def main():
    global _JSON_FD
    a = parse_args()
    # Libraries write to fd 1 behind Python's back (NCCL prints "NCCL version ..." there on
    # rank 0 when the box sets NCCL_DEBUG): everything but the JSON line goes to stderr.
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    if a.impl == "reference":
        return run_reference(a)
    return run_graft(a)


if __name__ == "__main__":
    sys.exit(main())
