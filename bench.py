#!/usr/bin/env python3
"""bench.py -- IQ Msample/s demodulated on B200, beside the reference's CPU path.

One "step" = one pass of the whole hot path (demodulate -> sync search -> fp64
confirm/slice/CRC -> carry) over one batch of synthetic captures.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
                    [--streams S] [--blocks NB] [--snr DB]

N = 1 workload: BASELINE.json configs[1] -- 1,024 independent 10 s captures
(328 blocks of 8192 samples each), mixed station IDs, +-50 ppm clock error,
+-3 kHz residual carrier.  N > 1 (torchrun): the same 1,024 streams on every
rank, sharded by stream, no data-path collective ("weak" scaling).

value   device-resident throughput (inputs already in HBM), CUDA-event timed
e2e     same metric through BatchDemodulator.submit/fetch with PINNED HOST input:
        H2D of every byte and D2H of the packet records inside the timed region
roofline  the demodulation kernel alone, 2 algorithmic bytes per complex sample
cpu_baseline  the reference algorithm on this box's host cores (oracle port)
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "iq_msamples_per_s_demodulated"
UNIT = "Msample/s"
BLOCK = 8192


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="graft", choices=["graft", "reference"])
    ap.add_argument("--streams", type=int, default=1024, help="streams per GPU")
    ap.add_argument("--blocks", type=int, default=328, help="8192-sample blocks per stream (328 = 10 s)")
    ap.add_argument("--snr", type=float, default=20.0)
    ap.add_argument("--packets", type=int, default=40)
    ap.add_argument("--seed", type=int, default=2024)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    return ap.parse_args()


def workload_name(a):
    return (f"{a.streams} streams x {a.blocks} blocks x {BLOCK} samples "
            f"({a.blocks * BLOCK / 268800.0:.2f} s @268.8 kS/s) per GPU, {a.packets} packets/stream, "
            f"stations 0-7, SNR {a.snr:g} dB, df +-3 kHz, clock +-50 ppm")


# ---------------------------------------------------------------------------
# CPU baseline: the oracle's literal restatement of the reference, one process
# per core, on a bounded sample of the same workload.
# ---------------------------------------------------------------------------
def _cpu_worker(job):
    iq, seconds, literal = job
    from oracle import davis_oracle as O
    o = O.CallOracle(O.OracleConfig(), literal_quantize=literal)
    step = 2 * BLOCK
    n_blocks = iq.size // step
    done, packets = 0, 0
    t0 = time.perf_counter()
    while True:
        k = done % n_blocks
        if k == 0 and done:
            o.reset()
        packets += len(o.demodulate(iq[k * step:(k + 1) * step]))
        done += 1
        if time.perf_counter() - t0 >= seconds:
            break
    return done * BLOCK, time.perf_counter() - t0, packets


def cpu_reference_rate(sample_iq: np.ndarray, seconds: float, cores: int, literal: bool = True):
    """Aggregate Msample/s of `cores` processes each demodulating its own stream
    of the sample for ~`seconds` (block by block, as the reference is driven)."""
    import multiprocessing as mp
    ctx = mp.get_context("fork")
    jobs = [(sample_iq[i % sample_iq.shape[0]], seconds, literal) for i in range(cores)]
    t0 = time.perf_counter()
    with ctx.Pool(cores) as pool:
        res = pool.map(_cpu_worker, jobs)
    wall = time.perf_counter() - t0
    samples = sum(r[0] for r in res)
    busy = max(r[1] for r in res)
    return samples / busy / 1e6, samples, wall


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
    power and throttle reasons every ~0.5 ms (the timed region is tens of ms, too
    short for `nvidia-smi -lms`)."""
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
# reference arm: the reference's CPU algorithm (oracle port) on all host cores
# ---------------------------------------------------------------------------
def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from rtldavis_b200 import synth
    cores = usable_cores()
    # bounded sample of the same workload: a handful of distinct streams (numpy
    # generator, same signal model), each process loops over its stream
    n_distinct = min(cores, 8)
    sample_blocks = min(a.blocks, 41)
    iq, _ = synth.make_batch(n_distinct, sample_blocks * BLOCK,
                             max(1, a.packets * sample_blocks // 328), a.snr, seed0=a.seed,
                             df_range_hz=3000.0, ppm_range=50.0)
    per_step = max(1.0, min(a.cpu_seconds, 150.0 / max(1, a.steps + a.warmup)))  # whole run <= ~2.5 min
    rates = []
    for i in range(a.warmup + a.steps):
        rate, samples, wall = cpu_reference_rate(iq, per_step, cores, literal=True)
        if i >= a.warmup:
            rates.append((rate, samples, wall))
    value = float(np.mean([r[0] for r in rates]))
    samples = int(np.mean([r[1] for r in rates]))
    sample_txt = (f"{cores} processes x ~{per_step:.1f} s each looping over a {sample_blocks}-block "
                  f"capture of the workload, block by block (unmodified algorithm incl. per-sample quantize)")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": a.gpus,
        "steps": a.steps, "warmup": a.warmup,
        "ms_per_step": float(np.mean([r[2] for r in rates])) * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": workload_name(a), "sample": sample_txt},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": sample_txt, "samples_per_step": samples},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)
    return 0


# ---------------------------------------------------------------------------
# graft arm
# ---------------------------------------------------------------------------
def run_graft(a):
    import ctypes as C

    import torch
    import torch.distributed as dist

    from rtldavis_b200 import _lib, dsp, protocol

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

    L = _lib.load()
    n_samples = a.blocks * BLOCK
    row = 2 * n_samples
    dev = torch.empty((a.streams, row), dtype=torch.uint8, device="cuda")
    seed = a.seed + 7919 * rank                         # every rank gets its own streams
    _lib.check(L.davis_synth_fill(dev.data_ptr(), a.streams, row, n_samples, seed, a.packets,
                                  a.snr, 3000.0, 50.0, local, None))
    torch.cuda.synchronize()

    cfg = protocol.new_packet_config(14)
    bd = dsp.BatchDemodulator(cfg, a.streams, device=local, max_blocks=a.blocks)
    # a non-default stream: the library takes a NULL stream argument as "use the handle's own"
    stream = torch.cuda.Stream(device=local)
    sptr = stream.cuda_stream
    assert sptr != 0

    # Two handles alternate (each carries its own copy of the per-stream state): batch i is
    # submitted before batch i-1 is fetched, so the device->host copy of one batch's packet
    # records (the handle's own copy stream) and the host-side bookkeeping overlap the next
    # batch's kernels.  Every kernel of both handles runs on ONE stream, in order, so the
    # per-kernel event timers are those of kernels running alone.
    bd_b = dsp.BatchDemodulator(cfg, a.streams, device=local, max_blocks=a.blocks)
    bds = (bd, bd_b)

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
    ms_total = e0.elapsed_time(e1)

    # Extra, not the headline: the same loop with each handle on its own stream, so the
    # latency-bound finish kernels of one batch run beside the FP32-bound k_demod of the
    # other.  Faster as a whole, but the per-kernel timers then measure kernels that share
    # the GPU, so `value` and `roofline` above come from the single-stream run.
    run_steps(2, lambda st: None, on_stream=None)
    barrier()
    t0 = time.perf_counter()
    run_steps(a.steps, lambda st: None, on_stream=None)
    torch.cuda.synchronize()
    dt2 = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(dt2, op=dist.ReduceOp.MAX)
    two_streams_ms = float(dt2.item()) / a.steps * 1e3
    t = torch.tensor([ms_total], dtype=torch.float64, device="cuda")
    n_valid = torch.tensor([int(hits["crc_ok"].sum())], dtype=torch.int64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(n_valid, op=dist.ReduceOp.SUM)   # host-side style gather of a count
    ms_step = float(t.item()) / a.steps
    samples_step = a.streams * n_samples * world
    value = samples_step / (ms_step * 1e-3) / 1e6

    # ---- roofline of the demodulation kernel (CUDA events inside the library,
    #      on the launching stream, around that kernel only) ----
    peak, peak_src = measured_peak()
    demod_avg = float(np.mean(demod_ms))
    ach = 2.0 * a.streams * n_samples / (demod_avg * 1e-3) / 1e9
    roofline = {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                "traffic": None, "peak_source": peak_src, "kernel": "k_demod",
                "kernel_ms": demod_avg, "search_ms": float(np.mean(search_ms)),
                "finish_ms": float(np.mean(finish_ms)),
                "frac_of_nominal_8TBs": ach / 8000.0,
                "algorithmic_bytes_per_launch": 2 * a.streams * n_samples}
    prof = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(prof):
        try:
            with open(prof) as f:
                cap = json.load(f)
            # the ncu capture is of the default workload; any other size has no capture
            if cap.get("algorithmic_bytes_per_launch") == roofline["algorithmic_bytes_per_launch"]:
                roofline["traffic"] = cap.get("k_demod_dram_bytes_per_launch")
        except Exception:
            pass

    # ---- end to end: pinned host input, H2D + records D2H inside the timed region ----
    e2e = None
    if not a.no_e2e:
        host = torch.empty((a.streams, row), dtype=torch.uint8, pin_memory=True)
        host.copy_(dev)
        torch.cuda.synchronize()
        host_np = host.numpy()
        bd2 = dsp.BatchDemodulator(cfg, a.streams, device=local, max_blocks=a.blocks)
        for _ in range(max(1, min(a.warmup, 2))):
            bd2.submit(host_np)
            h2 = bd2.fetch(copy=False)
        barrier()
        t0 = time.perf_counter()
        h2d = d2h = 0
        for _ in range(a.steps):
            bd2.submit(host_np)
            h2 = bd2.fetch(copy=False)
            st2 = bd2.stats()
            h2d, d2h = int(st2.h2d_bytes), int(st2.d2h_bytes)
        torch.cuda.synchronize()
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        e2e_val = samples_step * a.steps / float(dt.item()) / 1e6
        e2e = {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": h2d * world,
               "d2h_bytes_per_step": d2h * world,
               "ms_per_step": float(dt.item()) / a.steps * 1e3,
               "valid_packets_last_step": int(h2["crc_ok"].sum())}
        bd2.close()

    # ---- CPU baseline on rank 0, N = 1 only ----
    cpu = None
    if rank == 0 and world == 1 and not a.no_cpu:
        cores = usable_cores()
        n_distinct = min(cores, a.streams, 8)
        sample_blocks = min(a.blocks, 41)
        sample = dev[:n_distinct, : 2 * BLOCK * sample_blocks].cpu().numpy()
        rate, samples, wall = cpu_reference_rate(sample, a.cpu_seconds, cores, literal=True)
        c_rate = c_port_rate(dev[: min(a.streams, 2 * cores), :].cpu().numpy(), cores)
        cpu = {"value": rate, "unit": UNIT, "cores": cores, "kind": "port",
               "sample": (f"first {n_distinct} streams x {sample_blocks} blocks of this workload; "
                          f"{cores} processes x ~{a.cpu_seconds:g} s, block by block through the "
                          f"oracle's literal restatement of dsp.Demodulator (per-sample quantize kept)"),
               "samples_timed": samples, "wall_s": wall,
               "c_port_value": c_rate,
               "c_port_note": f"oracle/davis_oracle.c, {cores} threads, whole captures (not the reference's speed; a stronger CPU bar)"}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(a), "streams_per_gpu": a.streams,
                       "blocks_per_stream": a.blocks, "block_size": BLOCK,
                       "l2": "inputs (%.2f GB per GPU) far exceed the 126 MB L2" % (a.streams * row / 1e9),
                       "pipelining": "two handles alternate: batch i is submitted before batch i-1 is fetched "
                                     "(its record copy overlaps the next batch's kernels); all kernels on one stream",
                       "decision_arithmetic": "fp32 streaming kernel; every bit of every reported packet re-derived with an error bound (fp32 outside the guard band, fp64 inside), RSSI/SNR/discriminator mean in fp64"},
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e,
            "two_streams": {"value": samples_step / (two_streams_ms * 1e-3) / 1e6, "unit": UNIT,
                            "ms_per_step": two_streams_ms,
                            "note": "same loop, one stream per handle (kernels of consecutive batches overlap); "
                                    "host wall clock; not used for value/roofline"},
            "gpu_launches": int(launches), "clocks": clocks,
            "valid_packets_per_step": int(n_valid.item()),
            "embedded_packets_per_step": a.streams * a.packets * world,
        }
        emit(line)
    bd.close()
    bd_b.close()
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
