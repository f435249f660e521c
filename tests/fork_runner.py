"""rtldavis's own launch sequence over the drop-in, in a fresh interpreter.

Started by tests/test_fork_boundary.py with ``python tests/fork_runner.py u8|c128``.  It does
what ``runners/rtlsdr.py`` does (src/rtldavis/runners/rtlsdr.py:30,58-65; start method "fork",
__main__.py:277), with INTEGRATION.md section 2's import swap in place:

  1. the parent builds a ``protocol.Parser`` (which builds a ``dsp.Demodulator``) and reads
     ``p.cfg`` and a hop from it;
  2. it forks ``worker.worker_main`` (worker.py:10-58) as a ``multiprocessing.Process`` with real
     ``multiprocessing.Queue``s; the worker builds its own Parser and demodulates;
  3. the parent feeds BASELINE config 1's 328 blocks -- uint8, or complex128 (the type pyrtlsdr's
     read_samples hands the worker, dsp.py:144-151) -- and collects the ``Message``s.

A CUDA context in the parent before the fork would break every CUDA call of the child; the
drop-in's Demodulator therefore creates its handle at the first call.  The Messages must equal
those of the all-reference run on the same input, and the parent must never have loaded the
CUDA driver.  Exit code 0 = all of that holds.
"""
import importlib
import multiprocessing
import os
import queue
import random
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (ROOT, HERE):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402

B = 8192


def driver_loaded() -> bool:
    with open("/proc/self/maps") as fh:
        return "libcuda.so" in fh.read()


def main(mode: str) -> int:
    from conftest import load_seeded
    from oracle import ref_loader
    from test_reference_parser_drop_in import _assert_same_messages
    _, _, ref_protocol = ref_loader.load()
    worker = importlib.import_module("rtldavis.worker")
    g, iq = load_seeded("config1_10s")
    n_blocks = int(g["n_blocks"])
    if mode == "c128":
        u = iq.astype(np.float64)
        samples = ((u[0::2] - 127.4) / 127.6) + 1j * ((u[1::2] - 127.4) / 127.6)   # dsp.py:26
        blocks = [samples[B * k:B * (k + 1)] for k in range(n_blocks)]
        assert blocks[0].dtype == np.complex128
    else:
        blocks = [iq[2 * B * k:2 * B * (k + 1)] for k in range(n_blocks)]

    # the all-reference answer (CPU), before anything is swapped
    random.seed(20261105)
    p_ref = ref_protocol.Parser(symbol_length=14)
    want = [m for blk in blocks for m in p_ref.parse(p_ref.demodulator.demodulate(blk))]
    assert len(want) == 40, len(want)

    from rtldavis_b200 import dsp as gpu_dsp
    ref_protocol.dsp = gpu_dsp                                   # the import swap
    multiprocessing.set_start_method("fork")                     # __main__.py:277
    p = ref_protocol.Parser(symbol_length=14)                    # runners/rtlsdr.py:30
    assert isinstance(p.demodulator, gpu_dsp.Demodulator)
    assert p.cfg.sample_rate == 268800
    p.rand_hop()
    assert not driver_loaded(), "the parent touched CUDA before the fork"

    data_q, result_q = multiprocessing.Queue(), multiprocessing.Queue()
    proc = multiprocessing.Process(target=worker.worker_main, args=(data_q, result_q, None, 14, 30))
    proc.start()
    for blk in blocks:
        data_q.put(blk)
    data_q.put(None)
    got = []
    while True:
        try:
            got.append(result_q.get(timeout=0.5))
        except queue.Empty:
            if not proc.is_alive():
                break
    while True:                                                  # what arrived with the exit
        try:
            got.append(result_q.get(timeout=0.5))
        except queue.Empty:
            break
    proc.join(30)
    assert proc.exitcode == 0, proc.exitcode
    assert not driver_loaded(), "the parent loaded the CUDA driver"
    print(f"{mode}: {len(got)} messages from the forked worker, {len(want)} from the reference")
    assert got and isinstance(got[0].packet, gpu_dsp.Packet)
    _assert_same_messages([(0, m) for m in got], [(0, m) for m in want])
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv[1] if len(sys.argv) > 1 else "u8"))
