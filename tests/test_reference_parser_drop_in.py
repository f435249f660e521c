"""The reference's own caller on top of the GPU demodulator.

``worker.worker_main`` (src/rtldavis/worker.py:10-58) builds a ``protocol.Parser``
(protocol.py:115-116 constructs its ``dsp.Demodulator``), calls
``p.demodulator.demodulate(samples)`` and ``p.parse(packets)`` per block (worker.py:49-50), and
``Parser.parse`` reads ``self.demodulator.discriminated[index:index+224]`` (protocol.py:304-311).
These tests run that UNMODIFIED code (byte-compiled into oracle/_ref by oracle/stage_ref.py)
twice over BASELINE config 1 -- once as it is, once with ``rtldavis_b200.dsp`` in place of
``rtldavis.dsp`` as INTEGRATION.md section 2 describes -- and compare the ``Message`` lists.
"""
import math
import queue
import random

import numpy as np
import pytest

from conftest import load_seeded
from oracle import ref_loader

pytestmark = pytest.mark.gpu

B = 8192


def _msg_key(m):
    vals = {k: v for k, v in m.sensor_values.items() if k not in ("rssi", "snr")}
    return (m.id, m.sensor_type, m.raw_sensor_id, m.raw_msg_type3, int(m.packet.index),
            bytes(np.asarray(m.packet.data, dtype=np.uint8).tolist()), repr(sorted(vals.items())))


def _assert_same_messages(got, want):
    assert [(k, _msg_key(m)) for k, m in got] == [(k, _msg_key(m)) for k, m in want]
    for (_, a), (_, b) in zip(got, want):
        for key in ("rssi", "snr"):                     # fp64 both sides, summation order: 1e-9 dB
            assert math.isclose(a.sensor_values[key], b.sensor_values[key], rel_tol=0, abs_tol=1e-6)
        assert math.isclose(a.packet.rssi, b.packet.rssi, rel_tol=0, abs_tol=1e-9)
        assert math.isclose(a.packet.snr, b.packet.snr, rel_tol=0, abs_tol=1e-9)


def _parser(ref_protocol):
    random.seed(20261105)            # Parser.__post_init__ draws its first hop index (protocol.py:226)
    return ref_protocol.Parser(symbol_length=14)


def _reference_run(ref_protocol, iq, n_blocks):
    p = _parser(ref_protocol)
    out = []
    for k in range(n_blocks):
        for m in p.parse(p.demodulator.demodulate(iq[2 * B * k:2 * B * (k + 1)])):
            out.append((k, m))
    return p, out


@pytest.fixture(scope="module")
def ref():
    if not ref_loader.available():
        pytest.skip("oracle/_ref not staged (python oracle/stage_ref.py in the build container)")
    return ref_loader.load()


def test_unmodified_parser_with_the_gpu_demodulator(ref):
    """parser.demodulator swapped for the drop-in; everything else is the reference."""
    from rtldavis_b200 import dsp as gpu_dsp
    from rtldavis_b200 import protocol as gpu_protocol
    _, _, ref_protocol = ref
    g, iq = load_seeded("config1_10s")
    n_blocks = int(g["n_blocks"])
    p_ref, want = _reference_run(ref_protocol, iq, n_blocks)
    assert len(want) == 40                              # BASELINE config 1: 40 packets, station 0

    p = _parser(ref_protocol)
    p.demodulator = gpu_dsp.Demodulator(gpu_protocol.new_packet_config(p.symbol_length))
    got = []
    for k in range(n_blocks):
        for m in p.parse(p.demodulator.demodulate(iq[2 * B * k:2 * B * (k + 1)])):
            got.append((k, m))
    p.demodulator.close()
    _assert_same_messages(got, want)
    # the parser's own state after the run: frequency errors from demodulator.discriminated
    assert p.transmitter == p_ref.transmitter
    for tr, per_ch in p_ref.freq_err_tr_ch_list.items():
        for ch, errs in per_ch.items():
            assert p.freq_err_tr_ch_list[tr][ch] == errs
    assert any(any(e != 0 for e in errs) for per_ch in p_ref.freq_err_tr_ch_list.values()
               for errs in per_ch.values())


def test_unmodified_worker_loop_with_the_one_line_import_swap(ref, monkeypatch):
    """`from rtldavis_b200 import dsp` instead of `from . import dsp`
    in protocol.py, nothing else.  worker_main runs unmodified on in-process queues."""
    import importlib

    from rtldavis_b200 import dsp as gpu_dsp
    _, _, ref_protocol = ref
    worker = importlib.import_module("rtldavis.worker")
    g, iq = load_seeded("config1_10s")
    n_blocks = int(g["n_blocks"])
    _, want = _reference_run(ref_protocol, iq, n_blocks)

    monkeypatch.setattr(ref_protocol, "dsp", gpu_dsp)    # the import swap
    data_q, result_q = queue.Queue(), queue.Queue()
    for k in range(n_blocks):
        data_q.put(iq[2 * B * k:2 * B * (k + 1)])
    data_q.put(None)
    worker.worker_main(data_q, result_q, None, 14, 30)
    got = []
    while not result_q.empty():
        got.append(result_q.get())
    assert isinstance(got[0].packet, gpu_dsp.Packet)
    _assert_same_messages([(0, m) for m in got], [(0, m) for _, m in want])
