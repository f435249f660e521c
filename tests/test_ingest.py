"""Fan-in of several rtl_sdr-format byte sources into the batched demodulator
(worker.py:34-58, runners/rtlsdr.py:100-103) and the batch form of the
Parser.parse glue (protocol.py:282-332)."""
import io
import math

import numpy as np
import pytest

from conftest import load_golden
from rtldavis_b200 import ingest


class _Cfg:
    block_size = 16


class _Recorder:
    """Stands in for BatchDemodulator on the CPU: keeps what was submitted."""

    def __init__(self, n_streams, max_blocks=8):
        self.n_streams, self.max_blocks, self.cfg = n_streams, max_blocks, _Cfg()
        self.batches = []

    def submit(self, iq):
        self.batches.append(np.array(iq, copy=True))

    def fetch(self):
        return len(self.batches)


class _Dribble:
    """A pipe that returns at most 7 bytes per read()."""

    def __init__(self, data):
        self.f = io.BytesIO(data)

    def read(self, n):
        return self.f.read(min(n, 7))


def _bytes(n, seed):
    return np.random.default_rng(seed).integers(0, 256, n, dtype=np.uint8).tobytes()


def test_read_exact_and_iter_blocks_short_reads_and_tail():
    data = _bytes(32 * 5 + 9, 1)            # 5 whole blocks of 16 samples + a tail
    assert ingest.read_exact(_Dribble(data), 50) == data[:50]
    chunks = list(ingest.iter_blocks(_Dribble(data), 16, n_blocks=2))
    assert [c.size for c in chunks] == [64, 64, 32]      # the 9-byte tail is dropped
    assert b"".join(c.tobytes() for c in chunks) == data[:160]
    assert list(ingest.iter_blocks(io.BytesIO(b""), 16)) == []


@pytest.mark.parametrize("prefetch", [False, True])
def test_fan_in_ragged_sources_keep_stream_index_and_order(prefetch):
    step = 32
    a, b, c = _bytes(step * 7, 2), _bytes(step * 3 + 5, 3), _bytes(step * 5, 4)
    rec = _Recorder(3)
    fan = ingest.FanIn(rec, [io.BytesIO(a), _Dribble(b), io.BytesIO(c)], n_blocks=2, prefetch=prefetch)
    assert list(fan.run()) == list(range(1, len(rec.batches) + 1))
    fan.close()
    per_stream = [b"".join(x[i].tobytes() for x in rec.batches) for i in range(3)]
    for got, src, blocks in zip(per_stream, (a, b, c), (7, 3, 5)):
        assert got[: blocks * step] == src[: blocks * step]          # every whole block, in order
        assert set(got[blocks * step:]) <= {ingest.FanIn.IDLE}      # then "no signal"
    assert fan.blocks_read == [7, 3, 5]
    assert all(x.shape[1] % step == 0 for x in rec.batches)


def test_fan_in_argument_checks():
    with pytest.raises(ValueError):
        ingest.FanIn(_Recorder(2), [io.BytesIO(b"")], n_blocks=1)
    with pytest.raises(ValueError):
        ingest.FanIn(_Recorder(1, max_blocks=4), [io.BytesIO(b"")], n_blocks=5)


@pytest.mark.gpu
@pytest.mark.parametrize("prefetch", [False, True])
def test_fan_in_files_equal_direct_submit_and_parse(tmp_path, prefetch):
    from rtldavis_b200 import dsp, protocol
    names = ["hi_snr_8192", "lo_snr_8192"]
    caps = [np.asarray(load_golden(n)["iq"], dtype=np.uint8).ravel() for n in names]
    n_blocks = min(c.size for c in caps) // 16384
    paths = []
    for n, c in zip(names, caps):
        p = tmp_path / (n + ".u8")
        c[: n_blocks * 16384].tofile(p)     # equal lengths: no stream is fed idle blocks
        paths.append(str(p))
    cfg = protocol.new_packet_config(14)
    with dsp.BatchDemodulator(cfg, 2, max_blocks=n_blocks) as direct:
        direct.submit(np.stack([c[: n_blocks * 16384] for c in caps]))
        want = direct.fetch()
    with dsp.BatchDemodulator(cfg, 2, max_blocks=8) as bd:
        fan = ingest.FanIn(bd, paths, n_blocks=5, prefetch=prefetch)
        got = np.concatenate(list(fan.run()))
        fan.close()
    key = lambda h: (int(h["stream"]), int(h["call"]), int(h["index"]), h["data"].tobytes(),
                     bool(h["crc_ok"]))
    assert sorted(map(key, got)) == sorted(map(key, want))
    parsed = protocol.parse_hits(want, cfg)
    valid = [h for h in want if h["crc_ok"]]
    assert len(parsed) == len(valid) > 0
    for p, h in zip(parsed, valid):
        assert p["data"] == bytes(protocol.swap_bit_order(x) for x in h["data"])[2:]
        assert p["id"] == p["data"][0] & 7
        assert p["freq_err"] == -int(float(h["disc_mean"]) * 268800 / (2 * math.pi))
    # the discriminator mean is the reference's mean(discriminated[index:index+224]),
    # i.e. D over the 224 samples from the preamble's global position (protocol.py:304-310)
    from oracle import davis_oracle as O
    disc = [O.stream_discriminated(c[: n_blocks * 16384]) for c in caps]
    for h in valid:
        g = int(h["position"])
        np.testing.assert_allclose(float(h["disc_mean"]), disc[int(h["stream"])][g:g + 224].mean(),
                                   rtol=1e-9, atol=1e-13)
    assert protocol.parse_hits(want, cfg, station_id=5) == [p for p in parsed if p["id"] == 5]
