"""BASELINE.json configs 1-5 at the sizes they state, every record against the oracle.

All through the C ABI.  Config 1 goes through the drop-in ``Demodulator`` block by block
(328 calls) against fixtures the unmodified reference produced (oracle/gen_golden.py);
configs 2-5 are generated on the device and compared record by record -- stream, call, index,
10 payload bytes, CRC verdict, order -- with the C oracle run on a host copy of the same bytes
(the oracle itself is pinned on whole 10 s captures by tests/test_oracle_golden.py).
"""
import numpy as np
import pytest

from conftest import TEN_SECOND_FIXTURES, golden_rows, load_seeded
from oracle import c_oracle as CO

pytestmark = pytest.mark.gpu

B = 8192
N_BLOCKS = 328                       # 10 s at 268.8 kS/s: 2 686 976 of 2 688 000 samples


@pytest.fixture(scope="module")
def dsp():
    from rtldavis_b200 import dsp as d
    return d


@pytest.fixture(scope="module")
def protocol():
    from rtldavis_b200 import protocol as p
    return p


def _rows(rec):
    return [(int(r["stream"]), int(r["call"]), int(r["index"]), r["data"].tobytes(), bool(r["crc_ok"]))
            for r in rec]


def _device_batch(n_streams, n_blocks, seed, snr, n_packets=40, df=3000.0, ppm=50.0):
    """uint8[n_streams, 2 * n_samples] on the device + the messages the generator embedded."""
    import torch
    from rtldavis_b200 import _lib
    L = _lib.load()
    n_samples = n_blocks * B
    buf = torch.empty((n_streams, 2 * n_samples), dtype=torch.uint8, device="cuda")
    _lib.check(L.davis_synth_fill(buf.data_ptr(), n_streams, 2 * n_samples, n_samples, seed, n_packets,
                                  float(snr), float(df), float(ppm), 0, None))
    torch.cuda.synchronize()
    msgs = np.zeros((n_streams, n_packets, 8), dtype=np.uint8)
    starts = np.zeros((n_streams, n_packets), dtype=np.int64)
    _lib.check(L.davis_synth_truth(n_streams, n_samples, seed, n_packets, float(snr),
                                   msgs.ctypes.data, starts.ctypes.data))
    return buf, msgs


def _need_gb(gb):
    import torch
    free, _ = torch.cuda.mem_get_info()
    if free < gb * 2 ** 30:
        pytest.skip(f"needs {gb} GB of free device memory")


# ---------------------------------------------------------------------------
# config 1: one 10 s capture through Demodulator.demodulate, block by block
# ---------------------------------------------------------------------------
@pytest.mark.parametrize("name", TEN_SECOND_FIXTURES)
def test_config1_10s_capture_through_the_drop_in(dsp, protocol, name):
    """328 calls of the drop-in Demodulator (worker.py:49) on the capture the reference was run
    on: every dsp.Packet (call, index, 10 bytes), the CRC verdict of Parser.parse, RSSI and SNR."""
    g, iq = load_seeded(name)
    dem = dsp.Demodulator(protocol.new_packet_config(14))
    got, rssi, snr = [], [], []
    for k in range(int(g["n_blocks"])):
        pk = dem.demodulate(iq[2 * B * k:2 * B * (k + 1)])
        for p, h in zip(pk, dem.last_hits):
            got.append((k, p.index, p.data.tobytes(), bool(h["crc_ok"])))
            rssi.append(p.rssi)
            snr.append(p.snr)
    dem.close()
    assert got == golden_rows(g)
    np.testing.assert_allclose(rssi, g["pk_rssi"], rtol=0, atol=1e-9)      # dB, summation order only
    np.testing.assert_allclose(snr, g["pk_snr"], rtol=0, atol=1e-9)
    if name == "config1_10s":
        assert sum(r[3] for r in got) >= 40 and len({r[2] for r in got if r[3]}) == 40


def test_config1_10s_capture_as_one_batched_call_and_streaming_bits(dsp, protocol):
    """The same capture as ONE call of 328 blocks, and the streaming kernel's decision bits
    against the reference's quantized stream (gate ii: they may differ only next to a zero of the
    discriminator; counted)."""
    g, iq = load_seeded("snr6_10s")
    bd = dsp.BatchDemodulator(protocol.new_packet_config(14), 1, max_blocks=N_BLOCKS)
    bd.submit(iq.reshape(1, -1))
    hits = bd.fetch()
    assert [r[1:] for r in _rows(hits)] == golden_rows(g)
    first, bits = bd.debug_bits(0)
    want = np.unpackbits(g["bits"])[: int(g["n_bits"])]
    lo = max(first, 16)
    got = bits[lo - first: lo - first + want.size - lo]
    bad = int(np.sum(got != want[lo:lo + got.size]))
    print(f"streaming decisions differing from the reference's quantized bits: {bad} of {got.size}")
    assert got.size > 2_600_000 and bad <= 1e-5 * got.size
    bd.close()


# ---------------------------------------------------------------------------
# config 2: 1 024 x 10 s on one GPU (the bench workload)
# ---------------------------------------------------------------------------
def test_config2_1024_streams_10s_every_record(dsp, protocol):
    _need_gb(14)
    n_streams = 1024
    buf, msgs = _device_batch(n_streams, N_BLOCKS, 777, 20.0)
    with dsp.BatchDemodulator(protocol.new_packet_config(14), n_streams, max_blocks=N_BLOCKS) as bd:
        bd.submit(buf)
        hits = bd.fetch()
    host = buf.cpu().numpy()
    del buf
    want = CO.batch_records(host)
    assert want.shape[0] > 100_000
    assert _rows(hits) == _rows(want)
    valid = {(int(h["stream"]), bytes(protocol.swap_bit_order(b) for b in h["data"][2:]))
             for h in hits[hits["crc_ok"] != 0]}
    back = sum((s, msgs[s, k].tobytes()) in valid for s in range(n_streams) for k in range(40))
    print(f"config 2: {want.shape[0]} records equal to the oracle's; {back} of {n_streams * 40} embedded recovered")
    assert back >= 0.995 * n_streams * 40


def test_config2_1024_streams_10s_from_pinned_host_memory(dsp, protocol):
    """The end-to-end leg's path at full size: the same 1 024 x 10 s batch as
    pinned HOST memory, so it goes through the handle's pipeline stages (groups of ~24 streams
    cycling through three staging buffers) instead of being used in place.  Every record against
    the C oracle on the same bytes, and against the device-resident call."""
    import torch
    _need_gb(14)
    n_streams = 1024
    buf, _ = _device_batch(n_streams, N_BLOCKS, 778, 20.0)
    cfg = protocol.new_packet_config(14)
    with dsp.BatchDemodulator(cfg, n_streams, max_blocks=N_BLOCKS) as bd:
        bd.submit(buf)
        on_device = bd.fetch()
    pinned = torch.empty(buf.shape, dtype=torch.uint8, pin_memory=True)
    pinned.copy_(buf)
    torch.cuda.synchronize()
    del buf
    host = pinned.numpy()
    with dsp.BatchDemodulator(cfg, n_streams, max_blocks=N_BLOCKS) as bd:
        bd.submit(host)
        hits = bd.fetch()
        st = bd.stats()
    assert st.h2d_bytes == host.size                       # every byte crossed the host link, once
    want = CO.batch_records(host)
    assert want.shape[0] > 100_000 and hits.shape[0] == want.shape[0]
    for k in ("stream", "call", "index", "crc_ok"):
        assert np.array_equal(hits[k], want[k]), k
    assert np.array_equal(hits["data"], want["data"])
    for k in ("stream", "call", "index", "data", "crc_ok", "position", "rssi", "snr"):
        assert np.array_equal(hits[k], on_device[k]), k      # the two input paths agree to the last bit
    print(f"config 2 from pinned host memory: {want.shape[0]} records equal to the oracle's and to the "
          f"device-resident call's")


# ---------------------------------------------------------------------------
# config 3: the 51-channel hop set, 10 s, block by block and in ragged chunks
# ---------------------------------------------------------------------------
def test_config3_hop_set_51_channels_10s(dsp, protocol):
    from rtldavis_b200 import synth
    hop = int(round(2.5625 * 268800))              # dwell time, protocol.py:118
    iq, truth = synth.make_hop_set(51, N_BLOCKS * B, n_transmitters=8, hop_samples=hop, snr_db=25.0, seed=51)
    want = _rows(CO.batch_records(iq))
    cfg = protocol.new_packet_config(14)
    rng = np.random.default_rng(3)
    ragged, k = [], 0
    while k < N_BLOCKS:
        c = int(min(N_BLOCKS - k, rng.integers(1, 24)))
        ragged.append(c)
        k += c
    for chunks in ([1] * N_BLOCKS, ragged):
        bd = dsp.BatchDemodulator(cfg, 51, max_blocks=max(chunks))
        got, k = [], 0
        for c in chunks:
            bd.submit(np.ascontiguousarray(iq[:, k * 2 * B:(k + c) * 2 * B]))
            got += _rows(bd.fetch())
            k += c
        bd.close()
        assert sorted(got) == sorted(want)
    valid = {(r[0], bytes(protocol.swap_bit_order(b) for b in r[3][2:])) for r in want if r[4]}
    sent = [(ch, e.message) for ch, pk in enumerate(truth) for e in pk
            if e.start + synth.BURST_BITS * synth.SPS < (N_BLOCKS - 1) * B]   # last block never flushed
    assert len(sent) >= 24 and all(p in valid for p in sent)


# ---------------------------------------------------------------------------
# config 4: 4 096 streams, 0..20 dB
# ---------------------------------------------------------------------------
def test_config4_snr_sweep_4096_streams_10s(dsp, protocol):
    _need_gb(8)
    snrs = list(range(0, 21, 2))
    per = 372
    cfg = protocol.new_packet_config(14)
    total = bad = streams = 0
    recovered = {}
    for i, snr in enumerate(snrs):
        ns = per + (4096 - per * len(snrs) if i == len(snrs) - 1 else 0)
        buf, msgs = _device_batch(ns, N_BLOCKS, 7000 + i, snr)
        with dsp.BatchDemodulator(cfg, ns, max_blocks=N_BLOCKS) as bd:
            bd.submit(buf)
            hits = bd.fetch()
        host = buf.cpu().numpy()
        del buf
        want = CO.batch_records(host)
        same = _rows(hits) == _rows(want)
        bad += 0 if same else 1
        total += want.shape[0]
        streams += ns
        valid = {(int(h["stream"]), bytes(protocol.swap_bit_order(b) for b in h["data"][2:]))
                 for h in hits[hits["crc_ok"] != 0]}
        recovered[snr] = sum((s, msgs[s, k].tobytes()) in valid for s in range(ns) for k in range(40)) / (ns * 40)
        print(f"config 4: {snr:2d} dB  {ns} streams  {want.shape[0]} records  "
              f"{'equal' if same else 'MISMATCH'}  recovered {100 * recovered[snr]:.2f} %")
    assert streams == 4096 and total > 400_000 and bad == 0
    assert recovered[0] == 0.0 and recovered[20] > 0.99
    assert all(recovered[a] <= recovered[b] + 0.01 for a, b in zip(snrs, snrs[1:]))


# ---------------------------------------------------------------------------
# config 5: one GPU's shard of the 65 536-stream run (8 192 streams x 10 s, 44 GB)
# ---------------------------------------------------------------------------
def test_config5_one_gpu_shard_8192_streams_sampled(dsp, protocol):
    _need_gb(52)
    import torch
    n_streams = 8192
    buf, msgs = _device_batch(n_streams, N_BLOCKS, 4242, 20.0)
    with dsp.BatchDemodulator(protocol.new_packet_config(14), n_streams, max_blocks=N_BLOCKS) as bd:
        bd.submit(buf)
        hits = bd.fetch()
    pick = np.arange(0, n_streams, 8)
    host = buf[torch.from_numpy(pick).cuda()].cpu().numpy()
    del buf
    want = CO.batch_records(host)
    want["stream"] = pick[want["stream"]]
    sel = hits[hits["stream"] % 8 == 0]
    assert want.shape[0] > 100_000 and _rows(sel) == _rows(want)
    valid = hits[hits["crc_ok"] != 0]
    got = {(int(s), bytes(protocol.swap_bit_order(x) for x in d[2:])) for s, d in zip(valid["stream"], valid["data"])}
    back = sum((s, msgs[s, k].tobytes()) in got for s in range(n_streams) for k in range(40))
    print(f"config 5 shard: {want.shape[0]} records of every 8th stream equal to the oracle's; "
          f"{back} of {n_streams * 40} embedded recovered")
    assert back >= 0.995 * n_streams * 40


# ---------------------------------------------------------------------------
# gate (ii) at scale: decision bits of whole 10 s streams
# ---------------------------------------------------------------------------
def test_decision_bits_of_16_whole_streams(dsp, protocol):
    total = bad_total = outside = 0
    for i, snr in enumerate([4.0, 10.0, 20.0, 30.0]):
        buf, _ = _device_batch(4, N_BLOCKS, 9100 + i, snr)
        bd = dsp.BatchDemodulator(protocol.new_packet_config(14), 4, max_blocks=N_BLOCKS)
        bd.submit(buf)
        bd.fetch()
        host = buf.cpu().numpy()
        for s in range(4):
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
                outside += int(np.sum(np.abs(ref_disc[lo + bad]) > 1e-5 * scale[lo + bad]))
            total += want.size
            bad_total += bad.size
        bd.close()
    print(f"{total} samples, {bad_total} fp32 decisions differ from the float64 sign bit, "
          f"{outside} of them outside the 1e-5 band")
    assert total > 40_000_000 and outside == 0 and bad_total <= 1e-6 * total + 2
