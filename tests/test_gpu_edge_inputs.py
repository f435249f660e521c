"""Inputs the reference answers and a list sized for noise would not: a transmitter that repeats
the sync word, constant bytes, a full-scale tone, alternating bytes; and the
state bookkeeping around per-stream reset and a caller-sized packet buffer that overflows
(ADVICE r01).  Every record against the C oracle (src/rtldavis/dsp.py:171-205)."""
import numpy as np
import pytest

from oracle import c_oracle as CO

pytestmark = pytest.mark.gpu

B = 8192


@pytest.fixture(scope="module")
def dsp():
    from rtldavis_b200 import dsp as d
    return d


@pytest.fixture(scope="module")
def cfg():
    from rtldavis_b200 import protocol
    return protocol.new_packet_config(14)


def _rows(rec):
    return [(int(r["stream"]), int(r["call"]), int(r["index"]), r["data"].tobytes(), bool(r["crc_ok"]))
            for r in rec]


def _oracle(iq):
    return _rows(CO.batch_records(iq, per_stream_cap=1 << 17))


def sync_word_forever(n_samples, seed, snr_db=25.0, df_hz=800.0):
    from rtldavis_b200 import synth
    return synth.make_sync_flood(n_samples, seed, snr_db, df_hz)


def test_a_stream_that_repeats_the_sync_word_is_answered_not_refused(dsp, cfg):
    """~1 460 repetitions x several sample phases in one stream: candidate, confirmed and record
    lists sized for noise overflow; the chain runs again with larger ones."""
    n_blocks = 64
    iq = sync_word_forever(n_blocks * B, 5).reshape(1, -1)
    want = _oracle(iq)
    assert len(want) > 500                                   # after the per-call dedupe of identical payloads
    bd = dsp.BatchDemodulator(cfg, 1, max_blocks=n_blocks)
    bd.submit(iq)
    hits = bd.fetch()
    st = bd.stats()
    assert _rows(hits) == want
    # ~2 300 repetitions x several sample phases: far beyond 4 096 + 4 x the noise expectation
    assert st.reruns >= 1 and st.confirmed > 6000
    # the stream goes on afterwards as if nothing had happened: the next call continues the state
    from rtldavis_b200 import synth
    more = synth.make_capture(77, 8 * B, 2, 30.0).iq.reshape(1, -1)
    bd.submit(more)
    tail = _rows(bd.fetch())
    both = _oracle(np.concatenate([iq, more], axis=1))
    assert tail == [r for r in both if r[1] >= n_blocks]
    bd.close()


def test_dense_stream_inside_a_device_resident_batch_with_two_chains(dsp, cfg):
    """70 device-resident streams (one demodulation launch, two search/finish chains): one dense
    stream in each half."""
    import torch
    from rtldavis_b200 import synth
    n_blocks = 48
    iq, _ = synth.make_batch(70, n_blocks * B, 2, 20.0, seed0=6100)
    iq[3] = sync_word_forever(n_blocks * B, 11)
    iq[58] = sync_word_forever(n_blocks * B, 12, df_hz=-1200.0)
    want = _oracle(iq)
    bd = dsp.BatchDemodulator(cfg, 70, max_blocks=n_blocks)
    bd.submit(torch.from_numpy(iq).cuda())
    got = _rows(bd.fetch())
    assert bd.stats().reruns >= 2 and got == want
    bd.close()


def test_dense_stream_in_host_pipeline_groups(dsp, cfg):
    """Host input, groups of 2 streams cycling through 2 stages: the dense group's staging buffer
    has been reused by a later group when its overflow is seen."""
    from rtldavis_b200 import synth
    n_blocks = 48
    iq, _ = synth.make_batch(9, n_blocks * B, 2, 20.0, seed0=6200)
    iq[2] = sync_word_forever(n_blocks * B, 13)
    want = _oracle(iq)
    bd = dsp.BatchDemodulator(cfg, 9, max_blocks=n_blocks, host_slots=2, host_group=2)
    bd.submit(iq)
    got = _rows(bd.fetch())
    assert bd.stats().reruns >= 1 and got == want
    bd.close()


@pytest.mark.parametrize("kind", ["zeros", "ones", "mid", "alternating", "tone_lo", "tone_hi"])
def test_degenerate_inputs_match_the_oracle(dsp, cfg, kind):
    n_blocks = 6
    n = n_blocks * B
    if kind == "zeros":
        iq = np.zeros(2 * n, dtype=np.uint8)
    elif kind == "ones":
        iq = np.full(2 * n, 255, dtype=np.uint8)
    elif kind == "mid":
        iq = np.full(2 * n, 127, dtype=np.uint8)
    elif kind == "alternating":
        iq = np.tile(np.array([0, 255], dtype=np.uint8), n)
    else:
        from rtldavis_b200 import synth
        f = -synth.FS / 4.0 + (-4800.0 if kind == "tone_lo" else 4800.0)
        iq = synth.to_u8(np.exp(2j * np.pi * f * np.arange(n) / synth.FS))       # full scale, clips at 255
    iq = iq.reshape(1, -1)
    bd = dsp.BatchDemodulator(cfg, 1, max_blocks=n_blocks)
    bd.submit(iq)
    hits = bd.fetch()
    assert _rows(hits) == _oracle(iq)
    first, bits = bd.debug_bits(0)
    ref = CO.stream_bits(iq[0])
    lo = max(first, 16)
    got = bits[lo - first: lo - first + ref.size - lo]
    if kind in ("tone_lo", "tone_hi"):
        assert np.array_equal(got, ref[lo:lo + got.size])      # a clean tone: the decision is never marginal
    bd.close()


def test_reset_of_one_stream_keeps_the_call_bookkeeping(dsp, cfg):
    """BatchDemodulator.reset(stream) restarts one stream (history and call counter, dsp.py:248-253)
    while the others go on; demodulate() must keep filing every packet under the right call."""
    from rtldavis_b200 import synth
    iq, _ = synth.make_batch(3, 12 * B, 5, 30.0, seed0=8100)
    a, b = iq[:, : 2 * B * 5], iq[:, 2 * B * 5:]
    bd = dsp.BatchDemodulator(cfg, 3, max_blocks=7)
    first = bd.demodulate(a)
    bd.reset(1)
    second = bd.demodulate(b)
    whole = CO.batch_packets(iq)
    fresh = CO.batch_packets(b[1])[0]                         # stream 1 starts over at call 0
    for s in (0, 2):
        want = [[(p.index, p.data) for p in whole[s] if p.call == k] for k in range(12)]
        got = [[(p.index, p.data.tobytes()) for p in call] for call in first[s] + second[s]]
        assert got == want
    want1 = [[(p.index, p.data) for p in fresh if p.call == k] for k in range(7)]
    assert [[(p.index, p.data.tobytes()) for p in call] for call in second[1]] == want1
    assert bd.calls_done.tolist() == [12, 7, 12]
    bd.close()


def test_going_on_after_a_caller_sized_buffer_overflowed(dsp, cfg):
    """max_hits too small: the fetch raises (never truncates); the streams have advanced, and the
    next calls are filed correctly."""
    from rtldavis_b200 import _lib, synth
    iq, _ = synth.make_batch(4, 16 * B, 8, 30.0, seed0=8200)
    bd = dsp.BatchDemodulator(cfg, 4, max_blocks=8, max_hits=8)
    bd.submit(iq[:, : 2 * B * 8])
    with pytest.raises(_lib.DavisError) as e:
        bd.fetch()
    assert e.value.code == _lib.E_OVERFLOW
    assert bd.calls_done.tolist() == [8] * 4
    # second half in calls small enough for the buffer
    whole = CO.batch_packets(iq)
    for k in range(8, 16):
        out = bd.demodulate(iq[:, 2 * B * k: 2 * B * (k + 1)])
        for s in range(4):
            assert [(p.index, p.data.tobytes()) for p in out[s][0]] == \
                   [(p.index, p.data) for p in whole[s] if p.call == k]
    bd.close()


def test_small_calls_replay_a_cached_graph_and_stay_exact(dsp, cfg):
    """One block of one stream is launch-bound: its chain is captured into a CUDA graph once per
    carry buffer and replayed.  Same packets as the oracle, block by block, across a buffer
    re-allocation (a larger call in between) and a flood that forces a direct re-run."""
    from rtldavis_b200 import synth
    iq = synth.make_capture(4100, 40 * B, 10, 20.0, df_hz=1200.0).iq
    iq[2 * B * 20: 2 * B * 30] = synth.make_sync_flood(10 * B, 7)[: 2 * B * 10]
    want = _oracle(iq.reshape(1, -1))
    bd = dsp.BatchDemodulator(cfg, 1, max_blocks=10)
    got, replays, k = [], 0, 0
    for take in [1] * 12 + [8] + [1] * 20:                     # 12 + 8 + 20 = 40 blocks
        bd.submit(iq[2 * B * k: 2 * B * (k + take)].reshape(1, -1))
        got += _rows(bd.fetch())
        replays += int(bd.stats().graph_replays)
        k += take
    assert k == 40 and got == want
    assert replays >= 30
    bd.close()


def test_host_input_may_be_overwritten_when_its_staging_copy_is_still_there(dsp, cfg):
    """a re-run after a list overflow used to read the caller's host buffer a second
    time at the fetch.  It now does so only when a later group has taken the staging buffer; a
    one-group call re-runs from the library's own copy, so scribbling over the (pageable) input
    between submit and fetch changes nothing."""
    n_blocks = 32
    iq = sync_word_forever(n_blocks * B, 21).reshape(1, -1)
    want = _oracle(iq)
    bd = dsp.BatchDemodulator(cfg, 1, max_blocks=n_blocks)
    mine = iq.copy()
    bd.submit(mine)
    mine[:] = 0x55
    got = _rows(bd.fetch())
    st = bd.stats()
    assert st.reruns >= 1 and got == want
    assert st.h2d_bytes == iq.size                               # the input crossed the link once
    bd.close()


def test_counters_of_a_rerun_call_are_those_of_the_pass_that_was_kept(dsp, cfg):
    """candidates / confirmed / records were summed over discarded passes too.  The
    same call repeated on the same handle (lists already grown, fewer re-runs) must count the same."""
    n_blocks = 48
    iq = sync_word_forever(n_blocks * B, 22).reshape(1, -1)
    bd = dsp.BatchDemodulator(cfg, 1, max_blocks=n_blocks)
    bd.submit(iq)
    first = _rows(bd.fetch())
    a = bd.stats()
    bd.reset()
    bd.submit(iq)
    again = _rows(bd.fetch())
    b = bd.stats()
    # (the packet buffer is sized per call, so the repeat still re-runs for that one list)
    assert a.reruns >= 2 and b.reruns < a.reruns and first == again
    assert (a.candidates, a.confirmed, a.records) == (b.candidates, b.confirmed, b.records)
    assert a.confirmed >= len(first)
    bd.close()


def test_packets_beyond_max_hits_are_reported_once_and_kept(dsp, cfg):
    """the fetch that finds more packets than davis_cfg.max_hits still raises, but
    the packets are no longer thrown away: the next fetch hands out every one of them."""
    from rtldavis_b200 import _lib, synth
    iq, _ = synth.make_batch(4, 10 * B, 6, 30.0, seed0=1500)
    want = _oracle(iq)
    assert len(want) > 8
    bd = dsp.BatchDemodulator(cfg, 4, max_blocks=10, max_hits=8)
    bd.submit(iq)
    with pytest.raises(_lib.DavisError) as e:
        bd.fetch()
    assert e.value.code == _lib.E_OVERFLOW and "kept" in str(e.value)
    assert _rows(bd.fetch()) == want
    assert _rows(bd.fetch()) == want
    bd.close()


def test_small_calls_from_ever_new_device_buffers_stop_capturing_graphs(dsp, cfg):
    """the graph cache is keyed on the input pointer.  A caller whose every block
    sits in a fresh device buffer never hits it; after the cache has filled and missed 16 times
    in a row such calls are launched directly (no capture per call), on the caller's stream or
    the default one, and stay exact.  A buffer the cache knows is replayed again."""
    import torch
    from rtldavis_b200 import synth
    n = 72
    iq = synth.make_capture(4300, (n + 1) * B, 18, 20.0, df_hz=-900.0).iq
    want = _oracle(iq.reshape(1, -1))
    bd = dsp.BatchDemodulator(cfg, 1, max_blocks=1)
    side = torch.cuda.Stream()
    bufs, got, replays = [], [], []
    for k in range(n):
        with torch.cuda.stream(side if k % 2 else torch.cuda.current_stream()):
            d = torch.from_numpy(iq[2 * B * k: 2 * B * (k + 1)].copy()).cuda().reshape(1, -1)
            bufs.append(d)                                        # kept alive: every pointer is new
            bd.submit(d, cuda_stream=side.cuda_stream if k % 2 else None)
        got += _rows(bd.fetch())
        replays.append(int(bd.stats().graph_replays))
    # block 72 arrives in the buffer of call 30 (same carry parity), which the cache still holds
    bufs[30].copy_(torch.from_numpy(iq[2 * B * n:].copy()).reshape(1, -1))
    torch.cuda.synchronize()
    bd.submit(bufs[30])
    got += _rows(bd.fetch())
    replays.append(int(bd.stats().graph_replays))
    assert got == want
    assert sum(replays[:32]) >= 30 and sum(replays[40:60]) == 0 and replays[-1] == 1, replays
    bd.close()
