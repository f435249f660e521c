"""The Go mirror's numbers: no epsilon in the discriminator (dsp/dsp.go:93),
block size 512 (dsp/dsp.go:212), the parser's frequency error from symbols 8..24 with the
9600 Hz bias (protocol/protocol.go:156-168).

The Go restatement in oracle/davis_oracle.py is parity-unpinned (no Go toolchain in the build
container); what these tests pin is (a) that it agrees with the pinned Python restatement on
everything the two implementations share and (b) that the CUDA path, with
``davis_cfg.flavor = DAVIS_FLAVOR_GO``, returns that restatement's numbers."""
import math

import numpy as np
import pytest

from conftest import golden_rows, load_golden
from oracle import davis_oracle as O

GO_CFG = O.OracleConfig(block_size=512)


def _go_capture():
    g = load_golden("go_block_512")
    return g, g["iq"]


# ---------------------------------------------------------------------------
# CPU: the restatement itself
# ---------------------------------------------------------------------------
def test_go_fir_is_the_python_fir_up_to_rounding():
    rng = np.random.default_rng(5)
    iq = rng.standard_normal(4096 + 9) + 1j * rng.standard_normal(4096 + 9)
    a = O.fir9_valid(iq, 4096)
    b = O.go_fir9(iq, 4096)
    np.testing.assert_allclose(b, a, rtol=0, atol=4e-16 * np.abs(iq).max())


def test_go_discriminator_has_no_epsilon_and_nan_on_zero_history():
    f = np.array([0, 1 + 1j, 1e-6 - 2e-6j, -3 + 0.5j], dtype=np.complex128)
    d = O.go_discriminate(f)
    assert math.isnan(d[0])
    a, b = f[1:-1], f[2:]
    np.testing.assert_array_equal(d[1:], (a.imag * b.real - a.real * b.imag) / (a.real ** 2 + a.imag ** 2))
    # the Python form differs by the epsilon only (dsp.py:88)
    p = O.discriminate(f)
    assert p[0] == 0.0 and abs(p[2] - d[2]) > 0 and abs(p[1] - d[1]) < 1e-9
    assert O.go_quantize(d).tolist() == [0, int(d[1] < 0), int(d[2] < 0)]


def test_go_call_oracle_returns_the_reference_packets():
    """Decision bits and packets do not depend on the flavor: the Go restatement must give the
    rows the unmodified Python reference gave for the same capture at block size 512."""
    g, iq = _go_capture()
    o = O.GoCallOracle(GO_CFG)
    got = []
    for k in range(iq.size // 1024):
        got.extend(o.demodulate(iq[1024 * k:1024 * (k + 1)]))
    assert [(p.call, p.index, p.data, p.crc_ok) for p in got] == golden_rows(g)


def test_go_freq_error_formula():
    disc = np.zeros(1024)
    disc[100 + 112:100 + 336] = 0.25                       # symbols 8..24 after index 100
    disc[:100 + 112] = 9.0
    disc[100 + 336:] = -9.0
    mean, err = O.go_freq_error(disc, 100, GO_CFG)
    assert mean == 0.25
    assert err == -int(9600 + 0.25 * 268800 / (2 * math.pi))
    mean, err = O.go_freq_error(np.full(1024, np.nan), 0, GO_CFG)
    assert math.isnan(mean) and err is None


def test_packet_config_flavor_reaches_the_c_struct():
    from rtldavis_b200 import dsp, protocol
    c = protocol.new_go_packet_config()
    assert (c.block_size, c.buffer_length, c.flavor) == (512, 2048, "go")
    assert c._to_c(4).flavor == 1
    assert protocol.new_packet_config(14)._to_c(4).flavor == 0
    with pytest.raises(ValueError):
        dsp.PacketConfig(19200, 14, 16, 80, "1100101110001001", flavor="rust")


# ---------------------------------------------------------------------------
# GPU: the CUDA path with flavor = go against the restatement
# ---------------------------------------------------------------------------
def _go_want(iq, n_blocks, cfg=GO_CFG):
    B = cfg.block_size
    o = O.GoCallOracle(cfg)
    want = {}
    last_disc = None
    for k in range(n_blocks):
        for p in o.demodulate(iq[2 * B * k:2 * B * (k + 1)]):
            want[(k, p.index, p.data)] = (p.crc_ok,) + O.go_freq_error(o.discriminated, p.index, cfg)
        last_disc = o.discriminated.copy()
    return want, last_disc


@pytest.mark.gpu
@pytest.mark.parametrize("chunk", [96, 7])
def test_go_flavor_disc_mean_and_freq_error(chunk):
    from rtldavis_b200 import dsp, protocol
    g, iq = _go_capture()
    n_blocks = iq.size // 1024
    want, last_disc = _go_want(iq, n_blocks)
    assert sum(w[0] for w in want.values()) >= 3
    cfg = protocol.new_go_packet_config()
    bd = dsp.BatchDemodulator(cfg, 1, max_blocks=chunk)
    hits = []
    for k0 in range(0, n_blocks, chunk):
        n = min(chunk, n_blocks - k0)
        bd.submit(iq[None, 1024 * k0:1024 * (k0 + n)])
        hits.extend(bd.fetch())
    got = {(int(h["call"]), int(h["index"]), h["data"].tobytes()): h for h in hits}
    assert set(got) == set(want)
    for key, (ok, mean, err) in want.items():
        h = got[key]
        assert bool(h["crc_ok"]) == ok
        if ok:
            # fp64 on both sides; the device sums in the Python tap order, Go folds pairs
            assert abs(float(h["disc_mean"]) - mean) <= 1e-9 * abs(mean) + 1e-12
        else:
            assert math.isnan(float(h["disc_mean"]))
    parsed = {(p["call"], p["index"]): p["freq_err"] for p in protocol.parse_hits(np.array(hits), cfg)}
    assert parsed == {(k[0], k[1]): w[2] for k, w in want.items() if w[0]}
    # the buffer a Go caller reads (Demodulator.Discriminated, dsp/dsp.go:279,286)
    d = bd.discriminated(0)
    np.testing.assert_allclose(d, last_disc, rtol=1e-9, atol=1e-13)
    bd.close()


@pytest.mark.gpu
def test_go_flavor_zero_history_is_nan_like_go():
    """Call 0 of a fresh stream: Discriminated[512] = D[0] and Discriminated[513] = D[1] are 0/0
    (Filtered[0] is the zero carry and Filtered[1] = F[0] sees only the zero FIR history)."""
    from rtldavis_b200 import dsp, protocol
    _, iq = _go_capture()
    want, last_disc = _go_want(iq, 1)
    assert np.isnan(last_disc[512:514]).all() and not np.isnan(last_disc[514:]).any()
    bd = dsp.BatchDemodulator(protocol.new_go_packet_config(), 1, max_blocks=1)
    bd.submit(iq[None, :1024])
    bd.fetch()
    d = bd.discriminated(0)
    assert np.isnan(d[512:514]).all() and (d[:512] == 0).all()
    np.testing.assert_allclose(d[514:], last_disc[514:], rtol=1e-9, atol=1e-13)
    bd.close()


@pytest.mark.gpu
def test_go_flavor_at_block_8192_batch():
    """The flavor is independent of the block size: a seeded batch at 8192 (the finish
    kernel's integer path) against the Go restatement, and against the Python flavor for
    everything but the mean."""
    from rtldavis_b200 import dsp, synth
    n_streams, n_blocks, B = 6, 10, 8192
    iq, _ = synth.make_batch(n_streams, n_blocks * B, 8, 16.0, seed0=4100, df_range_hz=2500.0,
                             ppm_range=30.0)
    cfg8 = O.OracleConfig(block_size=B)
    want = {}
    for s in range(n_streams):
        w, _ = _go_want(iq[s], n_blocks, cfg8)
        want.update({(s,) + k: v for k, v in w.items()})
    rows = {}
    for flavor in ("go", "python"):
        c = dsp.PacketConfig(19200, 14, 16, 80, "1100101110001001", block_size=B, flavor=flavor)
        bd = dsp.BatchDemodulator(c, n_streams, max_blocks=n_blocks)
        bd.submit(iq)
        rows[flavor] = bd.fetch().copy()
        bd.close()
    go, py = rows["go"], rows["python"]
    for f in ("position", "stream", "call", "index", "data", "crc_ok", "rssi", "snr"):
        np.testing.assert_array_equal(go[f], py[f])
    got = {(int(h["stream"]), int(h["call"]), int(h["index"]), h["data"].tobytes()): h for h in go}
    assert set(got) == set(want)
    n_ok = 0
    for key, (ok, mean, _) in want.items():
        if ok:
            n_ok += 1
            assert abs(float(got[key]["disc_mean"]) - mean) <= 1e-9 * abs(mean) + 1e-12
    assert n_ok >= 20
