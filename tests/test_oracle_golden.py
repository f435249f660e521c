"""Pins oracle/ (numpy and C restatements) to the reference.

Golden fixtures under tests/golden were produced by the unmodified reference
(oracle/gen_golden.py); the known-answer vectors below are the reference's own
test vectors, cited per test.
"""
import numpy as np
import pytest

from conftest import TEN_SECOND_FIXTURES, golden_rows, load_golden, load_seeded
from oracle import c_oracle as CO
from oracle import davis_oracle as O

U8_FIXTURES = ["hi_snr_8192", "lo_snr_8192", "go_block_512"]


def _cfg(g):
    return O.OracleConfig(block_size=int(g["block"]))


# ---- reference tests/test_dsp.py:4-33 ------------------------------------
def test_quantize_polarity_kat():
    vals = np.array([-5.0, 5.0, -0.1, 0.1, 0.0])
    assert O.quantize_literal(vals).tolist() == [1, 0, 1, 0, 0]
    assert O.quantize_fast(vals).tolist() == [1, 0, 1, 0, 0]
    assert O.quantize_literal([-0.0]).tolist() == [1]        # bit 63, dsp.py:98
    assert O.quantize_fast(np.array([-0.0])).tolist() == [1]


def test_quantize_random_1000():
    vals = np.random.default_rng(42).uniform(-10, 10, 1000)
    want = (vals < 0).astype(np.uint8)
    assert (O.quantize_literal(vals) == want).all()
    assert (O.quantize_fast(vals) == want).all()


# ---- reference tests/test_protocol.py:3-13 --------------------------------
@pytest.mark.parametrize("a,b", [(0x00, 0x00), (0xFF, 0xFF), (0x01, 0x80), (0x80, 0x01),
                                 (0x55, 0xAA), (0xAA, 0x55)])
def test_swap_bit_order_kat(a, b):
    assert O.swap_bit_order(a) == b
    assert CO.lib().dor_swap_bits(a) == b


# ---- reference tests/test_protocol.py:15-36 -------------------------------
def test_crc_rain_packet_kat():
    air = bytes([0x07, 0xC0, 0x2B, 0x0B, 0x80, 0x40, 0x8E, 0xFF])
    host = bytes(O.swap_bit_order(b) for b in air)
    assert host.hex() == "e003d4d0010271ff"
    assert O.crc16(host) == 0 and CO.crc16(host) == 0
    bad = bytes(O.swap_bit_order(b) for b in air[:-1] + b"\xFE")
    assert O.crc16(bad) != 0 and CO.crc16(bad) != 0
    assert O.crc_gate(b"\xcb\x89" + air) == (b"\xd3\x91" + host, True)


# ---- frames held by the reference's decoder tests -------------------------
# decoders/temperature_test.py:15, humidity_test.py:15,20, debug_tools/scratch_crc.py:10
@pytest.mark.parametrize("frame", ["80052c2cf90b649e", "a006528338005ac8", "a00435d12b00703a",
                                   "81058D33CB0FF1DD"])
def test_crc_real_frames(frame):
    raw = bytes.fromhex(frame)
    assert O.crc16(raw) == 0
    assert CO.crc16(raw) == 0


def test_crc_table_spot_values():
    t = O.crc_table(0x1021)
    assert int(t[0]) == 0 and int(t[1]) == 0x1021 and int(t[255]) == 0x1EF0


# ---- crc/crc_test.go:22-41 (identity property, same three parameter sets) --
@pytest.mark.parametrize("init,poly", [(0, 0x8005), (0, 0x6F63), (0xFFFF, 0x1021), (0, 0x1021)])
def test_crc_identity_property(init, poly):
    rng = np.random.default_rng(poly)
    tbl = O.crc_table(poly)
    for _ in range(512):
        n = (int(rng.integers(0, 32)) & 0xFE) + 8
        body = bytes(rng.integers(0, 256, n - 2, dtype=np.uint8).tolist())
        c = O.crc16(body, init, tbl)
        whole = body + bytes([c >> 8, c & 0xFF])
        assert O.crc16(whole, init, tbl) == 0


# ---- end to end against the reference's Demodulator -----------------------
@pytest.mark.parametrize("name", U8_FIXTURES + ["complex_in_8192"])
def test_call_oracle_matches_reference(name):
    g = load_golden(name)
    cfg = _cfg(g)
    o = O.CallOracle(cfg)
    iq, B = g["iq"], cfg.block_size
    got, bits = [], []
    for k in range(iq.size // (2 * B)):
        blk = iq[2 * B * k:2 * B * (k + 1)]
        if bool(g["as_complex"]):
            blk = O.bytes_to_complex(blk).astype(np.complex64)
        got.extend(o.demodulate(blk))
        bits.append(o.quantized[-B:].copy())
    assert [(p.call, p.index, p.data, p.crc_ok) for p in got] == golden_rows(g)
    np.testing.assert_allclose([p.rssi for p in got], g["pk_rssi"], rtol=0, atol=1e-9)
    np.testing.assert_allclose([p.snr for p in got], g["pk_snr"], rtol=0, atol=1e-9)
    want_bits = np.unpackbits(g["bits"])[: int(g["n_bits"])]
    assert (np.concatenate(bits) == want_bits).all()


def test_call_oracle_literal_quantize_is_the_same():
    g = load_golden("go_block_512")
    got = O.run_calls(g["iq"], _cfg(g), literal_quantize=True)
    assert [(p.call, p.index, p.data, p.crc_ok) for p in got] == golden_rows(g)


@pytest.mark.parametrize("name", U8_FIXTURES)
def test_stream_formulation_matches_reference(name):
    g = load_golden(name)
    cfg = _cfg(g)
    want_bits = np.unpackbits(g["bits"])[: int(g["n_bits"])]
    assert (O.stream_bits(g["iq"]) == want_bits).all()
    lo, hi = int(g["disc_lo"]), int(g["disc_hi"])
    np.testing.assert_array_equal(O.stream_discriminated(g["iq"])[lo:hi], g["disc"])
    got = O.stream_oracle(g["iq"], cfg)
    assert [(p.call, p.index, p.data, p.crc_ok) for p in got] == golden_rows(g)


@pytest.mark.parametrize("name", U8_FIXTURES)
def test_c_oracle_matches_reference(name):
    g = load_golden(name)
    want_bits = np.unpackbits(g["bits"])[: int(g["n_bits"])]
    bits, disc = CO.stream_bits(g["iq"], want_disc=True)
    assert (bits == want_bits).all()
    lo, hi = int(g["disc_lo"]), int(g["disc_hi"])
    ref = g["disc"]
    assert np.all(np.abs(disc[lo:hi] - ref) <= 1e-12 * np.maximum(np.abs(ref), 1.0))
    got = CO.batch_packets(g["iq"], block=int(g["block"]))[0]
    assert [(p.call, p.index, p.data, p.crc_ok) for p in got] == golden_rows(g)


# ---- whole 10 s captures: BASELINE config 1 as stated, 6 dB, noise only, +-45.75 kHz ----
@pytest.mark.parametrize("name", TEN_SECOND_FIXTURES)
def test_oracles_match_reference_on_10s_captures(name):
    """328 blocks through the reference (oracle/gen_golden.py): every packet row, RSSI/SNR and
    all 2 686 976 quantized bits, against the numpy call-by-call restatement, the numpy stream
    formulation and the C restatement."""
    g, iq = load_seeded(name)
    cfg = _cfg(g)
    B = cfg.block_size
    want_rows = golden_rows(g)
    want_bits = np.unpackbits(g["bits"])[: int(g["n_bits"])]
    o = O.CallOracle(cfg)
    got = []
    for k in range(iq.size // (2 * B)):
        got.extend(o.demodulate(iq[2 * B * k:2 * B * (k + 1)]))
    assert [(p.call, p.index, p.data, p.crc_ok) for p in got] == want_rows
    np.testing.assert_allclose([p.rssi for p in got], g["pk_rssi"], rtol=0, atol=1e-9)
    np.testing.assert_allclose([p.snr for p in got], g["pk_snr"], rtol=0, atol=1e-9)
    assert (O.stream_bits(iq) == want_bits).all()
    assert (CO.stream_bits(iq) == want_bits).all()
    got_c = CO.batch_packets(iq, block=B)[0]
    assert [(p.call, p.index, p.data, p.crc_ok) for p in got_c] == want_rows


@pytest.mark.parametrize("name", ["config1_10s", "noise_10s"])
def test_fixtures_reproduce_from_the_unmodified_reference(name):
    """The fixtures are what the reference returns: run it again (oracle/_ref or /root/reference)
    on the regenerated bytes and compare every row and every quantized bit with the stored ones."""
    from oracle import ref_loader
    if not ref_loader.available():
        pytest.skip("the reference is neither staged in oracle/_ref nor at /root/reference")
    ref_dsp, ref_crc, ref_protocol = ref_loader.load()
    g, iq = load_seeded(name)
    B = int(g["block"])
    dem = ref_dsp.Demodulator(ref_protocol.new_packet_config(14))
    gate = ref_crc.CRC("CCITT-16", 0, 0x1021, 0)
    rows, bits = [], []
    for k in range(int(g["n_blocks"])):
        for p in dem.demodulate(iq[2 * B * k:2 * B * (k + 1)]):
            host = bytes(ref_protocol.swap_bit_order(int(b)) for b in p.data)
            rows.append((k, int(p.index), bytes(p.data), int(gate.checksum(host[2:])) == 0))
        bits.append(dem.quantized[dem.cfg.buffer_length - B:].copy())
    assert rows == golden_rows(g)
    assert (np.concatenate(bits) == np.unpackbits(g["bits"])[: int(g["n_bits"])]).all()


def test_config1_fixture_is_what_baseline_states():
    """BASELINE configs[0]: 10 s, station 0, 40 embedded packets, all recovered."""
    g = load_golden("config1_10s")
    rows = golden_rows(g)
    assert int(g["n_blocks"]) == 328 and g["truth_msg"].shape[0] == 40
    valid = {O.crc_gate(r[2])[0][2:] for r in rows if r[3]}
    assert {bytes(m.tolist()) for m in g["truth_msg"]} == valid
    assert all(m[0] & 7 == 0 for m in g["truth_msg"])
    assert not any(r[3] for r in golden_rows(load_golden("noise_10s")))
    assert not any(r[3] for r in golden_rows(load_golden("rf50ppm_10s")))


def test_embedded_packets_are_recovered_at_high_snr():
    g = load_golden("hi_snr_8192")
    valid = {O.crc_gate(r[2])[0][2:] for r in golden_rows(g) if r[3]}
    assert {bytes(m.tolist()) for m in g["truth_msg"]} <= valid


def test_boundary_duplicate_is_in_the_fixture():
    """the same payload is reported by consecutive calls when its
    candidate positions straddle a block boundary."""
    rows = golden_rows(load_golden("hi_snr_8192"))
    by_data = {}
    for c, i, d, ok in rows:
        by_data.setdefault(d, []).append((c, i))
    assert any(len({c for c, _ in v}) > 1 for v in by_data.values())


def test_wrong_size_raises():
    o = O.CallOracle()
    with pytest.raises(ValueError):
        o.demodulate(np.zeros(100, dtype=np.uint8))
    with pytest.raises(ValueError):
        o.demodulate(np.zeros(100, dtype=np.complex64))


def test_signed_zero_after_reset():
    """dsp.py:98 takes bit 63, so the two samples computed from zero history
    follow IEEE signed-zero rules (hard part 3)."""
    for i, q in [(10, 200), (200, 10), (10, 10), (200, 200)]:
        iq = np.full(2 * 8192, 128, dtype=np.uint8)
        iq[0], iq[1] = i, q
        bits = O.stream_bits(iq)
        assert bits[0] == 0
        assert bits[1] == (1 if (i < 127.4 and q > 127.4) else 0)
        assert (CO.stream_bits(iq)[:16] == bits[:16]).all()


def test_staged_reference_of_another_python_counts_as_absent(tmp_path, monkeypatch):
    """oracle/_ref is byte-code of the build container's Python.  On a box with
    another minor version the loader must say "not available" (callers then skip or time the
    port) instead of failing inside the import with "bad magic number"."""
    import importlib.util

    from oracle import ref_loader
    fake = tmp_path / "_ref" / "rtldavis"
    fake.mkdir(parents=True)
    (fake / "dsp.rpyc").write_bytes(b"\x00\x00\r\n" + b"\x00" * 12)        # not this interpreter's magic
    monkeypatch.setattr(ref_loader, "_REF_DIR", str(tmp_path / "_ref"))
    monkeypatch.setattr(ref_loader, "_SRC_DIR", str(tmp_path / "nowhere"))
    assert not ref_loader.available() and ref_loader.kind() == "port"
    with pytest.raises(ImportError):
        ref_loader.load()
    (fake / "dsp.rpyc").write_bytes(importlib.util.MAGIC_NUMBER + b"\x00" * 12)
    assert ref_loader.available() and ref_loader.kind() == "ref"
