"""The guard band of the fp32 decisions (davis_search.cu: guarded_bits).

k_exact keeps an fp32 sign only when |num32| exceeds tau = 1e-4*S1 + 2^-22*S2 + 1e-6 and
sends everything else to the fp64 reference-phase form.  This models the kernel's fp32
arithmetic in numpy (same operation order; an FMA is emulated by rounding the float64
result of the exact float32 product plus the addend) and checks the claim the design
rests on: outside the band the fp32 sign equals the reference's float64 sign bit, on
inputs built to sit near zero (weak noise, strong tones, a real capture)."""
import numpy as np
import pytest

from conftest import load_golden
from oracle import davis_oracle as O

C = [0.017682261285, 0.048171339939, 0.122424706672, 0.197408519126, 0.228626345955]
SRE, SIM = [-0.4, 0.4, 0.4, -0.4], [-0.4, -0.4, 0.4, 0.4]
f32 = np.float32


def _fma(a, b, c):
    return (a.astype(np.float64) * b.astype(np.float64) + c.astype(np.float64)).astype(f32)


def _residual(k, sgn):
    taps = C + C[3::-1]
    return f32(sum(t * sgn[(k + m) & 3] for m, t in enumerate(taps)))


def _model(u8):
    """(bits, sure) for samples g >= 10 of a byte stream, as guarded_bits computes them
    (rotation phase = global sample index, which the kernel may shift by a common j^p)."""
    i = u8[0::2].astype(f32) - f32(127)
    q = u8[1::2].astype(f32) - f32(127)
    n = i.size
    ph = np.arange(n) & 3
    re = np.select([ph == 0, ph == 1, ph == 2], [i, -q, -i], q).astype(f32)
    im = np.select([ph == 0, ph == 1, ph == 2], [q, i, -q], -i).astype(f32)
    k = np.arange(n - 9)                       # output k uses samples k .. k+8: F[g = k + 9], g < n
    c = [f32(x) for x in C]

    def fir(r, sgn):
        off = np.array([_residual(p, sgn) for p in range(4)], dtype=f32)[k & 3]
        a = _fma(np.full(k.size, c[4]), r[k + 4], off)
        for m in range(4):
            a = _fma(np.full(k.size, c[m]), (r[k + m] + r[k + 8 - m]).astype(f32), a)
        return a

    fr, fi = fir(re, SRE), fir(im, SIM)
    ar, ai, br, bi = fr[:-1], fi[:-1], fr[1:], fi[1:]
    t1 = (ar * bi).astype(f32)
    num = _fma(ai, br, -t1)
    s1 = (np.abs(ar) + np.abs(ai)) + (np.abs(br) + np.abs(bi))
    s2 = np.abs((ai * br).astype(f32)) + np.abs(t1)
    tau = f32(1e-4) * s1 + f32(2.3841858e-7) * s2 + f32(1e-6)
    return np.signbit(num), np.abs(num) > tau   # entry j decides sample g = j + 10


def _check(u8, expect_unsure_below=0.05):
    bits, sure = _model(u8)
    ref = O.stream_bits(u8)[10:10 + bits.size].astype(bool)
    wrong = np.flatnonzero(sure & (bits != ref))
    assert wrong.size == 0, f"fp32 sign kept but wrong at {wrong[:5] + 10}"
    assert (~sure).mean() < expect_unsure_below
    return int((~sure).sum()), int((bits != ref).sum())


def test_guard_band_on_noise_floors_and_tones():
    rng = np.random.default_rng(7)
    n = 400_000
    total_unsure = total_diff = 0
    # weak noise around mid-scale: |F| ~ 1, numerators near zero all the time
    for sigma in (0.6, 1.5, 6.0, 40.0):
        u8 = np.clip(np.rint(127.4 + sigma * rng.standard_normal(2 * n)), 0, 255).astype(np.uint8)
        a, b = _check(u8, 0.05)
        total_unsure += a
        total_diff += b
    # constant input: every F is the same number, every numerator exactly cancels
    _check(np.full(2 * 5000, 127, dtype=np.uint8), 1.01)
    _check(np.full(2 * 5000, 255, dtype=np.uint8), 1.01)
    # full-scale tone at the filter centre (-Fs/4) with a slow drift through zero frequency
    t = np.arange(n)
    phase = -np.pi / 2 * t + 0.2 * np.sin(2 * np.pi * t / 5000.0)
    x = 127.4 + 127.0 * np.stack([np.cos(phase), np.sin(phase)], axis=1).ravel()
    a, b = _check(np.clip(np.rint(x), 0, 255).astype(np.uint8), 0.05)
    assert total_unsure > 0                    # the band is exercised, not vacuous


@pytest.mark.parametrize("name", ["hi_snr_8192", "lo_snr_8192"])
def test_guard_band_on_captures(name):
    u8 = np.asarray(load_golden(name)["iq"], dtype=np.uint8).ravel()
    unsure, diff = _check(u8, 0.01)
    assert diff <= unsure
