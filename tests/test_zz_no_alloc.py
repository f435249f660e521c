"""davis_disc32_static, the entry point for callers that must not allocate on the GPU.

Some hosts of this library cannot call cudaMalloc while they work: a receive thread on the
dongle's per-block deadline (cudaMalloc and cudaFree can stall until the device is idle), a
plug-in whose process has given all device memory to another allocator, a process held to a
fixed memory budget.  For them the library has one entry point whose working memory is a
fixed array inside the library image (include/davis_b200.h, rtldavis_b200/csrc/davis_static.cu).

This file checks, through the C ABI, that it reports the same fp32 discriminator as the
allocating entry point it shadows (davis_demod_u8 + davis_debug_disc32 on a fresh handle), on
calls that follow one another and therefore find each other's leftovers in the shared array,
and that it turns away a size its array was not built for.

The file name sorts last on purpose: pytest -x stops at the first failure, and a failure
here must not hide the parity tests.  It passes alone in a fresh process and after the
other tests (nothing here depends on earlier state: the first call does its own set-up).
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

BLOCK = 8192
# the tolerance tests/test_gpu_parity.py puts on the fp32 discriminator: 1e-5 relative (the two
# entry points launch the same kernel on the same bytes, so in practice they agree bit for bit)
RTOL = 1e-5


@pytest.fixture(scope="module")
def torch_cuda():
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch


def _capture(seed, n_samples):
    from rtldavis_b200 import synth
    return synth.make_capture(seed, n_samples, 2, snr_db=15.0, df_hz=800.0).iq


def _allocating_entry_point(iq_bytes):
    """fp32 discriminator of a fresh stream's first two blocks: davis_demod_u8, then
    davis_debug_disc32 (which cudaMallocs its working memory on every call)."""
    from rtldavis_b200 import dsp, protocol
    bd = dsp.BatchDemodulator(protocol.new_packet_config(14), 1, max_blocks=2)
    try:
        bd.submit(np.ascontiguousarray(iq_bytes).reshape(1, -1))
        bd.fetch()
        return bd.discriminated_fp32(0).copy()
    finally:
        bd.close()


def _static_entry_point(torch, iq_bytes, n_samples, out=None, expect=0):
    from rtldavis_b200 import _lib
    L = _lib.load()
    d_in = torch.from_numpy(np.ascontiguousarray(iq_bytes)).cuda()
    d_out = out if out is not None else torch.full((max(n_samples, 1),), float("nan"),
                                                   dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    err = C.create_string_buffer(256)
    stream = torch.cuda.current_stream().cuda_stream
    rc = L.davis_disc32_static(d_in.data_ptr(), n_samples, d_out.data_ptr(), 0, stream, err, len(err))
    torch.cuda.synchronize()
    assert rc == expect, (rc, err.value.decode())
    return d_out.cpu().numpy()[:n_samples], err.value.decode()


def _same(got, want):
    assert got.shape == want.shape and np.all(np.isfinite(got))
    scale = float(np.max(np.abs(want)))
    np.testing.assert_allclose(got, want, rtol=RTOL, atol=RTOL * scale)
    return int(np.count_nonzero(got != want))


def test_static_scratch_matches_the_allocating_entry_point_twice_in_a_row(torch_cuda):
    torch = torch_cuda
    n = 2 * BLOCK
    first, second = _capture(9001, n), _capture(9002, n)
    want_first, want_second = _allocating_entry_point(first), _allocating_entry_point(second)
    assert not np.array_equal(want_first, want_second)            # the two calls are told apart

    got_first, _ = _static_entry_point(torch, first, n)
    got_second, _ = _static_entry_point(torch, second, n)         # finds the first call's bytes
    bits_off = _same(got_first, want_first) + _same(got_second, want_second)

    # a shorter call after a longer one: what the longer call left past its end is not read
    want_tail = _allocating_entry_point(np.concatenate([np.full(2 * BLOCK, 127, np.uint8), first[: 2 * BLOCK]]))
    got_short, _ = _static_entry_point(torch, first[: 2 * BLOCK], BLOCK)
    # (the allocating tap sees these samples at another offset inside the kernel's tiles, so
    #  rounding may differ in the last place; the first ten outputs still see its history)
    _same(got_short[16:], want_tail[BLOCK + 16:])

    # and once more at full size, into the same output buffer
    again, _ = _static_entry_point(torch, first, n)
    bits_off += _same(again, want_first)
    print(f"davis_disc32_static vs davis_debug_disc32: {bits_off} of {3 * n} values differ at all")


def test_static_scratch_refuses_what_its_array_cannot_hold(torch_cuda):
    torch = torch_cuda
    from rtldavis_b200 import _lib
    limit = 262144                                                # DAVIS_STATIC_MAX_SAMPLES
    out = torch.zeros(16, dtype=torch.float32, device="cuda")
    iq = np.full(32, 127, np.uint8)
    _, text = _static_entry_point(torch, iq, limit + 8, out=out, expect=_lib.E_SIZE)
    assert "DAVIS_STATIC_MAX_SAMPLES" in text and str(limit) in text
    _, text = _static_entry_point(torch, iq, 12, out=out, expect=_lib.E_SIZE)
    assert "multiple of 8" in text
    assert float(out.abs().sum()) == 0.0                          # nothing was run
    # the largest size it accepts runs and matches the tail the allocating tap reports
    big = _capture(9003, limit)
    got, _ = _static_entry_point(torch, big, limit)
    assert np.all(np.isfinite(got))
    head = _allocating_entry_point(big[: 4 * BLOCK])
    _same(got[: 2 * BLOCK], head)
