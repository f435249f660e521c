"""The drop-in boundary without a GPU: libdavis_b200.so loads, exports every entry point
include/davis_b200.h declares, and the product path fails loudly (no CPU fallback) when
there is no CUDA device.  No compute call is made here."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT

HEADER = os.path.join(ROOT, "include", "davis_b200.h")


def _declared():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"DAVIS_API\s+[\w\s\*]*?\b(davis_\w+)\s*\(", text)))


def test_header_declares_the_boundary():
    names = _declared()
    for must in ("davis_create", "davis_destroy", "davis_reset", "davis_demod_u8", "davis_demod_c64",
                 "davis_demod_c128", "davis_fetch_hits", "davis_fetch_hits_view", "davis_last_error"):
        assert must in names


def test_library_loads_and_exports_every_declared_symbol():
    from rtldavis_b200 import _lib
    L = _lib.load()
    missing = [n for n in _declared() if not hasattr(L, n)]
    assert not missing, f"declared in include/davis_b200.h but not exported: {missing}"
    L.davis_abi_version.restype = C.c_int
    assert L.davis_abi_version() == 3          # 2: davis_cfg.flavor; 3: davis_stats.reruns
    L.davis_strerror.restype = C.c_char_p
    assert L.davis_strerror(0) is not None


def test_struct_layouts_match_the_python_mirror():
    from rtldavis_b200 import dsp
    # davis_hit: int64, 3 x 32-bit, 10 + 2 bytes, 3 doubles = 56 bytes (include/davis_b200.h)
    assert dsp.HIT_DTYPE.itemsize == 56
    assert dsp.HIT_DTYPE.fields["data"][1] == 20 and dsp.HIT_DTYPE.fields["rssi"][1] == 32


def test_davis_cfg_layout_and_argument_checks_need_no_gpu():
    """davis_create validates the configuration before it touches CUDA: a flavor (or a
    symbol length) the library does not implement is DAVIS_E_INVAL on any machine."""
    from rtldavis_b200 import _lib
    L = _lib.load()
    assert C.sizeof(_lib.DavisCfg) == 44                       # eleven 32-bit fields
    assert _lib.DavisCfg.flavor.offset == 40
    h = C.c_void_p()
    bad_flavor = _lib.DavisCfg(19200, 14, 16, 80, 0xCB89, 8192, 1, 0, 0, 0, 7)
    assert L.davis_create(C.byref(bad_flavor), 1, 0, C.byref(h)) == _lib.E_INVAL and not h.value
    bad_sps = _lib.DavisCfg(19200, 13, 16, 80, 0xCB89, 8192, 1, 0, 0, 0, 0)
    assert L.davis_create(C.byref(bad_sps), 1, 0, C.byref(h)) == _lib.E_INVAL
    assert L.davis_create(None, 1, 0, C.byref(h)) == _lib.E_INVAL


def _no_gpu():
    try:
        import torch
        return not torch.cuda.is_available()
    except Exception:
        return True


@pytest.mark.skipif(not _no_gpu(), reason="needs a machine without a CUDA device")
def test_no_cpu_fallback_without_a_gpu():
    from rtldavis_b200 import dsp, protocol
    # the single-stream mirror creates its device handle at the first call (the reference builds
    # a Demodulator in a parent process that must stay free of CUDA before it forks its worker)
    dem = dsp.Demodulator(protocol.new_packet_config(14))
    with pytest.raises(Exception) as e:
        dem.demodulate(np.full(2 * 8192, 127, dtype=np.uint8))
    assert not isinstance(e.value, (AttributeError, TypeError, NameError))
    with pytest.raises(Exception):
        dem.discriminated
    with pytest.raises(Exception):
        dsp.BatchDemodulator(protocol.new_packet_config(14), 4)


def test_product_package_does_not_import_the_oracle():
    pkg = os.path.join(ROOT, "rtldavis_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M), f
