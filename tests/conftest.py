import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))


TEN_SECOND_FIXTURES = ["config1_10s", "snr6_10s", "noise_10s", "rf50ppm_10s"]


def load_seeded(name):
    """(fixture, iq) of a ``*_10s`` fixture: the IQ is rebuilt from the stored generator recipe
    and must hash to what the reference was fed (oracle/gen_golden.py:save_seeded)."""
    import hashlib
    import json

    from rtldavis_b200 import synth
    g = load_golden(name)
    cap = synth.make_capture(**json.loads(str(g["recipe"])))
    assert hashlib.sha256(cap.iq.tobytes()).hexdigest() == str(g["sha256"]), \
        "the generator no longer reproduces the bytes the reference was run on"
    B, n = int(g["block"]), int(g["n_blocks"])
    return g, cap.iq[: 2 * B * n]


def golden_rows(g):
    """(call, index, data, crc_ok) tuples of a fixture, in the reference's order."""
    return [(int(c), int(i), bytes(d.tolist()), bool(k))
            for c, i, d, k in zip(g["pk_call"], g["pk_index"], g["pk_data"], g["pk_crc"])]


@pytest.fixture(scope="session")
def golden_names():
    return ["hi_snr_8192", "lo_snr_8192", "go_block_512", "complex_in_8192"]
