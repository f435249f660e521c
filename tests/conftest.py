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


def golden_rows(g):
    """(call, index, data, crc_ok) tuples of a fixture, in the reference's order."""
    return [(int(c), int(i), bytes(d.tolist()), bool(k))
            for c, i, d, k in zip(g["pk_call"], g["pk_index"], g["pk_data"], g["pk_crc"])]


@pytest.fixture(scope="session")
def golden_names():
    return ["hi_snr_8192", "lo_snr_8192", "go_block_512", "complex_in_8192"]
