"""The drop-in inside rtldavis's real process layout: Parser in the parent, worker forked.

runners/rtlsdr.py:30 builds a Parser (hence a Demodulator) in the parent, rtlsdr.py:61-65 forks
``worker_main`` (start method "fork", __main__.py:277), worker.py:28 builds another in the child.
If the parent's Demodulator created a CUDA context, the child's first CUDA call would fail
("Failed to initialize worker").  The script runs in a fresh interpreter because this pytest
process has long since initialised CUDA itself.  tests/fork_runner.py has the details.
"""
import os
import subprocess
import sys

import pytest

from conftest import ROOT
from oracle import ref_loader

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("mode", ["u8", "c128"])
def test_parent_parser_then_forked_worker_matches_the_reference(mode):
    if not ref_loader.available():
        pytest.skip("oracle/_ref not staged (python oracle/stage_ref.py in the build container)")
    res = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "fork_runner.py"), mode],
                         capture_output=True, text=True, timeout=600, cwd=ROOT)
    print(res.stdout[-2000:], res.stderr[-4000:])
    assert res.returncode == 0, res.stderr[-4000:]
    assert "40 messages from the forked worker" in res.stdout
