"""N GPUs, one global batch, gathered records == oracle (needs >= 2 CUDA devices; the same code
path at N = 1 is every other GPU test, and bench.py's parity_check covers N = 1, 2, 4, 8)."""
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu


def test_sharded_batch_over_all_visible_gpus_equals_the_oracle():
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least two GPUs")
    n = 8 if n >= 8 else 4 if n >= 4 else 2
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    here = os.path.dirname(os.path.abspath(__file__))
    res = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
                          "--master-addr", "127.0.0.1", "--master-port", str(port),
                          os.path.join(here, "sharded_worker.py")],
                         capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stderr[-2000:]
    assert "SHARDED OK" in res.stdout, res.stdout[-2000:] + res.stderr[-2000:]
