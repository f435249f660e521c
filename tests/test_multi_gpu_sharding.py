"""Stream sharding + host-side gather with world_size 2 over gloo on
the CPU.  The per-rank "demodulator" is the oracle (the CUDA library needs a GPU);
what is under test is the partitioning and the gather of rtldavis_b200.shard."""
import os
import socket

import numpy as np
import pytest

from rtldavis_b200 import shard
from rtldavis_b200._lib import HIT_DTYPE


def test_partition_covers_everything_once():
    for n in (0, 1, 7, 8, 1024, 65536):
        for world in (1, 2, 3, 4, 8):
            parts = shard.partition(n, world)
            assert len(parts) == world
            assert sum(c for _, c in parts) == n
            at = 0
            for first, cnt in parts:
                assert first == at
                at += cnt
            sizes = [c for _, c in parts]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard.partition(4, 0)


def test_weighted_partition_follows_the_link_rates():
    """The 8-GPU box of profiles/r01_h2d_ceiling_8gpu.txt: 23.3 GB/s to GPUs 0-3, 36.0 to 4-7."""
    w = [23.3] * 4 + [36.0] * 4
    parts = shard.partition(8192, 8, weights=w)
    counts = [c for _, c in parts]
    assert sum(counts) == 8192 and [f for f, _ in parts] == [sum(counts[:r]) for r in range(8)]
    assert counts[0] == counts[3] and counts[4] == counts[7]
    assert abs(counts[4] / counts[0] - 36.0 / 23.3) < 0.01
    # the slowest rank's time (streams / rate) drops from 1024 / 23.3 to the balanced figure
    assert max(c / r for c, r in zip(counts, w)) < 0.80 * (1024 / 23.3)
    assert shard.partition(7, 3, weights=[1, 1, 1]) == shard.partition(7, 3)
    assert [c for _, c in shard.partition(5, 2, weights=[1e-9, 1.0])] == [0, 5]
    with pytest.raises(ValueError):
        shard.partition(8, 2, weights=[1.0])
    with pytest.raises(ValueError):
        shard.partition(8, 2, weights=[1.0, 0.0])


def test_gather_is_identity_without_a_process_group():
    h = np.zeros(3, dtype=HIT_DTYPE)
    h["stream"] = [0, 1, 1]
    out = shard.gather_packets(h, 10)
    assert out["stream"].tolist() == [10, 11, 11]
    assert h["stream"].tolist() == [0, 1, 1]          # input untouched


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _oracle_hits(iq_rows):
    """Records of the C oracle in the library's record layout (local stream ids)."""
    from oracle import c_oracle as CO
    rec = CO.batch_records(iq_rows)
    out = np.zeros(len(rec), dtype=HIT_DTYPE)
    for k in ("stream", "call", "index", "data", "crc_ok"):
        out[k] = rec[k]
    return out


def _worker(rank, world, port, n_streams, n_blocks, q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank),
                      WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from rtldavis_b200 import synth
        iq, _ = synth.make_batch(n_streams, n_blocks * 8192, 2, 25.0, seed0=900)
        first, cnt = shard.partition(n_streams, world)[rank]
        assert shard.rank_world()[:2] == (rank, world)
        mine = _oracle_hits(iq[first:first + cnt]) if cnt else np.empty(0, dtype=HIT_DTYPE)
        dist.barrier()
        got = shard.gather_packets(mine, first, dst=0)
        if rank == 0:
            want = _oracle_hits(iq)
            ok = (got is not None and len(got) == len(want)
                  and all(np.array_equal(got[k], want[k]) for k in ("stream", "call", "index", "data", "crc_ok")))
            q.put((ok, len(want), int(want["crc_ok"].sum())))
        else:
            assert got is None
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_streams", [5, 2, 1])
def test_two_ranks_equal_one_demodulator(n_streams):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_streams, 4, q)) for r in range(2)]
    for p in procs:
        p.start()
    ok, n, valid = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert ok, "gathered shards differ from the single-demodulator result"
    assert n > 0 and valid > 0
