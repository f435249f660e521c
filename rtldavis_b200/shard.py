"""Sharding of independent IQ streams over the GPUs of one box.

Streams never exchange data, so the demodulation path needs no collective: every
rank (one process per GPU) owns a contiguous range of streams and its own
``BatchDemodulator``.  The only communication is the host-side gather of the packet
records, a few hundred bytes per stream, done here with ``torch.distributed``
(NCCL ranks gather through CPU tensors on a gloo group or with ``gather_object``).
"""
from __future__ import annotations

import os
from typing import List, Optional, Tuple

import numpy as np

from ._lib import HIT_DTYPE


def partition(n_streams: int, world: int) -> List[Tuple[int, int]]:
    """Contiguous ``(first_stream, count)`` per rank; the first ``n % world`` ranks
    get one stream more.  Ranks beyond the stream count get ``(n_streams, 0)``."""
    if n_streams < 0 or world < 1:
        raise ValueError("n_streams >= 0 and world >= 1 required")
    base, extra = divmod(n_streams, world)
    out, at = [], 0
    for r in range(world):
        cnt = base + (1 if r < extra else 0)
        out.append((at, cnt))
        at += cnt
    return out


def rank_world() -> Tuple[int, int, int]:
    """(rank, world, local_rank) from the torchrun environment (1 process if unset)."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")),
            int(os.environ.get("LOCAL_RANK", "0")))


def globalize(hits: np.ndarray, first_stream: int) -> np.ndarray:
    """Copy of a rank's records with ``stream`` renumbered into the global range."""
    out = np.array(hits, dtype=HIT_DTYPE, copy=True)
    out["stream"] += np.uint32(first_stream)
    return out


def gather_packets(hits: np.ndarray, first_stream: int, dst: int = 0, group=None) -> Optional[np.ndarray]:
    """Host-side gather of every rank's packet records on rank ``dst``.

    Each rank passes the records of its own streams (local numbering) and the
    global index of its first stream.  Rank ``dst`` receives one array ordered by
    global stream, then call, then the reference's per-call order -- i.e. exactly
    what a single demodulator over all streams would have returned; other ranks
    get ``None``.  Without an initialised process group this is the identity."""
    import torch.distributed as dist

    mine = globalize(hits, first_stream)
    if not (dist.is_available() and dist.is_initialized()):
        return mine
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    payload = (int(first_stream), mine.tobytes())
    gathered = [None] * world if rank == dst else None
    dist.gather_object(payload, gathered, dst=dst, group=group)
    if rank != dst:
        return None
    parts = sorted(gathered, key=lambda p: p[0])           # ranges are disjoint and contiguous
    arrays = [np.frombuffer(b, dtype=HIT_DTYPE) for _, b in parts]
    return np.concatenate(arrays) if arrays else np.empty(0, dtype=HIT_DTYPE)


class ShardedDemodulator:
    """The rank-local part of a batch of ``n_streams`` streams sharded over the
    ranks of the current process group.  ``submit`` takes this rank's rows only
    (``uint8[count, n_blocks*2*block_size]``); ``fetch_gathered`` returns the whole
    batch's packets on rank 0."""

    def __init__(self, cfg, n_streams: int, max_blocks: int = 1, device: Optional[int] = None, **kw):
        from . import dsp
        rank, world, local = rank_world()
        self.rank, self.world = rank, world
        self.first_stream, self.count = partition(n_streams, world)[rank]
        self.n_streams = n_streams
        self.local = None
        if self.count:
            self.local = dsp.BatchDemodulator(cfg, self.count, device=local if device is None else device,
                                              max_blocks=max_blocks, **kw)

    def submit(self, iq_rows, cuda_stream=None) -> None:
        if self.local is not None:
            self.local.submit(iq_rows, cuda_stream=cuda_stream)

    def fetch(self) -> np.ndarray:
        if self.local is None:
            return np.empty(0, dtype=HIT_DTYPE)
        return self.local.fetch(copy=False)

    def fetch_gathered(self, dst: int = 0, group=None) -> Optional[np.ndarray]:
        return gather_packets(self.fetch(), self.first_stream, dst=dst, group=group)

    def reset(self) -> None:
        if self.local is not None:
            self.local.reset()

    def close(self) -> None:
        if self.local is not None:
            self.local.close()
