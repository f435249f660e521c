"""Sharding of independent IQ streams over the GPUs of one box.

Streams never exchange data, so the demodulation path needs no collective: every
rank (one process per GPU) owns a contiguous range of streams and its own
``BatchDemodulator``.  The only communication is the host-side gather of the packet
records, a few hundred bytes per stream, done here with ``torch.distributed``
(NCCL ranks gather through CPU tensors on a gloo group or with ``gather_object``).
"""
from __future__ import annotations

import os
from typing import List, Optional, Sequence, Tuple

import numpy as np

from ._lib import HIT_DTYPE


def partition(n_streams: int, world: int, weights: Optional[Sequence[float]] = None
              ) -> List[Tuple[int, int]]:
    """Contiguous ``(first_stream, count)`` per rank.

    Without ``weights`` the split is uniform: the first ``n % world`` ranks get one stream
    more, ranks beyond the stream count get ``(n_streams, 0)``.  With ``weights`` (one
    positive number per rank, e.g. the host->device copy rate each GPU gets while all of them
    copy, ``probe_h2d_weights``) rank r gets ``n * w_r / sum(w)`` streams, rounded by largest
    remainder: the end-to-end step ends with the slowest rank, and on a box whose GPUs hang
    off unequal host links a uniform split leaves the faster links idle."""
    if n_streams < 0 or world < 1:
        raise ValueError("n_streams >= 0 and world >= 1 required")
    if weights is None:
        base, extra = divmod(n_streams, world)
        counts = [base + (1 if r < extra else 0) for r in range(world)]
    else:
        w = [float(x) for x in weights]
        if len(w) != world or any(not (x > 0.0) for x in w):
            raise ValueError("weights: one positive number per rank")
        total = sum(w)
        exact = [n_streams * x / total for x in w]
        counts = [int(e) for e in exact]
        left = n_streams - sum(counts)
        order = sorted(range(world), key=lambda r: (counts[r] - exact[r], r))
        for r in order[:left]:
            counts[r] += 1
    out, at = [], 0
    for cnt in counts:
        out.append((at, cnt))
        at += cnt
    return out


def probe_h2d_weights(device: int, seconds: float = 0.5, nbytes: int = 256 << 20) -> List[float]:
    """Pinned host -> device GB/s of every rank while ALL ranks copy at once (the situation of
    the end-to-end path), gathered on every rank: the ``weights`` of ``partition``.  Needs an
    initialised process group whose backend can all_gather CUDA tensors; one rank: ``[rate]``."""
    import time

    import torch
    import torch.distributed as dist

    multi = dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1
    h = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
    d = torch.empty(nbytes, dtype=torch.uint8, device=f"cuda:{device}")
    d.copy_(h, non_blocking=True)
    torch.cuda.synchronize(device)
    if multi:
        dist.barrier()
    t0 = time.perf_counter()
    n = 0
    while time.perf_counter() - t0 < seconds:
        d.copy_(h, non_blocking=True)
        torch.cuda.synchronize(device)
        n += 1
    rate = n * nbytes / (time.perf_counter() - t0) / 1e9
    if not multi:
        return [rate]
    mine = torch.tensor([rate], dtype=torch.float64, device=f"cuda:{device}")
    every = [torch.zeros_like(mine) for _ in range(dist.get_world_size())]
    dist.all_gather(every, mine)
    return [float(x.item()) for x in every]


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

    def __init__(self, cfg, n_streams: int, max_blocks: int = 1, device: Optional[int] = None,
                 weights: Optional[Sequence[float]] = None, **kw):
        from . import dsp
        rank, world, local = rank_world()
        self.rank, self.world = rank, world
        self.first_stream, self.count = partition(n_streams, world, weights)[rank]
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
