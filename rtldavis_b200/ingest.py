"""Feeding the batched demodulator from several sources at once.

The reference runs one process per SDR: ``runners/rtlsdr.py:100-103`` reads
``block_size`` samples at a time and ``worker.py:34-58`` hands every block to
``Parser.demodulator.demodulate``.  ``FanIn`` is that loop for many sources: it reads
one chunk of ``n_blocks`` blocks from every source (files or pipes in the ``rtl_sdr``
format -- interleaved uint8 I,Q -- or any object with ``read(nbytes)``), places them
in one pinned batch buffer and makes one ``BatchDemodulator.submit`` per chunk.  A
source that cannot fill a whole block is finished: like the reference
(``ByteToCmplxLUT.execute``, dsp.py:32-36, rejects a short block) its tail is not
demodulated.  Streams that ended are fed mid-scale bytes (127/128: no signal) so the
others keep their stream index.
"""
from __future__ import annotations

import ctypes as C
from typing import BinaryIO, Iterator, List, Optional, Sequence, Tuple, Union

import numpy as np

Source = Union[str, BinaryIO]


def read_exact(f, nbytes: int) -> bytes:
    """Up to ``nbytes`` from a file or pipe (short only at end of input)."""
    parts, got = [], 0
    while got < nbytes:
        b = f.read(nbytes - got)
        if not b:
            break
        parts.append(b)
        got += len(b)
    return b"".join(parts)


def iter_blocks(f, block_size: int, n_blocks: int = 1) -> Iterator[np.ndarray]:
    """uint8 chunks of up to ``n_blocks`` whole blocks (2*block_size bytes each)."""
    step = 2 * block_size
    while True:
        raw = read_exact(f, step * n_blocks)
        whole = len(raw) // step
        if whole == 0:
            return
        yield np.frombuffer(raw, dtype=np.uint8, count=whole * step)
        if whole < n_blocks:
            return


class FanIn:
    """Drives a ``BatchDemodulator`` from ``len(sources)`` byte sources."""

    IDLE = 127          # byte fed for a stream that has ended

    def __init__(self, demod, sources: Sequence[Source], n_blocks: int = 8, prefetch: bool = False):
        """``prefetch``: read chunk i+1 from the sources while the GPU works on chunk i
        (two pinned buffers).  For files; on a live pipe it would hold the packets of a
        chunk back until the next chunk has arrived."""
        if len(sources) != demod.n_streams:
            raise ValueError("one source per stream")
        if n_blocks < 1 or n_blocks > demod.max_blocks:
            raise ValueError("n_blocks must be in 1..max_blocks")
        self.demod = demod
        self.n_blocks = n_blocks
        self.block_bytes = 2 * demod.cfg.block_size
        self._files = [open(s, "rb") if isinstance(s, str) else s for s in sources]
        self._own = [isinstance(s, str) for s in sources]
        self.alive = [True] * len(sources)
        self.blocks_read = [0] * len(sources)
        self.prefetch = bool(prefetch)
        self._pins: list = []
        shape = (len(sources), n_blocks * self.block_bytes)
        self._batches = [self._pinned(shape) for _ in range(2 if self.prefetch else 1)]

    def _pinned(self, shape: Tuple[int, int]) -> np.ndarray:
        """Page-locked host buffer from the library (falls back to pageable memory
        only if the allocation is refused; the library stages either)."""
        from . import _lib
        L = _lib.load()
        n = int(shape[0]) * int(shape[1])
        p = C.c_void_p()
        if L.davis_host_alloc(C.byref(p), n) == 0 and p.value:
            self._pins.append((L, p))
            buf = (C.c_uint8 * n).from_address(p.value)
            return np.frombuffer(buf, dtype=np.uint8).reshape(shape)
        return np.empty(shape, dtype=np.uint8)

    def fill(self, slot: int = 0) -> int:
        """Reads the next chunk of every live source into buffer ``slot``; returns how many
        blocks the batch holds (the minimum any live source delivered; 0 = all ended)."""
        step = self.block_bytes
        got: List[Optional[np.ndarray]] = []
        usable = self.n_blocks
        for i, f in enumerate(self._files):
            if not self.alive[i]:
                got.append(None)
                continue
            raw = read_exact(f, step * self.n_blocks)
            whole = len(raw) // step
            if whole < self.n_blocks:
                self.alive[i] = False
            if whole == 0:
                got.append(None)
                continue
            got.append(np.frombuffer(raw, dtype=np.uint8, count=whole * step))
            usable = min(usable, whole)
        if all(g is None for g in got):
            return 0
        # every stream advances by the same number of blocks per call
        for i, g in enumerate(got):
            row = self._batches[slot][i, : usable * step]
            if g is None:
                row[:] = self.IDLE
            else:
                row[:] = g[: usable * step]
                self.blocks_read[i] += usable
                if g.size > usable * step:
                    self._pushback(i, g[usable * step:])
        return usable

    def _pushback(self, i: int, rest: np.ndarray) -> None:
        """A source delivered more whole blocks than the batch could take (another
        source ran short): keep them for the next chunk."""
        import io
        f = self._files[i]
        self._files[i] = _Chain(io.BytesIO(rest.tobytes()), f)
        self.alive[i] = True

    def run(self) -> Iterator[np.ndarray]:
        """Yields the packet records of every chunk until all sources end."""
        slot = 0
        n = self.fill(slot)
        while n:
            self.demod.submit(self._batches[slot][:, : n * self.block_bytes])
            if self.prefetch:
                # the other buffer: this one is being copied to the device until fetch() returns
                slot ^= 1
                nxt = self.fill(slot)
                yield self.demod.fetch()
                n = nxt
            else:
                yield self.demod.fetch()
                n = self.fill(slot)

    def close(self) -> None:
        for f, own in zip(self._files, self._own):
            if own:
                try:
                    f.close()
                except Exception:
                    pass
        self._batches = []
        while self._pins:
            L, p = self._pins.pop()
            L.davis_host_free(p)


class _Chain:
    """read() from ``first`` until it is empty, then from ``second``."""

    def __init__(self, first, second):
        self.first, self.second = first, second

    def read(self, n: int) -> bytes:
        b = self.first.read(n) if self.first is not None else b""
        if len(b) < n:
            self.first = None
            b += self.second.read(n - len(b))
        return b

    def close(self) -> None:
        if hasattr(self.second, "close"):
            self.second.close()
