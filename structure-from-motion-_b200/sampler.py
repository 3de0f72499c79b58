"""RANSAC sample streams drawn from CPython's GLOBAL `random` generator, exactly as the
reference consumes it: one random.sample(range(n), min(k, n)) per hypothesis
(GetInliersRANSAC.py:55, PnPRANSAC.py:49). The fast path replays the Mersenne Twister in C
(csrc/mt_sample.cpp) and writes the advanced state back, so code that interleaves on the same
stream (the wrapper's homography RANSAC, GetHomographyInliers.py:126) sees identical draws."""
from __future__ import annotations

import ctypes as C
import random

import numpy as np

from . import _lib


def _python_stream(n: int, k: int, count: int) -> np.ndarray:
    out = np.empty((count, k), dtype=np.int32)
    for i in range(count):
        out[i] = random.sample(range(n), k)
    return out


def sample_stream(n: int, k: int, count: int, out: np.ndarray | None = None) -> np.ndarray:
    """[count, min(k,n)] int32 index sets; advances the global `random` state like the reference.
    `out` (optional): a C-contiguous int32 [count, min(k,n)] array to fill (e.g. pinned memory)."""
    k = min(k, n)
    if count <= 0 or k <= 0:
        return np.empty((max(count, 0), max(k, 0)), dtype=np.int32)
    if out is not None and (out.shape != (count, k) or out.dtype != np.int32 or not out.flags.c_contiguous):
        raise ValueError("out must be a C-contiguous int32 array of shape (count, min(k, n))")
    inst = getattr(random, "_inst", None)
    fast = (type(inst) is random.Random and getattr(random.sample, "__self__", None) is inst
            and n < 2 ** 31 and k <= 64 and count >= 16)
    if fast:
        st = random.getstate()
        fast = st[0] == 3 and len(st[1]) == 625
    if not fast:
        res = _python_stream(n, k, count)
        if out is not None:
            out[...] = res
            return out
        return res
    mt = np.array(st[1][:624], dtype=np.uint32)
    idx = C.c_int32(st[1][624])
    if out is None:
        out = np.empty((count, k), dtype=np.int32)
    rc = _lib.load().sfm_mt_sample_stream(mt.ctypes.data, C.addressof(idx), n, k, count, out.ctypes.data)
    _lib.check(rc, "sfm_mt_sample_stream")
    random.setstate((3, tuple(mt.tolist()) + (int(idx.value),), st[2]))
    return out


class StreamReplayer:
    """Chunked replay of the same stream: the generator state is read once, advanced in C chunk by chunk
    (draw), and written back once (finish) -- the per-call getstate/setstate of sample_stream costs
    ~0.1 ms, which matters when the chunks are pipelined against GPU work. Falls back to
    random.sample itself when the global generator is not the stock Mersenne Twister."""

    def __init__(self, n: int, k: int):
        self.n, self.k = int(n), min(int(k), int(n))
        inst = getattr(random, "_inst", None)
        self.fast = (type(inst) is random.Random and getattr(random.sample, "__self__", None) is inst
                     and self.n < 2 ** 31 and 0 < self.k <= 64)
        if self.fast:
            st = random.getstate()
            self.fast = st[0] == 3 and len(st[1]) == 625
        if self.fast:
            self._gauss = st[2]
            self._mt = np.array(st[1][:624], dtype=np.uint32)
            self._idx = C.c_int32(st[1][624])
            self._fn = _lib.load().sfm_mt_sample_stream

    def draw(self, count: int, out: np.ndarray, n: int | None = None) -> np.ndarray:
        """Fill out [count, k] (C-contiguous int32) with the next `count` index sets. n: draw from range(n) instead
        of the replayer's own population (consecutive registrations of different sizes on ONE stream; n >= k)."""
        n = self.n if n is None else int(n)
        if n < self.k or n >= 2 ** 31:
            raise ValueError("StreamReplayer.draw: the population must hold at least k and fewer than 2^31 elements")
        if out.shape != (count, self.k) or out.dtype != np.int32 or not out.flags.c_contiguous:
            raise ValueError("out must be a C-contiguous int32 array of shape (count, min(k, n))")
        if count <= 0:
            return out
        if not self.fast:
            out[...] = _python_stream(n, self.k, count)
            return out
        rc = self._fn(self._mt.ctypes.data, C.addressof(self._idx), n, self.k, count, out.ctypes.data)
        _lib.check(rc, "sfm_mt_sample_stream")
        return out

    def finish(self) -> None:
        """Write the advanced state back into the global generator (call exactly once)."""
        if self.fast:
            random.setstate((3, tuple(self._mt.tolist()) + (int(self._idx.value),), self._gauss))
            self.fast = False
