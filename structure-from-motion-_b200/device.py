"""Device plumbing: torch is used for device memory, streams and H2D/D2H copies only."""
from __future__ import annotations

import numpy as np
import torch

from ._lib import SfmError


def require_cuda() -> torch.device:
    if not torch.cuda.is_available():
        raise SfmError("a CUDA device is required: libsfm_b200 has no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


def to_device(a, dtype=np.float64) -> torch.Tensor:
    """Contiguous host array -> device tensor (one H2D copy on the current stream)."""
    if isinstance(a, torch.Tensor):
        return a.to(device=require_cuda(), non_blocking=True).contiguous()
    a = np.ascontiguousarray(a, dtype=dtype)
    return torch.from_numpy(a).to(require_cuda(), non_blocking=True)


def empty(shape, dtype) -> torch.Tensor:
    return torch.empty(shape, dtype=dtype, device=require_cuda())


def ptr(t) -> int:
    return 0 if t is None else int(t.data_ptr())


def stream() -> int:
    return int(torch.cuda.current_stream().cuda_stream)


# ---- large host arrays: staged and compared in slices on a few threads (numpy releases the GIL for both) ----
_POOL = None
PAR_MIN_BYTES = 4 << 20
PAR_THREADS = 4


def _pool():
    global _POOL
    if _POOL is None:
        from concurrent.futures import ThreadPoolExecutor
        _POOL = ThreadPoolExecutor(max_workers=PAR_THREADS, thread_name_prefix="sfm-stage")
    return _POOL


def _slices(n: int, parts: int):
    step = -(-n // parts)
    return [(i, min(n, i + step)) for i in range(0, n, step)]


def upload_via_pinned(dst: torch.Tensor, src: np.ndarray, pinned: torch.Tensor) -> None:
    """dst (device) <- src (host, same shape) through `pinned`: small arrays in one piece; large ones in slices whose
    host copies run on a few threads, each slice's asynchronous upload issued as soon as it is staged (a 64 MB
    observation array: ~3 ms instead of ~10). All uploads go to the caller's current stream."""
    h = pinned.numpy()
    n = len(src)
    if src.nbytes < PAR_MIN_BYTES or n < 2 * PAR_THREADS:
        h[...] = src
        dst.copy_(pinned, non_blocking=True)
        return
    parts = _slices(n, 2 * PAR_THREADS)
    futs = [_pool().submit(np.copyto, h[a:b], src[a:b]) for a, b in parts]
    for (a, b), f in zip(parts, futs):
        f.result()
        dst[a:b].copy_(pinned[a:b], non_blocking=True)


def arrays_equal(a: np.ndarray, b: np.ndarray) -> bool:
    """np.array_equal for two contiguous arrays of one dtype and shape, sliced over a few threads when large."""
    if a.shape != b.shape:
        return False
    if a.nbytes < PAR_MIN_BYTES:
        return bool(np.array_equal(a, b))
    parts = _slices(len(a), PAR_THREADS)
    return all(_pool().map(lambda ab: bool(np.array_equal(a[ab[0]:ab[1]], b[ab[0]:ab[1]])), parts))

