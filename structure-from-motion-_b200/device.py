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
