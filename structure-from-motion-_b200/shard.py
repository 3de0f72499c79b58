"""Independent problems over several GPUs (second row): the per-point and per-registration solvers
-- NonLinearTriangulation.py:53-130, NonlinearPnP.py:47-123, LinearTriangulation -- have no coupling between
problems, so rank r of `world` takes a contiguous chunk, runs the single-GPU path on it, and the chunks are
all-gathered: "replicas of the kernel", no data-path collective. The reference is single-process; nothing here
has a counterpart there. One process per GPU (torch.distributed, NCCL on GPUs, gloo in the CPU tests of the
plumbing)."""
from __future__ import annotations

import numpy as np


def shard_range(n: int, rank: int, world: int):
    """Contiguous chunk [lo, hi) of 0..n-1 for `rank`; sizes differ by at most one, earlier ranks take the extra."""
    if world < 1 or not 0 <= rank < world or n < 0:
        raise ValueError("shard_range: need 0 <= rank < world and n >= 0")
    q, r = divmod(n, world)
    lo = rank * q + min(rank, r)
    return lo, lo + q + (1 if rank < r else 0)


def _dist():
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return None
    return dist


def map_sharded(fn, n: int, arrays, rank: int | None = None, world: int | None = None) -> np.ndarray:
    """fn(*[a[lo:hi] for a in arrays]) -> [hi - lo, ...] float64 array, evaluated on this rank's chunk; returns the
    concatenation over all ranks (every rank gets it). Without an initialised process group it is fn on everything."""
    import torch
    dist = _dist()
    if dist is None:
        return np.asarray(fn(*arrays), np.float64)
    rank = dist.get_rank() if rank is None else rank
    world = dist.get_world_size() if world is None else world
    lo, hi = shard_range(n, rank, world)
    local = np.ascontiguousarray(fn(*[a[lo:hi] for a in arrays]), np.float64) if hi > lo else None
    # the trailing shape is the same on every rank; an empty chunk learns it from the others
    shp = torch.tensor(list(local.shape[1:]) if local is not None else [], dtype=torch.int64)
    dims = torch.tensor([shp.numel()], dtype=torch.int64)
    dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")
    dims = dims.to(dev)
    dist.all_reduce(dims, op=dist.ReduceOp.MAX)
    tail = torch.zeros(int(dims.item()), dtype=torch.int64)
    tail[:shp.numel()] = shp
    tail = tail.to(dev)
    dist.all_reduce(tail, op=dist.ReduceOp.MAX)
    tail = tuple(int(v) for v in tail.cpu())
    width = int(np.prod(tail)) if tail else 1
    rows = max(shard_range(n, r, world)[1] - shard_range(n, r, world)[0] for r in range(world))
    buf = torch.zeros((rows, width), dtype=torch.float64)
    if local is not None:
        buf[:hi - lo] = torch.from_numpy(local.reshape(hi - lo, width))
    buf = buf.to(dev)
    parts = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(parts, buf)
    out = np.empty((n,) + tail, np.float64)
    for r in range(world):
        a, b = shard_range(n, r, world)
        out[a:b] = parts[r][:b - a].cpu().numpy().reshape((b - a,) + tail)
    return out


def nonlinear_triangulation_sharded(K, C1, R1, C2, R2, x1, x2, x0):
    """NonLinearTriangulation (NonLinearTriangulation.py:53-130) of N points split over the ranks -> X [N,3]."""
    from . import refine
    x1, x2, x0 = (np.ascontiguousarray(a, np.float64) for a in (x1, x2, x0))
    return map_sharded(lambda a, b, c: refine.NonLinearTriangulation(K, C1, R1, C2, R2, a, b, c), len(x0), (x1, x2, x0))


def linear_triangulation_sharded(K, C1, R1, C2, R2, x1, x2):
    """LinearTriangulation of N correspondences split over the ranks -> X [N,3]."""
    from . import refine
    x1, x2 = (np.ascontiguousarray(a, np.float64) for a in (x1, x2))
    return map_sharded(lambda a, b: refine.LinearTriangulation(K, C1, R1, C2, R2, a, b), len(x1), (x1, x2))


def nonlinear_pnp_sharded(X_list, x_list, K, C_list, R_list):
    """NonLinearPnP (NonlinearPnP.py:47-123) of B registrations split over the ranks -> (C [B,3], R [B,3,3])."""
    from . import refine
    B = len(X_list)
    idx = np.arange(B)

    def solve(ids):
        ids = [int(i) for i in ids]
        C, R, _ = refine.nonlinear_pnp_batch([X_list[i] for i in ids], [x_list[i] for i in ids], K,
                                             [C_list[i] for i in ids], [R_list[i] for i in ids])
        return np.hstack([np.asarray(C, np.float64).reshape(len(ids), 3), np.asarray(R, np.float64).reshape(len(ids), 9)])
    out = map_sharded(solve, B, (idx,))
    return out[:, :3].copy(), out[:, 3:].reshape(B, 3, 3).copy()
