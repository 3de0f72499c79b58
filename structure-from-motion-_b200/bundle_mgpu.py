"""Point-sharded bundle adjustment across the GPUs of one box: one process per GPU,
cameras replicated, each rank owns a contiguous range of the point-sorted observation list. The
collectives run inside the library on its own NCCL communicator (sfm_ba_solve_mgpu); torch.distributed
is used only to hand the 128-byte NCCL unique id from rank 0 to the other ranks.

The reference has no multi-process path at all; its BA is one scipy call (BundleAdjustment.py:205-212)."""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib, device
from .bundle import BA_MAX_NFEV, REF_TOL, _hptr


def shard_points(pt_idx: np.ndarray, n_pts: int, world: int):
    """Split points 0..n_pts-1 (observations sorted by point) into `world` contiguous ranges balanced
    by observation count. Returns a list of (p_lo, p_hi, o_lo, o_hi); ranges may be empty when there
    are fewer points than ranks."""
    pt_idx = np.asarray(pt_idx)
    n_obs = len(pt_idx)
    if n_obs > 1 and np.any(np.diff(pt_idx) < 0):
        raise ValueError("observations must be sorted by point")
    pt_ptr = np.searchsorted(pt_idx, np.arange(n_pts + 1), side="left")      # CSR pointer, pt_ptr[n_pts] = n_obs
    cuts = [0]
    for r in range(1, world):
        target = r * n_obs / world
        p = int(np.searchsorted(pt_ptr, target, side="left"))                # first point starting at/after target
        p = min(max(p, cuts[-1]), n_pts)
        cuts.append(p)
    cuts.append(n_pts)
    return [(cuts[r], cuts[r + 1], int(pt_ptr[cuts[r]]), int(pt_ptr[cuts[r + 1]])) for r in range(world)]


def exchange_unique_id(rank: int, make_id, group=None) -> bytes:
    """Rank 0 calls make_id() -> 128 bytes; every rank returns them (torch.distributed broadcast on
    whatever backend the group uses: nccl on the GPU box, gloo in the CPU tests)."""
    import torch.distributed as dist
    backend = dist.get_backend(group)
    dev = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")
    buf = torch.zeros(128, dtype=torch.uint8, device=dev)
    if rank == 0:
        raw = make_id()
        if len(raw) != 128:
            raise ValueError("an NCCL unique id is 128 bytes")
        buf.copy_(torch.frombuffer(bytearray(raw), dtype=torch.uint8))
    dist.broadcast(buf, src=0, group=group)
    return bytes(buf.cpu().numpy().tobytes())


class NcclComm:
    """The library's own NCCL communicator (one per process)."""

    def __init__(self, rank: int, world: int, group=None):
        self.rank, self.world, self.handle = int(rank), int(world), C.c_void_p(None)
        if world == 1:
            return
        lib = _lib.load()

        def make_id():
            raw = C.create_string_buffer(128)
            _lib.check(lib.sfm_nccl_unique_id(C.addressof(raw)), "sfm_nccl_unique_id")
            return raw.raw

        uid = exchange_unique_id(rank, make_id, group)
        raw = C.create_string_buffer(uid, 128)
        _lib.check(lib.sfm_nccl_comm_create(C.addressof(raw), self.rank, self.world, C.byref(self.handle)),
                   "sfm_nccl_comm_create")

    def ptr(self):
        return self.handle.value or 0

    def allreduce_max_u64(self, t: torch.Tensor):
        _lib.check(_lib.load().sfm_allreduce_max_u64(self.ptr(), device.ptr(t), t.numel(), device.stream()),
                   "sfm_allreduce_max_u64")

    def allreduce_sum_f64(self, t: torch.Tensor):
        _lib.check(_lib.load().sfm_allreduce_sum_f64(self.ptr(), device.ptr(t), t.numel(), device.stream()),
                   "sfm_allreduce_sum_f64")

    def destroy(self):
        if self.handle.value:
            _lib.check(_lib.load().sfm_nccl_comm_destroy(self.handle), "sfm_nccl_comm_destroy")
            self.handle = C.c_void_p(None)


class BAShardedProblem:
    """This rank's share of a sparse BA problem (global arrays in, local slices on the device)."""

    def __init__(self, n_cams, n_pts, cam_idx, pt_idx, obs, K, comm: NcclComm):
        cam_idx = np.ascontiguousarray(cam_idx, np.int32)
        pt_idx = np.ascontiguousarray(pt_idx, np.int32)
        obs = np.ascontiguousarray(obs, np.float64).reshape(-1, 2)
        self.comm, self.n_cams, self.n_pts_global = comm, int(n_cams), int(n_pts)
        self.p_lo, self.p_hi, self.o_lo, self.o_hi = shard_points(pt_idx, n_pts, comm.world)[comm.rank]
        self.n_pts, self.n_obs = self.p_hi - self.p_lo, self.o_hi - self.o_lo
        if self.n_pts < 1 or self.n_obs < 1:
            raise ValueError("every rank needs at least one point with an observation")
        self.K = np.ascontiguousarray(K, np.float64).reshape(3, 3)
        sl = slice(self.o_lo, self.o_hi)
        self.d_cam_idx = device.to_device(cam_idx[sl], np.int32)
        self.d_pt_idx = device.to_device(pt_idx[sl] - self.p_lo, np.int32)
        self.d_obs = device.to_device(obs[sl])
        self.ws_bytes = int(_lib.load().sfm_ba_workspace_bytes(self.n_cams, self.n_pts, self.n_obs))
        self.ws = device.empty((self.ws_bytes,), torch.uint8)
        self._indexed = False

    def solve_device(self, d_cam, d_pts_local, max_nfev=BA_MAX_NFEV, tol=REF_TOL, lambda0=1e-4, verbose=False,
                     profile=False):
        lib = _lib.load()
        opts = _lib.BAOptions()
        lib.sfm_ba_default_options(C.addressof(opts))
        opts.max_trials, opts.verbose = int(max_nfev), int(bool(verbose) and self.comm.rank == 0)
        opts.ftol = opts.xtol = opts.gtol = float(tol)
        opts.lambda0, opts.reuse_index = float(lambda0), int(self._indexed)
        opts.reserved = 1 if profile else 0
        res = _lib.BAResult()
        rc = lib.sfm_ba_solve_mgpu(
            device.ptr(d_cam), device.ptr(d_pts_local), device.ptr(self.d_obs), device.ptr(self.d_cam_idx),
            device.ptr(self.d_pt_idx), self.n_cams, self.n_pts, self.n_obs, _hptr(self.K), C.addressof(opts),
            C.addressof(res), device.ptr(self.ws), self.ws_bytes, self.comm.ptr(), self.comm.rank, self.comm.world,
            device.stream())
        _lib.check(rc, "sfm_ba_solve_mgpu")
        self._indexed = True
        return dict(cost0=res.cost0, cost=res.cost, gnorm=res.gnorm, lambda_=res.lambda_, status=res.status,
                    n_linearisations=res.n_linearisations, n_trials=res.n_trials, n_chol_fail=res.n_chol_fail)


def ba_solve_sharded(cams, pts, cam_idx, pt_idx, obs, K, comm: NcclComm, max_nfev=BA_MAX_NFEV, tol=REF_TOL):
    """Host arrays (the GLOBAL problem on every rank) in; returns (cams [Nc,6], this rank's refined
    points, (p_lo, p_hi), info)."""
    cams = np.ascontiguousarray(cams, np.float64).reshape(-1, 6)
    pts = np.ascontiguousarray(pts, np.float64).reshape(-1, 3)
    prob = BAShardedProblem(len(cams), len(pts), cam_idx, pt_idx, obs, K, comm)
    d_cam, d_pts = device.to_device(cams), device.to_device(pts[prob.p_lo:prob.p_hi])
    info = prob.solve_device(d_cam, d_pts, max_nfev, tol)
    return d_cam.cpu().numpy(), d_pts.cpu().numpy(), (prob.p_lo, prob.p_hi), info
