"""Host side of the F-matrix RANSAC path: mirrors GetInliersRANSAC.py / EstimateFundamentalMatrix.py
(names, argument meaning, return order, error behaviour) and drives K1..K4 through the C ABI."""
from __future__ import annotations

import numpy as np
import torch

from . import _lib, device, sampler


class FRansacDevice:
    """Device-resident outputs of one sfm_f_ransac call (all torch CUDA tensors)."""
    __slots__ = ("F", "counts", "best", "F_best", "mask", "err", "packed")


def _alloc_result(n: int, h: int, want_err: bool) -> FRansacDevice:
    r = FRansacDevice()
    r.F = device.empty((h, 9), torch.float64)
    r.counts = device.empty((h,), torch.int32)
    r.packed = device.empty((80 + n,), torch.uint8)
    r.F_best = r.packed[:72].view(torch.float64)
    r.best = r.packed[72:80].view(torch.int32)
    r.mask = r.packed[80:]
    r.err = device.empty((n,), torch.float64) if want_err else None
    return r


class FRansacPlan:
    """Preallocated buffers for repeated RANSAC runs of one shape (no allocation per step).

    Single GPU: run(d_x, d_s, thr) = sfm_f_ransac. Hypothesis-sharded multi-GPU:
    run_sharded(d_x, d_s_all, thr, rank, world): this rank scores hypotheses
    [rank*h, (rank+1)*h) of the replicated stream, one 8-byte NCCL max-allreduce of the packed
    (count, ~index) key picks the global first maximum, and every rank recomputes the winner's
    F and mask from the winning sample."""

    def __init__(self, n: int, h: int, want_err: bool = False):
        self.n, self.h = n, h
        self.r = _alloc_result(n, h, want_err)
        self.key = device.empty((1,), torch.int64)

    def run(self, d_x1x2, d_samples, threshold: float) -> FRansacDevice:
        r = self.r
        rc = _lib.load().sfm_f_ransac(
            device.ptr(d_x1x2), self.n, device.ptr(d_samples), self.h, float(threshold),
            device.ptr(r.F), device.ptr(r.counts), device.ptr(r.best), device.ptr(r.F_best),
            device.ptr(r.mask), device.ptr(r.err), device.stream())
        _lib.check(rc, "sfm_f_ransac")
        return r

    def run_sharded(self, d_x1x2, d_samples_all, threshold: float, rank: int, world: int,
                    all_reduce_max=None) -> FRansacDevice:
        lib, r, st = _lib.load(), self.r, device.stream()
        h, off = self.h, rank * self.h
        d_local = d_samples_all[off:off + h]
        _lib.check(lib.sfm_f8pt_batch(device.ptr(d_x1x2), self.n, device.ptr(d_local), h,
                                      device.ptr(r.F), st), "sfm_f8pt_batch")
        _lib.check(lib.sfm_f_score_count(device.ptr(d_x1x2), self.n, device.ptr(r.F), h,
                                         float(threshold), device.ptr(r.counts), st), "sfm_f_score_count")
        _lib.check(lib.sfm_argmax_first(device.ptr(r.counts), h, off, device.ptr(r.best),
                                        device.ptr(self.key), st), "sfm_argmax_first")
        if all_reduce_max is None:
            import torch.distributed as dist
            dist.all_reduce(self.key, op=dist.ReduceOp.MAX)       # key < 2^63: signed max is fine
        else:
            all_reduce_max(self.key)
        _lib.check(lib.sfm_f_winner_from_key(
            device.ptr(d_x1x2), self.n, device.ptr(d_samples_all), int(d_samples_all.shape[0]),
            device.ptr(self.key), float(threshold), device.ptr(r.best), device.ptr(r.F_best),
            device.ptr(r.mask), device.ptr(r.err), st), "sfm_f_winner_from_key")
        return r


def f_ransac_device(d_x1x2: torch.Tensor, d_samples: torch.Tensor, threshold: float,
                    want_err: bool = False) -> FRansacDevice:
    """K1+K2+K3+K4 on the current stream. d_x1x2 [n,4] f64, d_samples [h,8] i32 (device)."""
    lib = _lib.load()
    n, h = int(d_x1x2.shape[0]), int(d_samples.shape[0])
    r = FRansacDevice()
    r.F = device.empty((h, 9), torch.float64)
    r.counts = device.empty((h,), torch.int32)
    # one packed result buffer -> one D2H copy: [F_best 72 B | best 8 B | mask n B]
    r.packed = device.empty((80 + n,), torch.uint8)
    r.F_best = r.packed[:72].view(torch.float64)
    r.best = r.packed[72:80].view(torch.int32)
    r.mask = r.packed[80:]
    r.err = device.empty((n,), torch.float64) if want_err else None
    rc = lib.sfm_f_ransac(device.ptr(d_x1x2), n, device.ptr(d_samples), h, float(threshold),
                          device.ptr(r.F), device.ptr(r.counts), device.ptr(r.best),
                          device.ptr(r.F_best), device.ptr(r.mask), device.ptr(r.err), device.stream())
    _lib.check(rc, "sfm_f_ransac")
    return r


def f_ransac_host(points1, points2, samples, threshold, want_err=False, want_all=False):
    """Host arrays in, host results out: dict(best_idx, best_count, F_best[3,3], mask[n] bool,
    and with want_all also F[h,9], counts[h]; with want_err err[n])."""
    x1x2 = np.ascontiguousarray(np.hstack([np.asarray(points1, np.float64), np.asarray(points2, np.float64)]))
    d_x = device.to_device(x1x2)
    d_s = device.to_device(samples, np.int32)
    r = f_ransac_device(d_x, d_s, threshold, want_err)
    packed = r.packed.cpu().numpy()            # synchronises the stream
    out = dict(F_best=packed[:72].view(np.float64).reshape(3, 3).copy(),
               best_idx=int(packed[72:76].view(np.int32)[0]),
               best_count=int(packed[76:80].view(np.int32)[0]),
               mask=packed[80:].astype(bool))
    if want_all:
        out["F"] = r.F.cpu().numpy()
        out["counts"] = r.counts.cpu().numpy()
    if want_err:
        out["err"] = r.err.cpu().numpy()
    return out


_PINNED: dict = {}
STREAM_MIN_HYPOTHESES = 16384      # below this one sfm_f_ransac call is cheaper than chunking
STREAM_CHUNKS = 2
STREAM_FIRST_CHUNK = 4096
STREAM_GROWTH = 1.4


def _side_stream() -> torch.cuda.Stream:
    s = _PINNED.get("side_stream")
    if s is None:
        s = torch.cuda.Stream()
        _PINNED["side_stream"] = s
    return s


def _pinned_i32(rows: int, cols: int) -> torch.Tensor:
    """Grow-only pinned staging buffer for the sample index sets."""
    need = rows * cols
    buf = _PINNED.get("samples")
    if buf is None or buf.numel() < need:
        buf = torch.empty(need, dtype=torch.int32, pin_memory=True)
        _PINNED["samples"] = buf
    return buf[:need].view(rows, cols)


def _chunk_schedule(h: int, chunks: int = STREAM_CHUNKS) -> list:
    """Chunk sizes of f_ransac_stream: STREAM_FIRST_CHUNK, then growing by STREAM_GROWTH up to h / chunks; a last
    chunk under a third of its predecessor is merged into it (no tiny last launch). Sums to h."""
    big = max(1, -(-h // max(1, int(chunks))))
    sizes, c = [], min(big, STREAM_FIRST_CHUNK)
    while sum(sizes) < h:
        sizes.append(min(c, h - sum(sizes)))
        c = min(big, max(c + 1, int(c * STREAM_GROWTH)))
    if len(sizes) > 1 and sizes[-1] < sizes[-2] // 3:
        last = sizes.pop()
        sizes[-1] += last
    return sizes


def _stream_buffers(n: int, h: int):
    """Device buffers of f_ransac_stream, kept for the next call of the same shape on the same device (every
    result leaves as a host copy, so nothing the caller holds aliases them)."""
    k = (n, h, torch.cuda.current_device())
    hit = _PINNED.get("stream_buffers")
    if hit is None or hit[0] != k:
        hit = (k, _alloc_result(n, h, False), device.empty((1,), torch.int64), device.empty((h, 8), torch.int32),
               device.empty((2 * n, 2), torch.float64))
        _PINNED["stream_buffers"] = hit
    return hit[1:]


def _pinned_f64(rows: int, cols: int) -> torch.Tensor:
    """Grow-only pinned staging buffer for the correspondences (a call ends with a synchronising read, so the
    buffer is free again when the next call fills it)."""
    need = rows * cols
    buf = _PINNED.get("points")
    if buf is None or buf.numel() < need:
        buf = torch.empty(need, dtype=torch.float64, pin_memory=True)
        _PINNED["points"] = buf
    return buf[:need].view(rows, cols)


def f_ransac_stream(points1, points2, threshold, n_max, chunks=STREAM_CHUNKS):
    """GetInliersRANSAC's hot loop with the host work overlapped: the reference's RNG stream is
    replayed chunk by chunk into pinned memory while the GPU generates and scores the previous
    chunk (K1 + K2 per chunk, one K3 over all counts, K4 with the winner's row of F). The draws, their order and every result are the
    same as one sfm_f_ransac call over the whole stream. Returns the dict of f_ransac_host."""
    lib = _lib.load()
    p1, p2 = np.asarray(points1), np.asarray(points2)
    n, h = len(p1), int(n_max)
    # [n,4] rows (x1,y1,x2,y2). Both point sets go through ONE pinned buffer (two contiguous host copies, one
    # asynchronous upload that overlaps the replay of the first chunk; a pageable source blocks the host for
    # ~0.7 ms at 50k correspondences) and are interleaved on the device
    pin_x = _pinned_f64(2 * n, 2)
    hx = pin_x.numpy()
    np.copyto(hx[:n], p1, casting="unsafe")           # one pass whatever the caller's dtype and strides
    np.copyto(hx[n:], p2, casting="unsafe")
    st = device.stream()
    r, key, d_s, d_pp = _stream_buffers(n, h)
    d_pp.copy_(pin_x, non_blocking=True)
    d_x = torch.cat([d_pp[:n], d_pp[n:]], dim=1)
    pinned = _pinned_i32(h, 8)
    host = pinned.numpy()
    # chunk schedule: a small first chunk (the only one the GPU waits for), then growing by STREAM_GROWTH per
    # chunk -- the replay runs ~1.4x faster than K2 scores (18-22 vs 28 ns per hypothesis on the B200 box), so the
    # CPU stays ahead at a growth of 1.4 -- up to h / chunks (tools/fransac_e2e_trace.py prints the timeline)
    sizes = _chunk_schedule(h, chunks)
    rep = sampler.StreamReplayer(n, 8)
    # the sample upload and K1 of chunk i+1 go to a side stream, so they overlap the tail of K2 of chunk i
    main = torch.cuda.current_stream()
    side = _side_stream()
    side.wait_stream(main)                             # d_x, d_s, r.F were allocated / written on the main stream
    off = 0
    try:
        for c in sizes:
            rep.draw(c, host[off:off + c])
            with torch.cuda.stream(side):
                d_s[off:off + c].copy_(pinned[off:off + c], non_blocking=True)
                _lib.check(lib.sfm_f8pt_batch(device.ptr(d_x), n, device.ptr(d_s) + 32 * off, c,
                                              device.ptr(r.F) + 72 * off, side.cuda_stream), "sfm_f8pt_batch")
            main.wait_stream(side)
            _lib.check(lib.sfm_f_score_count(device.ptr(d_x), n, device.ptr(r.F) + 72 * off, c, float(threshold),
                                             device.ptr(r.counts) + 4 * off, st), "sfm_f_score_count")
            off += c
    finally:
        main.wait_stream(side)
        if off < h:                                    # an error mid-way: still consume the whole stream
            rep.draw(h - off, host[off:h])
        rep.finish()
    _lib.check(lib.sfm_argmax_first(device.ptr(r.counts), h, 0, device.ptr(r.best), device.ptr(key), st),
               "sfm_argmax_first")
    # F_best = F[winner] without a host round trip (a device-side row gather), then K4 with it
    r.F_best.copy_(r.F.index_select(0, r.best[:1].long()).view(9))
    _lib.check(lib.sfm_f_inlier_mask(device.ptr(d_x), n, device.ptr(r.F_best), float(threshold),
                                     device.ptr(r.mask), 0, st), "sfm_f_inlier_mask")
    packed = r.packed.cpu().numpy()            # synchronises the stream
    return dict(F_best=packed[:72].view(np.float64).reshape(3, 3).copy(),
                best_idx=int(packed[72:76].view(np.int32)[0]),
                best_count=int(packed[76:80].view(np.int32)[0]),
                mask=packed[80:].view(np.bool_))        # K4 writes 0 / 1 bytes


def GetInliersRANSAC(points1, points2, index, threshold=0.06, n_max=1000):
    """Drop-in for GetInliersRANSAC.py:5-106 -> (inlier_index, outlier_indices, F_best)."""
    points1 = np.asarray(points1)                      # the reference copies (np.array); nothing here writes to them
    points2 = np.asarray(points2)
    index = np.asarray(index)
    n_points = len(points1)
    if n_points < 8:                                   # GetInliersRANSAC.py:38-40 (no RNG draw)
        return np.array([]), index, None
    shape_ok = points1.ndim == 2 and points1.shape[1] == 2 and points2.shape == points1.shape
    if shape_ok and int(n_max) >= STREAM_MIN_HYPOTHESES:
        res = f_ransac_stream(points1, points2, threshold, int(n_max))
    else:
        samples = sampler.sample_stream(n_points, 8, int(n_max))   # one draw per iteration (:55)
        if n_max <= 0:
            return np.array([]), index, None
        if not shape_ok:
            # EstimateFundamentalMatrix raises inside the try of every iteration (:59,:90-92)
            return np.array([]), index, None
        res = f_ransac_host(points1, points2, samples, threshold)
    if res["best_count"] == 0:                         # :95-96
        return np.array([]), index, None
    mask = res["mask"]
    return np.where(mask)[0], index[~mask], res["F_best"]


def get_inliers_ransac(points1, points2, index, threshold=0.06, n_max=1000):
    """Drop-in for the wrapper alias, GetInliersRANSAC.py:109-121 -> (F_best, index[inliers])."""
    inlier_index, _, F_best = GetInliersRANSAC(points1, points2, index, threshold, n_max)
    if len(inlier_index) > 0:
        inlier_idx = np.array(index)[inlier_index]
    else:
        inlier_idx = np.array([])
    return F_best, inlier_idx


def EstimateFundamentalMatrix(points1, points2):
    """Drop-in for EstimateFundamentalMatrix.py:3-83 (any N >= 8; N == 8 uses K1)."""
    points1 = np.array(points1)
    points2 = np.array(points2)
    if points1.ndim != 2 or points1.shape[1] != 2:     # :80-81
        raise ValueError("Points must be in 2D format (N x 2)")
    n = points1.shape[0]
    if n < 8:
        # the reference still runs its SVD here and returns a rank-deficient guess that depends
        # on LAPACK internals; there is nothing well-defined to reproduce.
        raise ValueError("EstimateFundamentalMatrix needs at least 8 correspondences")
    lib = _lib.load()
    d_F = device.empty((9,), torch.float64)
    if n == 8:
        d_x = device.to_device(np.hstack([points1.astype(np.float64), points2.astype(np.float64)]))
        d_s = device.to_device(np.arange(8, dtype=np.int32).reshape(1, 8), np.int32)
        _lib.check(lib.sfm_f8pt_batch(device.ptr(d_x), 8, device.ptr(d_s), 1, device.ptr(d_F),
                                      device.stream()), "sfm_f8pt_batch")
    else:
        d_x1, d_x2 = device.to_device(points1), device.to_device(points2)
        nbytes = int(lib.sfm_fundamental_npt_workspace_bytes(n))
        ws = device.empty((nbytes,), torch.uint8)
        _lib.check(lib.sfm_fundamental_npt(device.ptr(d_x1), device.ptr(d_x2), n, device.ptr(d_F),
                                           device.ptr(ws), nbytes, device.stream()), "sfm_fundamental_npt")
    return d_F.cpu().numpy().reshape(3, 3)
