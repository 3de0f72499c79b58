"""Host timeline of one ransac.f_ransac_stream call (an instrumented copy of its body): where the milliseconds
beyond the kernels go."""
import random
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
import bench_config as cfg  # noqa: E402
from sfm_b200 import ransac, sampler, device, _lib  # noqa: E402

sc = cfg.two_view_scene()
N, H, T = len(sc["x1"]), 100000, 0.06
random.seed(1)
for _ in range(3):
    ransac.f_ransac_stream(sc["x1"], sc["x2"], T, H)
lib = _lib.load()


def traced():
    marks = []
    def mark(name):
        marks.append((name, time.perf_counter()))
    mark("start")
    p1 = np.asarray(sc["x1"]); p2 = np.asarray(sc["x2"])
    n, h = N, H
    pin_x = ransac._pinned_f64(2 * n, 2); hx = pin_x.numpy(); np.copyto(hx[:n], p1); np.copyto(hx[n:], p2)
    mark("points staged")
    st = device.stream()
    r, key, d_s, d_pp = ransac._stream_buffers(n, h)
    d_pp.copy_(pin_x, non_blocking=True)
    d_x = torch.cat([d_pp[:n], d_pp[n:]], dim=1)
    pinned = ransac._pinned_i32(h, 8); host = pinned.numpy()
    mark("allocated")
    big = max(1, -(-h // 2)); sizes, c = [], min(big, ransac.STREAM_FIRST_CHUNK)
    while sum(sizes) < h:
        sizes.append(min(c, h - sum(sizes))); c = min(big, int(c * ransac.STREAM_GROWTH))
    rep = sampler.StreamReplayer(n, 8)
    main = torch.cuda.current_stream(); side = ransac._side_stream(); side.wait_stream(main)
    mark("replayer ready")
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(len(sizes) + 3)]
    ev[0].record(main)
    off = 0
    for i, c in enumerate(sizes):
        rep.draw(c, host[off:off + c])
        mark(f"draw {c}")
        with torch.cuda.stream(side):
            d_s[off:off + c].copy_(pinned[off:off + c], non_blocking=True)
            lib.sfm_f8pt_batch(device.ptr(d_x), n, device.ptr(d_s) + 32 * off, c, device.ptr(r.F) + 72 * off, side.cuda_stream)
        main.wait_stream(side)
        lib.sfm_f_score_count(device.ptr(d_x), n, device.ptr(r.F) + 72 * off, c, float(T), device.ptr(r.counts) + 4 * off, st)
        ev[i + 1].record(main)
        mark(f"launched {c}")
        off += c
    main.wait_stream(side)
    rep.finish()
    mark("rng written back")
    lib.sfm_argmax_first(device.ptr(r.counts), h, 0, device.ptr(r.best), device.ptr(key), st)
    r.F_best.copy_(r.F.index_select(0, r.best[:1].long()).view(9))
    lib.sfm_f_inlier_mask(device.ptr(d_x), n, device.ptr(r.F_best), float(T), device.ptr(r.mask), 0, st)
    ev[len(sizes) + 1].record(main)
    mark("tail launched")
    packed = r.packed.cpu().numpy()
    mark("result on host")
    t0 = marks[0][1]
    for name, t in marks:
        print(f"{1e3 * (t - t0):8.3f} ms  {name}")
    for i in range(len(sizes) + 1):
        print(f"gpu segment {i}: {ev[i].elapsed_time(ev[i + 1]):.3f} ms")


traced()
for first, growth in [(4096, 1.4), (2048, 1.4), (3072, 1.5), (4096, 1.3), (6144, 1.3)]:
    ransac.STREAM_FIRST_CHUNK, ransac.STREAM_GROWTH = first, growth
    ts = []
    for _ in range(8):
        t = time.perf_counter(); ransac.f_ransac_stream(sc["x1"], sc["x2"], T, H); ts.append(time.perf_counter() - t)
    print("first", first, "growth", growth, "ms min", round(1e3 * min(ts), 3), "median", round(1e3 * sorted(ts)[4], 3))
ransac.STREAM_FIRST_CHUNK, ransac.STREAM_GROWTH = 4096, 1.4
print("----")
traced()
