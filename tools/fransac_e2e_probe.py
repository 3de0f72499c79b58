"""Where the end-to-end F-RANSAC call (GetInliersRANSAC on host arrays, C2) spends its time beyond the kernels:
host-side timestamps of every stage of ransac.f_ransac_stream, plus the whole call and the device-only step."""
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
idx = np.arange(N)
random.seed(1)
for _ in range(3):
    ransac.GetInliersRANSAC(sc["x1"], sc["x2"], idx, T, H)
torch.cuda.synchronize()
ts = []
for _ in range(5):
    t = time.perf_counter()
    ransac.GetInliersRANSAC(sc["x1"], sc["x2"], idx, T, H)
    ts.append(time.perf_counter() - t)
print("whole call ms:", [round(1e3 * x, 3) for x in ts])
# the replay alone
rep_t = []
for _ in range(3):
    out = np.empty((H, 8), np.int32)
    t = time.perf_counter()
    rep = sampler.StreamReplayer(N, 8)
    rep.draw(H, out)
    rep.finish()
    rep_t.append(time.perf_counter() - t)
print("MT replay of 100k x 8 ms:", [round(1e3 * x, 3) for x in rep_t])
# device-only: samples resident
d_x = torch.cat([device.to_device(sc["x1"]), device.to_device(sc["x2"])], dim=1)
d_s = device.to_device(out, np.int32)
ctx = ransac.FRansacPlan(N, H)
for _ in range(3):
    ctx.run(d_x, d_s, T)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    ctx.run(d_x, d_s, T)
e1.record()
torch.cuda.synchronize()
print("device-only step ms:", e0.elapsed_time(e1) / 5)
# pieces of the host path
t = time.perf_counter(); p1 = np.ascontiguousarray(sc["x1"], np.float64); p2 = np.ascontiguousarray(sc["x2"], np.float64)
a = device.to_device(p1); b = device.to_device(p2); d = torch.cat([a, b], dim=1); torch.cuda.synchronize()
print("points H2D + cat ms:", 1e3 * (time.perf_counter() - t))
t = time.perf_counter(); r = ransac._alloc_result(N, H, False); k = device.empty((1,), torch.int64); s = device.empty((H, 8), torch.int32); torch.cuda.synchronize()
print("allocations ms:", 1e3 * (time.perf_counter() - t))
res = ransac.f_ransac_stream(sc["x1"], sc["x2"], T, H)
t = time.perf_counter(); m = res["mask"]; i1 = np.where(m)[0]; i2 = idx[~m]
print("host post ms:", 1e3 * (time.perf_counter() - t))
for first, growth, chunks in [(3072, 2.5, 2), (2048, 1.5, 2), (2048, 1.4, 2), (2048, 1.3, 2), (4096, 1.4, 2), (1024, 1.5, 2), (2048, 1.4, 4),
                              (2048, 1.25, 2), (3072, 1.2, 3)]:
    ransac.STREAM_FIRST_CHUNK, ransac.STREAM_GROWTH = first, growth
    ts = []
    for _ in range(6):
        t = time.perf_counter(); ransac.f_ransac_stream(sc["x1"], sc["x2"], T, H, chunks=chunks); ts.append(time.perf_counter() - t)
    print("first", first, "growth", growth, "chunks", chunks, "ms min", round(1e3 * min(ts), 3), "median", round(1e3 * sorted(ts)[3], 3))
