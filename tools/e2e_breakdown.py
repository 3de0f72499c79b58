#!/usr/bin/env python
"""Where the end-to-end time of GetInliersRANSAC(…, n_max=100000) goes (diagnostic): host phases timed
with perf_counter, with and without device synchronisation after each."""
import os, random, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from sfm_b200 import ransac, sampler, device, _lib
from oracle import scenes

sc = scenes.two_view()
idx = np.arange(50000)
random.seed(1)
for _ in range(3):
    ransac.GetInliersRANSAC(sc["x1"], sc["x2"], idx, 0.06, 100000)
torch.cuda.synchronize()
ts = []
for _ in range(10):
    t = time.perf_counter()
    ransac.GetInliersRANSAC(sc["x1"], sc["x2"], idx, 0.06, 100000)
    ts.append(time.perf_counter() - t)
print("GetInliersRANSAC e2e ms: min %.2f median %.2f" % (1e3 * min(ts), 1e3 * sorted(ts)[5]))
# pieces
def T(f, n=10):
    torch.cuda.synchronize(); v = []
    for _ in range(n):
        t = time.perf_counter(); f(); torch.cuda.synchronize(); v.append(time.perf_counter() - t)
    return 1e3 * min(v)
p1 = np.ascontiguousarray(sc["x1"]); p2 = np.ascontiguousarray(sc["x2"])
print("H2D of both point arrays + cat: %.3f ms" % T(lambda: torch.cat([device.to_device(p1), device.to_device(p2)], dim=1)))
print("alloc result buffers: %.3f ms" % T(lambda: ransac._alloc_result(50000, 100000, False)))
pinned = ransac._pinned_i32(100000, 8); host = pinned.numpy()
def rep_all():
    r = sampler.StreamReplayer(50000, 8); r.draw(100000, host); r.finish()
print("replay of 100k index sets (one call): %.3f ms" % T(rep_all))
def rep_first():
    r = sampler.StreamReplayer(50000, 8); r.draw(2048, host[:2048]); r.finish()
print("replayer setup + first 2048 + setstate: %.3f ms" % T(rep_first))
d_x = torch.cat([device.to_device(p1), device.to_device(p2)], dim=1)
d_s = device.to_device(host, np.int32)
plan = ransac.FRansacPlan(50000, 100000)
print("device-resident K1..K4: %.3f ms" % T(lambda: plan.run(d_x, d_s, 0.06)))
r = plan.run(d_x, d_s, 0.06)
print("D2H packed + numpy post: %.3f ms" % T(lambda: (lambda pk: (np.where(pk[80:].astype(bool))[0], idx[~pk[80:].astype(bool)]))(r.packed.cpu().numpy())))
for first in (512, 1024, 2048, 4096):
    for chunks in (3, 4, 6):
        ransac.STREAM_FIRST_CHUNK = first
        v = []
        for _ in range(6):
            torch.cuda.synchronize(); t = time.perf_counter()
            ransac.f_ransac_stream(sc["x1"], sc["x2"], 0.06, 100000, chunks=chunks)
            v.append(time.perf_counter() - t)
        print("first", first, "chunks", chunks, "ms %.2f" % (1e3 * min(v)))
