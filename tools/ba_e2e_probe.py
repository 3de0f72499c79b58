"""Where a host-array BA call (bundle.ba_solve_sparse, C3) spends its time: cProfile of 5 calls after a warm one,
alone and after the device-resident solves bench_ba runs first."""
import cProfile
import pstats
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
import bench_ba  # noqa: E402
import bench_config as cfg  # noqa: E402
from sfm_b200 import bundle  # noqa: E402

sc, cams, pts = bench_ba._scene(cfg.C3)
args = (cams, pts, sc["cam_idx"], sc["pt_idx"], sc["obs"], sc["K"])


def run(tag):
    bundle.ba_solve_sparse(*args)
    torch.cuda.synchronize()
    ts = []
    for _ in range(5):
        t = time.perf_counter(); bundle.ba_solve_sparse(*args); ts.append(time.perf_counter() - t)
    print(tag, "ms per call:", [round(1e3 * x, 3) for x in ts])


run("alone")
prob = bundle.BAProblem(cfg.C3["n_cams"], cfg.C3["n_pts"], sc["cam_idx"], sc["pt_idx"], sc["obs"], sc["K"])
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
from sfm_b200 import device
for _ in range(4):
    d_c, d_p = device.to_device(cams), device.to_device(pts)
    flush.zero_()
    prob.solve_device(d_c, d_p)
torch.cuda.synchronize()
run("after device solves with another BAProblem")
pr = cProfile.Profile()
pr.enable()
for _ in range(5):
    bundle.ba_solve_sparse(*args)
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(18)
