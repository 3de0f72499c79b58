"""Host-array BA call at C5 (bundle.ba_solve_sparse): per-call time and a cProfile of where the host side goes."""
import cProfile
import pstats
import sys
import time

import torch

sys.path.insert(0, ".")
import bench_ba  # noqa: E402
import bench_config as cfg  # noqa: E402
from sfm_b200 import bundle  # noqa: E402

sc, cams, pts = bench_ba._scene(cfg.C5)
args = (cams, pts, sc["cam_idx"], sc["pt_idx"], sc["obs"], sc["K"])
for _ in range(2):
    bundle.ba_solve_sparse(*args)
torch.cuda.synchronize()
ts = []
for _ in range(4):
    t = time.perf_counter(); c, p, info = bundle.ba_solve_sparse(*args); ts.append(time.perf_counter() - t)
print("ms per call:", [round(1e3 * x, 2) for x in ts], "cost", info["cost"])
pr = cProfile.Profile()
pr.enable()
for _ in range(3):
    bundle.ba_solve_sparse(*args)
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(10)
