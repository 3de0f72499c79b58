"""Where pnp_ransac.pnp_ransac_batch (C4: 500 registrations x 2,000 correspondences x 1,000 hypotheses) spends its time."""
import cProfile
import pstats
import random
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from oracle import scenes  # noqa: E402
from sfm_b200 import pnp_ransac  # noqa: E402

regs = [scenes.pnp_scene(n=2000, seed=5000 + b) for b in range(500)]
Xs, xs = [r["X"] for r in regs], [r["x"] for r in regs]
random.seed(777)
pnp_ransac.pnp_ransac_batch(Xs[:8], xs[:8], scenes.K_REF)
torch.cuda.synchronize()
for _ in range(2):
    random.seed(777)
    t = time.perf_counter()
    C, R, counts = pnp_ransac.pnp_ransac_batch(Xs, xs, scenes.K_REF)
    torch.cuda.synchronize()
    print("batch of 500: %.2f ms -> %.0f registrations/s, mean inliers %.1f" % (1e3 * (time.perf_counter() - t), 500 / (time.perf_counter() - t), counts.mean()))
random.seed(777)
pr = cProfile.Profile()
pr.enable()
pnp_ransac.pnp_ransac_batch(Xs, xs, scenes.K_REF)
torch.cuda.synchronize()
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(14)
