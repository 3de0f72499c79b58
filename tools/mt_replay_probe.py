import sys, time, random, numpy as np
sys.path.insert(0, ".")
from sfm_b200 import sampler
random.seed(1)
out = np.empty((100000, 8), np.int32)
rep = sampler.StreamReplayer(50000, 8)
ts = []
for _ in range(12):
    t = time.perf_counter(); rep.draw(100000, out); ts.append(time.perf_counter() - t)
rep.finish()
print("draw 100k x 8: min ms", round(min(ts) * 1e3, 3), "median", round(sorted(ts)[6] * 1e3, 3))
