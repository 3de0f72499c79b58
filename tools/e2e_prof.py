"""Host-side cost of the chunk-pipelined GetInliersRANSAC path (diagnostic)."""
import sys, time, random, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from sfm_b200 import ransac
from oracle import scenes
sc = scenes.two_view()
idx = np.arange(50000)
random.seed(1)
for _ in range(2): ransac.GetInliersRANSAC(sc["x1"], sc["x2"], idx, 0.06, 100000)
for first in (1024, 4096, 8192):
    for chunks in (2, 3, 4, 6, 8):
        ransac.STREAM_FIRST_CHUNK = first
        ts = []
        for _ in range(5):
            torch.cuda.synchronize(); t = time.perf_counter()
            ransac.f_ransac_stream(sc["x1"], sc["x2"], 0.06, 100000, chunks=chunks)
            ts.append(time.perf_counter() - t)
        print("first", first, "chunks", chunks, "ms %.2f" % (1e3 * min(ts)))
