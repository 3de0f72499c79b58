"""K2 (f_score_packed_kernel) with and without the early reject across thresholds, at C2 sizes: which launch rule pays.
Each (mode) runs in its own process because the library reads SFM_F_SCORE_EARLY / SFM_F_SCORE_NO_EARLY once."""
import os
import subprocess
import sys

CHILD = r'''
import sys, random, numpy as np, torch
sys.path.insert(0, ".")
import bench_config as cfg
from sfm_b200 import ransac, device, sampler, _lib
sc = cfg.two_view_scene()
N, H = len(sc["x1"]), 100000
random.seed(1)
s = sampler.sample_stream(N, 8, H)
d_x = torch.cat([device.to_device(sc["x1"]), device.to_device(sc["x2"])], dim=1)
d_s = device.to_device(s, np.int32)
plan = ransac.FRansacPlan(N, H)
lib = _lib.load()
out = []
for thr in (0.06, 0.15, 0.3, 0.6, 1.0, 3.0):
    plan.run(d_x, d_s, thr)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(5):
        lib.sfm_f_score_count(device.ptr(d_x), N, device.ptr(plan.r.F), H, thr, device.ptr(plan.r.counts), device.stream())
    b.record(); torch.cuda.synchronize()
    out.append((thr, a.elapsed_time(b) / 5, int(plan.r.counts.max().item())))
print(" ".join(f"thr={t}: {ms:.3f} ms (best {c})" for t, ms, c in out))
'''
for mode, env in (("early", {"SFM_F_SCORE_EARLY": "1"}), ("no early", {"SFM_F_SCORE_NO_EARLY": "1"}), ("rule", {})):
    e = dict(os.environ); e.update(env)
    r = subprocess.run([sys.executable, "-c", CHILD], env=e, capture_output=True, text=True)
    print(f"{mode:9s}", r.stdout.strip().splitlines()[-1] if r.stdout.strip() else r.stderr[-400:])
