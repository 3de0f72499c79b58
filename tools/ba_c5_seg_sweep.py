"""C5 (or C3: first argument) BA on one GPU for several slice lengths of the K10g pair list (SFM_SCHUR_SEG): phase times per solve."""
import os
import subprocess
import sys

CHILD = r'''
import sys, torch
sys.path.insert(0, ".")
import bench_ba, bench_config as cfg
from sfm_b200 import bundle, device
import os
C = getattr(cfg, os.environ.get("SFM_SWEEP_CFG", "C5"))
sc, cams, pts = bench_ba._scene(C)
prob = bundle.BAProblem(C["n_cams"], C["n_pts"], sc["cam_idx"], sc["pt_idx"], sc["obs"], sc["K"])
for _ in range(2):
    prob.solve_device(device.to_device(cams), device.to_device(pts))
ms = []
for _ in range(8):
    d_c, d_p = device.to_device(cams), device.to_device(pts)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); info = prob.solve_device(d_c, d_p); b.record(); torch.cuda.synchronize()
    ms.append(a.elapsed_time(b))
prob.solve_device(device.to_device(cams), device.to_device(pts), profile=True)
p = bundle.last_profile()
print("solve ms min", round(min(ms), 4), "schur_ms", round(p["schur_ms"], 3), "k11", round(p["camera_solve_k11_ms"], 3), "cost", info["cost"])
'''
args = sys.argv[1:]
if args and args[0] in ("C3", "C5"):
    os.environ["SFM_SWEEP_CFG"] = args.pop(0)
for seg in args or ["48", "32", "24", "16", "12"]:
    e = dict(os.environ); e["SFM_SCHUR_SEG"] = seg
    r = subprocess.run([sys.executable, "-c", CHILD], env=e, capture_output=True, text=True)
    print("seg", seg, r.stdout.strip().splitlines()[-1] if r.stdout.strip() else r.stderr[-300:])
