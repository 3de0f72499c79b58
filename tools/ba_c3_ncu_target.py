"""One device-resident BA solve at C3 (50 cameras / 20k points / 200k observations) for ncu captures."""
import sys
import torch
sys.path.insert(0, ".")
import bench_ba
import bench_config as cfg
from sfm_b200 import bundle, device
sc, cams, pts = bench_ba._scene(cfg.C3)
prob = bundle.BAProblem(cfg.C3["n_cams"], cfg.C3["n_pts"], sc["cam_idx"], sc["pt_idx"], sc["obs"], sc["K"])
for _ in range(2):
    info = prob.solve_device(device.to_device(cams), device.to_device(pts))
torch.cuda.synchronize()
print("ok", info["cost"], info["n_linearisations"])
