import os, sys
sys.path.insert(0, "/root/repo")
import torch
import bench_ba
from sfm_b200 import bundle
sc5, cams5, pts5 = bench_ba._scene(bench_ba.C5)
prob5 = bundle.BAProblem(1000, 500000, sc5["cam_idx"], sc5["pt_idx"], sc5["obs"], sc5["K"])
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
print(bench_ba.time_pass(prob5, cams5, pts5, 1, 3, flush))
