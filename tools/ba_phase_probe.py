"""Phase times of device-resident BA solves at C3 and C5 (bundle.last_profile) and the solve time itself."""
import sys
import torch
sys.path.insert(0, ".")
import bench_ba
import bench_config as cfg
from sfm_b200 import bundle, device
for name, C in (("C3", cfg.C3), ("C5", cfg.C5)):
    sc, cams, pts = bench_ba._scene(C)
    prob = bundle.BAProblem(C["n_cams"], C["n_pts"], sc["cam_idx"], sc["pt_idx"], sc["obs"], sc["K"])
    for _ in range(3):
        prob.solve_device(device.to_device(cams), device.to_device(pts))
    ms = []
    for _ in range(5):
        d_c, d_p = device.to_device(cams), device.to_device(pts)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); info = prob.solve_device(d_c, d_p); b.record(); torch.cuda.synchronize()
        ms.append(a.elapsed_time(b))
    prob.solve_device(device.to_device(cams), device.to_device(pts), profile=True)
    p = bundle.last_profile()
    print(name, "solve ms", round(min(ms), 4), "LM it/s", round(info["n_linearisations"] / min(ms) * 1e3, 1), "cost", repr(info["cost"]),
          {k: round(float(v), 4) for k, v in p.items() if k.endswith("_ms")})
    prob.close()
