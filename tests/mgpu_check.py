#!/usr/bin/env python
"""Multi-GPU parity check, launched by torchrun (one process per GPU): point-sharded BA and
hypothesis-sharded F-RANSAC must reproduce the single-GPU results. Run on the GPU box:
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29511 tests/mgpu_check.py
Exit code 0 on success (every rank asserts)."""
import os
import random
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    from scipy.spatial.transform import Rotation
    from sfm_b200 import bundle, bundle_mgpu, device, ransac, sampler
    from oracle import scenes
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    comm = bundle_mgpu.NcclComm(rank, world)

    # ---- BA: sharded vs single GPU on the same scene ----
    for cfg in (dict(n_cams=12, n_pts=300, obs_per_pt=6, seed=9), dict(n_cams=50, n_pts=20000, obs_per_pt=10, seed=20260103)):
        sc = scenes.ba_scene(**cfg)
        cams0 = np.hstack([Rotation.from_matrix(sc["R0"]).as_rotvec(), -np.einsum("nij,nj->ni", sc["R0"], sc["C0"])])
        cams_s, pts_s, info_s = bundle.ba_solve_sparse(cams0, sc["pts0"], sc["cam_idx"], sc["pt_idx"], sc["obs"], sc["K"])
        cams_m, pts_m, (lo, hi), info_m = bundle_mgpu.ba_solve_sharded(cams0, sc["pts0"], sc["cam_idx"], sc["pt_idx"],
                                                                      sc["obs"], sc["K"], comm)
        rel = abs(info_m["cost"] - info_s["cost"]) / info_s["cost"]
        assert rel < 1e-9, (rank, info_m, info_s)
        assert info_m["n_linearisations"] == info_s["n_linearisations"], (info_m, info_s)
        assert np.max(np.abs(cams_m - cams_s)) < 1e-7 and np.max(np.abs(pts_m - pts_s[lo:hi])) < 1e-6
        # every rank ends with bit-identical cameras
        t = torch.from_numpy(cams_m).cuda()
        ref = t.clone()
        dist.broadcast(ref, src=0)
        assert torch.equal(t, ref), "cameras differ between ranks"
        if rank == 0:
            print(f"BA {cfg['n_cams']}/{cfg['n_pts']}: sharded cost {info_m['cost']:.9f} vs single {info_s['cost']:.9f} "
                  f"(rel {rel:.1e}), {info_m['n_linearisations']} LM iterations on {world} GPUs", flush=True)

    # ---- F-RANSAC: hypothesis-sharded winner == single-GPU winner over the whole stream ----
    n, h = 5000, 4096
    sc = scenes.two_view(n=n, seed=3)
    random.seed(11)
    samples = sampler.sample_stream(n, 8, h * world)
    x1x2 = np.ascontiguousarray(np.hstack([sc["x1"], sc["x2"]]))
    d_x, d_s_all = device.to_device(x1x2), device.to_device(samples, np.int32)
    single = ransac.f_ransac_host(sc["x1"], sc["x2"], samples, 0.5)
    plan = ransac.FRansacPlan(n, h)
    for reducer in (None, comm.allreduce_max_u64):          # torch.distributed and the library's communicator
        r = plan.run_sharded(d_x, d_s_all, 0.5, rank, world, all_reduce_max=reducer)
        best = r.best.cpu().numpy()
        assert int(best[0]) == single["best_idx"] and int(best[1]) == single["best_count"], (best, single["best_idx"])
        assert np.array_equal(r.mask.cpu().numpy().astype(bool), single["mask"])
        assert np.array_equal(r.F_best.cpu().numpy().reshape(3, 3), single["F_best"])     # the same bits as one GPU
    if rank == 0:
        print(f"F-RANSAC sharded winner {single['best_idx']} ({single['best_count']} inliers) on {world} GPUs: ok", flush=True)
    # ---- independent problems (second row): chunks per rank + all-gather == the single-GPU call ----
    from sfm_b200 import refine, shard
    tv = scenes.two_view(n=2001, outlier_frac=0.0, seed=21)
    args = (tv["K"], tv["C1"], tv["R1"], tv["C2"], tv["R2"], tv["x1"], tv["x2"])
    X0 = shard.linear_triangulation_sharded(*args)
    assert np.array_equal(X0, refine.LinearTriangulation(*args))
    X = shard.nonlinear_triangulation_sharded(*args, X0)
    assert np.array_equal(X, refine.NonLinearTriangulation(*args, X0))
    regs = [scenes.pnp_scene(n=200 + 10 * b, outlier_frac=0.0, seed=300 + b) for b in range(7)]
    Xs, xs = [r["X"] for r in regs], [r["x"] for r in regs]
    C0 = [r["C"] + 0.05 for r in regs]
    R0 = [r["R"] for r in regs]
    Cs, Rs = shard.nonlinear_pnp_sharded(Xs, xs, scenes.K_REF, C0, R0)
    C1, R1, _ = refine.nonlinear_pnp_batch(Xs, xs, scenes.K_REF, C0, R0)
    assert np.array_equal(Cs, np.array(C1)) and np.array_equal(Rs, np.array(R1))
    if rank == 0:
        print(f"sharded triangulation (2001 points) and NonLinearPnP (7 registrations) on {world} GPUs: bit-equal to one GPU", flush=True)
    comm.destroy()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
