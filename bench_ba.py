"""bench_ba.py -- the bundle-adjustment half of bench.py (BASELINE.json metric "BA LM iters/sec").

One step = one complete sfm_ba_solve (K8-K12 Levenberg-Marquardt to convergence, scipy's default
tolerances) of the synthetic scene from the same perturbed start; an "LM iteration" is one
linearisation (Jacobian / normal-equation build + Schur solve, with all of its trial steps).
  N = 1 : BASELINE configs[2]  (50 cameras, 20k points, 200k observations)  -> headline BA number
          plus BASELINE configs[4] (1,000 cameras, 500k points, 4M observations) on the one GPU for
          the HBM-roofline evidence of the residual/Jacobian pass K9p (C3 is launch-latency sized).
  N > 1 : configs[4] sharded by point across the ranks (see sfm_b200.bundle_mgpu), when available.
"""
from __future__ import annotations

import os
import time

import numpy as np

from bench_config import C3, C5  # noqa: E402


def _scene(cfg):
    from scipy.spatial.transform import Rotation
    from oracle import scenes
    sc = scenes.ba_scene(**cfg)
    cams0 = np.hstack([Rotation.from_matrix(sc["R0"]).as_rotvec(),
                       -np.einsum("nij,nj->ni", sc["R0"], sc["C0"])])
    return sc, np.ascontiguousarray(cams0), np.ascontiguousarray(sc["pts0"])


def _name(cfg):
    return f"{cfg['n_cams']} cameras, {cfg['n_pts']} points, {cfg['n_pts'] * cfg['obs_per_pt']} observations"


def time_solves(prob, cams0, pts0, steps, warmup, flush):
    """CUDA-event time of `steps` complete solves (inputs resident in HBM). Returns (ms, infos)."""
    import torch
    from sfm_b200 import device
    d_cam0, d_pts0 = device.to_device(cams0), device.to_device(pts0)
    d_cam, d_pts = torch.empty_like(d_cam0), torch.empty_like(d_pts0)
    infos = []
    for _ in range(warmup):
        d_cam.copy_(d_cam0); d_pts.copy_(d_pts0)
        prob.solve_device(d_cam, d_pts)
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for a, b in ev:
        flush.zero_()
        d_cam.copy_(d_cam0); d_pts.copy_(d_pts0)
        a.record()
        infos.append(prob.solve_device(d_cam, d_pts))
        b.record()
    torch.cuda.synchronize()
    return sum(a.elapsed_time(b) for a, b in ev), infos


def time_pass(prob, cams0, pts0, which, reps, flush):
    """Average CUDA-event duration (ms) of one isolated linearisation pass."""
    import torch
    from sfm_b200 import device
    d_cam, d_pts = device.to_device(cams0), device.to_device(pts0)
    prob.linearise_device(d_cam, d_pts, which)             # builds index + camera table
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for a, b in ev:
        flush.zero_()
        a.record()
        prob.linearise_device(d_cam, d_pts, which, reuse_table=True)
        b.record()
    torch.cuda.synchronize()
    return sum(a.elapsed_time(b) for a, b in ev) / reps


def device_copy(a):
    from sfm_b200 import device
    return device.to_device(np.ascontiguousarray(a))


def k9p_bytes(n_obs, n_pts):
    """Algorithmic bytes of the point pass K9p: per observation obs 16 + cam_idx 4 +
    pt_idx 4 in, W 144 out; per point X 24 in, V 48 + g_p 24 out (+ the memset of V, g_p: 72)."""
    return 168.0 * n_obs + 168.0 * n_pts


def k9c_bytes(n_obs):
    """Camera pass K9c: per observation obs 16 + point index 4 + gathered point 24."""
    return 44.0 * n_obs


def k8_bytes(n_obs):
    """Cost pass K8: obs 16 + cam_idx 4 + pt_idx 4 + gathered point 24 (camera table in shared memory)."""
    return 48.0 * n_obs


def cpu_baseline(seconds_budget=20.0):
    """The reference's BA algorithm (scipy least_squares 'lm', dense forward-difference Jacobian,
    BundleAdjustment.py:205-212) through the oracle's vectorised residual, at a size the host can
    finish; the reference cannot run configs[2] at all."""
    from scipy.spatial.transform import Rotation
    from oracle import ba_oracle as bo
    from oracle import scenes
    nc, npts, k = 12, 300, 6
    sc = scenes.ba_scene(n_cams=nc, n_pts=npts, obs_per_pt=k, seed=9)
    cams0 = np.hstack([Rotation.from_matrix(sc["R0"]).as_rotvec(), -np.einsum("nij,nj->ni", sc["R0"], sc["C0"])])
    x0 = np.concatenate([cams0.reshape(-1), sc["pts0"].reshape(-1)])
    t = time.perf_counter()
    res = bo.ba_solve(x0, nc, npts, sc["cam_idx"], sc["pt_idx"], sc["obs"], sc["K"], method="lm")
    dt = time.perf_counter() - t
    iters = int(res.njev) if getattr(res, "njev", None) else int(res.nfev)
    nobs, nvar = len(sc["obs"]), 6 * nc + 3 * npts
    per_iter = dt / max(iters, 1)
    # one LM iteration of the reference = one dense FD Jacobian = nvar residual calls of cost ~ nobs
    scale = (200_000 * 60_300) / (nobs * nvar)
    return {"value": iters / dt, "unit": "LM iterations/s", "cores": 1, "kind": "port",
            "sample": f"scipy least_squares('lm', dense 2-point Jacobian) on {nc} cameras / {npts} points / {nobs} "
                      f"observations ({nvar} parameters): {iters} Jacobian evaluations, {res.nfev} trial steps in {dt:.1f} s, "
                      f"final cost {res.cost:.6f}",
            "extrapolated_c3": {"value": 1.0 / (per_iter * scale), "unit": "LM iterations/s",
                                "how": "time per iteration scaled by (n_obs x n_params) to 200k x 60,300; the "
                                       "reference cannot run this size (193 GB dense Jacobian): an extrapolation, "
                                       "not a measurement"}}


def run_single_gpu(args, peaks, peak_src, flush):
    """N = 1: C3 headline + C5 roofline pass. Returns the `ba` object of the bench line."""
    import torch
    from sfm_b200 import _lib, bundle
    out = {"metric": "BA LM iterations/sec", "unit": "LM iterations/s", "dtype": "f64"}
    steps = max(3, min(args.steps, 10))
    # ---- C3 ----
    sc, cams0, pts0 = _scene(C3)
    prob = bundle.BAProblem(C3["n_cams"], C3["n_pts"], sc["cam_idx"], sc["pt_idx"], sc["obs"], sc["K"])
    l0 = _lib.launch_count()
    ms, infos = time_solves(prob, cams0, pts0, steps, 3, flush)
    launches = _lib.launch_count() - l0
    its = sum(i["n_linearisations"] for i in infos)
    nobs = len(sc["obs"])
    out.update(value=its / (ms * 1e-3), ms_per_solve=ms / steps, solves=steps,
               config={"workload": "C3 synthetic BA: " + _name(C3) + ", LM to convergence (ftol=xtol=gtol=1e-8)",
                       "l2": "256 MiB flush write between timed solves"},
               lm_iterations_per_solve=its / steps, trial_steps_per_solve=sum(i["n_trials"] for i in infos) / steps,
               status=infos[-1]["status"], cost0=infos[-1]["cost0"], cost=infos[-1]["cost"],
               rmse_px=float(np.sqrt(2 * infos[-1]["cost"] / nobs)), gpu_launches=int(launches * steps / (steps + 3)))
    # end to end through the host-array API (H2D of the problem, solve, D2H of the parameters)
    e2e_steps = 10
    for _ in range(3):      # warm: the first call builds the problem, the second allocates its pinned staging buffer
        bundle.ba_solve_sparse(cams0, pts0, sc["cam_idx"], sc["pt_idx"], sc["obs"], sc["K"])
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    e_its = 0
    for _ in range(e2e_steps):
        c, p, info = bundle.ba_solve_sparse(cams0, pts0, sc["cam_idx"], sc["pt_idx"], sc["obs"], sc["K"])
        e_its += info["n_linearisations"]
    dt = time.perf_counter() - t0
    out["e2e"] = {"value": e_its / dt, "unit": "LM iterations/s",
                  "h2d_bytes_per_step": int(cams0.nbytes + pts0.nbytes + sc["obs"].nbytes),   # the index arrays stay on the device (cache)
                  "d2h_bytes_per_step": int(cams0.nbytes + pts0.nbytes + 64),
                  "api": "sfm_b200.bundle.ba_solve_sparse(cams, pts, cam_idx, pt_idx, obs, K): host numpy in/out, "
                         "H2D of observations and parameters, solve, D2H every step; the per-problem preparation "
                         "(workspace, index, elimination plan) is cached by index structure after the first call",
                  "steps": e2e_steps}
    d_c, d_p = device_copy(cams0), device_copy(pts0)
    prob.solve_device(d_c, d_p, profile=True)
    out["phase_ms_per_solve"] = bundle.last_profile()
    del prob
    # ---- C5 on one GPU: roofline evidence for the streaming passes ----
    sc5, cams5, pts5 = _scene(C5)
    prob5 = bundle.BAProblem(C5["n_cams"], C5["n_pts"], sc5["cam_idx"], sc5["pt_idx"], sc5["obs"], sc5["K"])
    n5, p5 = len(sc5["obs"]), C5["n_pts"]
    ms5, infos5 = time_solves(prob5, cams5, pts5, 3, 2, flush)
    its5 = sum(i["n_linearisations"] for i in infos5)
    passes = {}
    for name, which, nbytes in (("K9p ba_point_pass_kernel", 1, k9p_bytes(n5, p5)),
                                ("K9c ba_camera_pass_kernel", 2, k9c_bytes(n5)),
                                ("K8 ba_cost_kernel", 4, k8_bytes(n5))):
        t_ms = time_pass(prob5, cams5, pts5, which, 10, flush)
        gbs = nbytes / (t_ms * 1e-3) / 1e9
        passes[name] = {"ms": t_ms, "algorithmic_bytes": nbytes, "achieved_gbs": gbs, "frac_of_measured_hbm": gbs / peaks["hbm_gbs"]}
    d_c, d_p = device_copy(cams5), device_copy(pts5)
    prob5.solve_device(d_c, d_p, profile=True)
    c5_phases = bundle.last_profile()
    # C5 end to end through the host-array API (the problem cache holds its own prepared problem after the first call)
    prob5.close()
    for _ in range(2):
        bundle.ba_solve_sparse(cams5, pts5, sc5["cam_idx"], sc5["pt_idx"], sc5["obs"], sc5["K"])
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    e5_its = 0
    for _ in range(3):
        _, _, i5 = bundle.ba_solve_sparse(cams5, pts5, sc5["cam_idx"], sc5["pt_idx"], sc5["obs"], sc5["K"])
        e5_its += i5["n_linearisations"]
    dt5 = time.perf_counter() - t0
    c5_e2e = {"value": e5_its / dt5, "unit": "LM iterations/s", "ms_per_call": dt5 / 3 * 1e3, "steps": 3,
              "h2d_bytes_per_step": int(cams5.nbytes + pts5.nbytes + sc5["obs"].nbytes),
              "d2h_bytes_per_step": int(cams5.nbytes + pts5.nbytes + 64),
              "api": "sfm_b200.bundle.ba_solve_sparse, host numpy in/out"}
    bundle.clear_problem_cache()
    k9 = passes["K9p ba_point_pass_kernel"]
    out["c5_single_gpu"] = {"e2e": c5_e2e, "workload": "C5 synthetic BAL-scale BA on ONE GPU: " + _name(C5),
                            "value": its5 / (ms5 * 1e-3), "unit": "LM iterations/s", "ms_per_solve": ms5 / 3,
                            "lm_iterations_per_solve": its5 / 3, "cost": infos5[-1]["cost"],
                            "rmse_px": float(np.sqrt(2 * infos5[-1]["cost"] / n5)), "passes": passes,
                            "phase_ms_per_solve": c5_phases}
    out["roofline"] = {"bound": "hbm", "kernel": "ba_point_pass_kernel (K9p) on C5", "achieved": k9["achieved_gbs"],
                       "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": k9["frac_of_measured_hbm"],
                       # dram__bytes_read.sum + dram__bytes_write.sum of this kernel on this workload, one
                       # ncu --set full capture (profiles/r01_k9p_c5_v3_ncu_details.txt): 110.3 + 552.4 MB
                       "traffic": 662.7e6,
                       "peak_source": peak_src, "kernel_ms": k9["ms"], "algorithmic_bytes_per_launch": k9["algorithmic_bytes"],
                       "note": "168 B per observation + 168 B per point (DESIGN.md); timed alone with CUDA events, "
                               "includes the 36 MB memset of V/g_p and the per-camera linearisation reuse"}
    return out


def run_sharded(args, rank, world, flush):
    """N > 1: BASELINE configs[4] (C5) sharded by point across the ranks; strong scaling (the problem
    is fixed, each rank holds 1/N of the points). Time = max over ranks of the CUDA-event time. Rank 0 first
    solves the WHOLE problem alone (outside the timed region): the one-GPU figure of the same run on the same
    box, and the parity reference for the sharded result."""
    import torch
    import torch.distributed as dist
    from sfm_b200 import _lib, bundle, bundle_mgpu, device
    comm = bundle_mgpu.NcclComm(rank, world)
    sc, cams0, pts0 = _scene(C5)
    steps = max(3, min(args.steps, 5))
    n1 = None
    if rank == 0:
        whole = bundle.BAProblem(C5["n_cams"], C5["n_pts"], sc["cam_idx"], sc["pt_idx"], sc["obs"], sc["K"])
        ms1, infos1 = time_solves(whole, cams0, pts0, steps, 2, flush)
        d_c1, d_p1 = device_copy(cams0), device_copy(pts0)
        whole.solve_device(d_c1, d_p1)
        n1 = dict(value=sum(i["n_linearisations"] for i in infos1) / (ms1 * 1e-3), ms_per_solve=ms1 / steps,
                  cost=infos1[-1]["cost"], cams=d_c1.cpu().numpy(), pts=d_p1.cpu().numpy())
        del whole, d_c1, d_p1
        bundle.release_cache()
        torch.cuda.empty_cache()
    dist.barrier()
    prob = bundle_mgpu.BAShardedProblem(C5["n_cams"], C5["n_pts"], sc["cam_idx"], sc["pt_idx"], sc["obs"], sc["K"], comm)
    d_cam0, d_pts0 = device.to_device(cams0), device.to_device(pts0[prob.p_lo:prob.p_hi])
    d_cam, d_pts = torch.empty_like(d_cam0), torch.empty_like(d_pts0)
    for _ in range(2):
        d_cam.copy_(d_cam0); d_pts.copy_(d_pts0)
        prob.solve_device(d_cam, d_pts)
    dist.barrier(); torch.cuda.synchronize()
    l0 = _lib.launch_count()
    ev, infos = [], []
    for _ in range(steps):
        flush.zero_()
        d_cam.copy_(d_cam0); d_pts.copy_(d_pts0)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        infos.append(prob.solve_device(d_cam, d_pts))
        b.record()
        ev.append((a, b))
    dist.barrier(); torch.cuda.synchronize()
    launches = int(_lib.launch_count() - l0)
    t = torch.tensor([sum(a.elapsed_time(b) for a, b in ev)], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    # one more solve with the phase profile on (not timed)
    d_cam.copy_(d_cam0); d_pts.copy_(d_pts0)
    prob.solve_device(d_cam, d_pts, profile=True)
    phases = bundle.last_profile()
    its = sum(i["n_linearisations"] for i in infos)
    n5 = len(sc["obs"])
    value = its / (ms * 1e-3)
    out = {"metric": "BA LM iterations/sec", "unit": "LM iterations/s", "dtype": "f64", "scaling": "strong",
           "value": value, "ms_per_solve": ms / steps, "solves": steps, "n_gpus": world,
           "config": {"workload": "C5 synthetic BAL-scale BA sharded by point: " + _name(C5),
                      "points_per_rank": prob.n_pts, "observations_per_rank": prob.n_obs,
                      "collectives_per_trial": "all-reduce of the blocks of S some rank touches + g_red (bytes below), "
                                               "6 scalars, gmax; per linearisation U|g_c|cost",
                      "s_allreduce_bytes_per_trial": phases["s_allreduce_bytes_per_trial"]},
           "lm_iterations_per_solve": its / steps, "cost": infos[-1]["cost"],
           "rmse_px": float(np.sqrt(2 * infos[-1]["cost"] / n5)), "status": infos[-1]["status"],
           "gpu_launches": launches, "phase_ms_per_solve_rank0": phases}
    if rank == 0:
        cams = d_cam.cpu().numpy()
        pts_local = d_pts.cpu().numpy()
        out["n1_value"] = n1["value"]
        out["n1_ms_per_solve"] = n1["ms_per_solve"]
        out["efficiency_vs_n1"] = value / (world * n1["value"])
        out["speedup_vs_n1"] = value / n1["value"]
        out["parity_vs_n1"] = {"cost_rel": abs(infos[-1]["cost"] - n1["cost"]) / n1["cost"],
                               "cameras_max_abs": float(np.max(np.abs(cams - n1["cams"]))),
                               "points_rank0_max_abs": float(np.max(np.abs(pts_local - n1["pts"][prob.p_lo:prob.p_hi]))),
                               "what": "the sharded solve against rank 0 solving the whole problem alone in this run"}
    comm.destroy()
    return out
