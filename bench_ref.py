"""bench_ref.py -- the reference arm of bench.py (`--impl reference`): the reference's own CPU implementation of the
hot path, timed on the host cores of the box. Never imports the product package (no libsfm_b200.so in this process).

What is timed: the UNMODIFIED reference staged under oracle/_ref by oracle/stage_ref.py (kind "reference") when it
travelled with the snapshot, else the oracle port (kind "port"). The reference is single-threaded; "all cores" means
one process per core, each replaying its own consecutive piece of the ONE seeded `random` stream the GPU arm scores.
A step is a bounded sample of the workload (the full C2 step is ~190 s of one core), so the arm is rate-normalised:
same metric, unit and inputs as the GPU arm, fewer hypotheses per step -- stated in config.sample.
"""
from __future__ import annotations

import io
import json
import os
import random
import time
from contextlib import redirect_stdout

import numpy as np

import bench_config as cfg

_REF = {}


def _ref_modules():
    """Staged reference modules of this process (loaded once), or None when oracle/_ref did not travel."""
    if "m" not in _REF:
        from oracle import stage_ref
        _REF["m"] = stage_ref.load_modules() if stage_ref.available() else None
    return _REF["m"]


def kind():
    from oracle import stage_ref
    return "reference" if stage_ref.available() else "port"


def _f_chunk(args):
    """One worker: `n_hyp` hypotheses of the stream from RNG state `state`, all correspondences."""
    x1, x2, state, n_hyp, thr = args
    os.environ.setdefault("OMP_NUM_THREADS", "1")
    random.setstate(state)
    m = _ref_modules()
    t = time.perf_counter()
    if m is not None:
        with redirect_stdout(io.StringIO()):
            inl, _, _ = m["GetInliersRANSAC"].GetInliersRANSAC(x1, x2, np.arange(len(x1)), thr, n_hyp)
        best = len(inl)
    else:
        from oracle import sfm_oracle as so
        best = int(so.f_ransac_detailed(x1, x2, thr, so.draw_samples(len(x1), 8, n_hyp))["best_count"])
    return time.perf_counter() - t, best


def f_ransac_leg(args, pool, cores):
    per_core = 48                                   # ~0.1 s per core per step at 1.9 ms/hypothesis
    h_step = per_core * cores
    sc = cfg.two_view_scene()
    n_steps = args.steps + args.warmup
    states = cfg.chunk_states(n_steps * cores, per_core)

    def step(i):
        jobs = [(sc["x1"], sc["x2"], states[i * cores + c], per_core, cfg.THRESH) for c in range(cores)]
        t = time.perf_counter()
        res = pool.map(_f_chunk, jobs)
        return time.perf_counter() - t, max(r[1] for r in res)
    for i in range(args.warmup):
        step(i)
    dts = [step(args.warmup + i)[0] for i in range(args.steps)]
    total = sum(dts)
    sample = (f"each step = {h_step} consecutive hypotheses of the seeded stream ({per_core} per core, {cores} "
              f"single-threaded processes) over all {cfg.N_PTS} correspondences; the GPU arm's step is "
              f"{cfg.N_HYP} hypotheses of the same stream: rate-normalised, not the same step size")
    return h_step * args.steps / total, 1e3 * total / args.steps, sample


# ------------------------------------------------------------------------------------------------
# BA leg: scipy least_squares(method='lm', 2-point Jacobian) as BundleAdjustment.py:205-212 calls it
# ------------------------------------------------------------------------------------------------
def _ba_case(args):
    """One bundle adjustment on one core. use_ref: the shipped per-observation Python residual
    (BundleAdjustment.py:43-110) instead of the oracle's vectorised restatement of it."""
    nc, npts, k, seed, max_nfev, use_ref = args
    os.environ.setdefault("OMP_NUM_THREADS", "1")
    from scipy.optimize import least_squares
    from scipy.spatial.transform import Rotation
    from oracle import ba_oracle as bo
    from oracle import scenes
    sc = scenes.ba_scene(n_cams=nc, n_pts=npts, obs_per_pt=k, seed=seed)
    cams0 = np.hstack([Rotation.from_matrix(sc["R0"]).as_rotvec(), -np.einsum("nij,nj->ni", sc["R0"], sc["C0"])])
    x0 = np.concatenate([cams0.reshape(-1), sc["pts0"].reshape(-1)])
    m = _ref_modules() if use_ref else None
    fn = m["BundleAdjustment"].bundle_adjustment_residuals if m is not None else bo.ba_residuals
    fargs = (nc, npts, sc["cam_idx"], sc["pt_idx"], sc["obs"], sc["K"])
    calls = [0]

    def counted(p, *a):
        calls[0] += 1
        return fn(p, *a)
    t = time.perf_counter()
    res = least_squares(counted, x0, method="lm", max_nfev=max_nfev, args=fargs)
    dt = time.perf_counter() - t
    nobs, nvar = len(sc["obs"]), len(x0)
    n_jac = max(1, (calls[0] - int(res.nfev)) // nvar)         # every Jacobian = nvar forward-difference calls
    return dict(cameras=nc, points=npts, observations=nobs, parameters=nvar, seconds=dt, residual_calls=calls[0],
                jacobians=n_jac, trial_steps=int(res.nfev), cost=float(res.cost), status=int(res.status),
                kind="reference" if m is not None else "port",
                s_per_lm_iteration=dt / n_jac, c_us=1e6 * dt / n_jac / (nobs * nvar))


def ba_leg(pool):
    """BA LM iterations/s of the reference at the three sizes SURVEY 8(d) names, the fitted cost model
    t_iter = c * Nobs * (6 Nc + 3 Np) and its extrapolation to C3 / C5 (labelled as such: the reference cannot
    run those sizes -- 193 GB dense Jacobian at C3)."""
    sizes = [(5, 60), (6, 120), (10, 300)]
    jobs = []
    if kind() == "reference":                     # shipped residual: a Python loop per observation, 25 us each
        jobs += [(5, 60, 4, 1, 100, True), (6, 120, 4, 2, 2, True), (10, 300, 4, 3, 1, True)]
    jobs += [(nc, npts, 4, 1 + i, 100, False) for i, (nc, npts) in enumerate(sizes)]
    runs = pool.map(_ba_case, jobs, chunksize=1)
    out = {"metric": "BA LM iterations/sec", "unit": "LM iterations/s", "cores": 1,
           "what": "scipy least_squares(method='lm', max_nfev as listed) with the 2-point finite-difference Jacobian, "
                   "one LM iteration = one Jacobian (n_params residual calls); one core per problem (the reference "
                   "is single-threaded)", "runs": runs}
    for tag in ("reference", "port"):
        rs = [r for r in runs if r["kind"] == tag]
        if not rs:
            continue
        c = float(np.mean([r["c_us"] for r in rs]))           # us per (observation x parameter) per Jacobian
        def ext(cfgd):
            nobs = cfgd["n_pts"] * cfgd["obs_per_pt"]
            nvar = 6 * cfgd["n_cams"] + 3 * cfgd["n_pts"]
            return 1.0 / (c * 1e-6 * nobs * nvar)
        out[tag] = {"fit_c_us_per_obs_param": c, "c_per_size": [r["c_us"] for r in rs],
                    "measured_lm_iterations_s": [1.0 / r["s_per_lm_iteration"] for r in rs],
                    "extrapolated_lm_iterations_s": {"C3": ext(cfg.C3), "C5": ext(cfg.C5)},
                    "note": "extrapolation by t_iter = c * Nobs * (6 Nc + 3 Np); NOT a measurement"}
    tag = "reference" if "reference" in out else "port"
    last = [r for r in runs if r["kind"] == tag][-1]
    out["value"] = 1.0 / last["s_per_lm_iteration"]
    out["config"] = {"workload": f"synthetic BA at the largest size timed on the host ({tag}): {last['cameras']} cameras, "
                                 f"{last['points']} points, {last['observations']} observations"}
    return out


# ------------------------------------------------------------------------------------------------
# C4 leg: triangulation / PnP with one process per core
# ------------------------------------------------------------------------------------------------
def _c4_tri(args):
    K, C1, R1, C2, R2, x1, x2 = args
    os.environ.setdefault("OMP_NUM_THREADS", "1")
    m = _ref_modules()
    from oracle import sfm_oracle as so
    with redirect_stdout(io.StringIO()):
        t = time.perf_counter()
        X0 = (m["LinearTriangulation"].LinearTriangulation if m else so.linear_triangulation)(K, C1, R1, C2, R2, x1, x2)
        t_lin = time.perf_counter() - t
        X0 = np.asarray(X0)[:, :3]
        t = time.perf_counter()
        (m["NonLinearTriangulation"].NonLinearTriangulation if m else so.nonlinear_triangulation)(K, C1, R1, C2, R2, x1, x2, X0)
        t_nl = time.perf_counter() - t
    return t_lin, t_nl


def _c4_pnp(args):
    seed, n_corr, n_reg = args
    os.environ.setdefault("OMP_NUM_THREADS", "1")
    m = _ref_modules()
    from oracle import scenes
    from oracle import sfm_oracle as so
    random.seed(777 + seed)
    t_r = t_n = 0.0
    with redirect_stdout(io.StringIO()):
        for b in range(n_reg):
            s = scenes.pnp_scene(n=n_corr, seed=20260105 + seed * n_reg + b)
            t = time.perf_counter()
            C, R = (m["PnPRANSAC"].PnPRANSAC if m else so.pnp_ransac)(s["X"], s["x"], s["K"], 200, 1000)
            t_r += time.perf_counter() - t
            t = time.perf_counter()
            (m["NonlinearPnP"].NonLinearPnP if m else so.nonlinear_pnp)(s["X"], s["x"], s["K"], C, R)
            t_n += time.perf_counter() - t
    return t_r, t_n


def c4_leg(pool, cores, pts_per_core=300, n_corr=2000, reg_per_core=1):
    from oracle import scenes
    sc = scenes.two_view(n=pts_per_core * cores, outlier_frac=0.0, noise=0.5, seed=20260104)
    jobs = [(sc["K"], sc["C1"], sc["R1"], sc["C2"], sc["R2"], sc["x1"][c * pts_per_core:(c + 1) * pts_per_core],
             sc["x2"][c * pts_per_core:(c + 1) * pts_per_core]) for c in range(cores)]
    t = time.perf_counter()
    tri = pool.map(_c4_tri, jobs, chunksize=1)
    wall_tri = time.perf_counter() - t
    t = time.perf_counter()
    pnp = pool.map(_c4_pnp, [(c, n_corr, reg_per_core) for c in range(cores)], chunksize=1)
    wall_pnp = time.perf_counter() - t
    n_pts, n_reg = pts_per_core * cores, reg_per_core * cores
    lin1, nl1 = np.mean([a for a, _ in tri]), np.mean([b for _, b in tri])
    r1, p1 = np.mean([a for a, _ in pnp]), np.mean([b for _, b in pnp])
    return {"kind": kind(), "cores": cores,
            "sample": f"{n_pts} points of the C4 two-view scene ({pts_per_core} per core), {n_reg} registrations of "
                      f"{n_corr} correspondences (PnPRANSAC threshold 200, 1000 hypotheses, then NonLinearPnP)",
            "all_cores": {"triangulation_linear_plus_nonlinear_points_s": n_pts / wall_tri,
                          "pnp_ransac_plus_nonlinear_registrations_s": n_reg / wall_pnp},
            "one_core": {"lintri_points_s": pts_per_core / lin1, "nltri_points_s": pts_per_core / nl1,
                         "pnp_ransac_registrations_s": reg_per_core / r1, "nlpnp_registrations_s": reg_per_core / p1}}


def run(args, rank, world):
    if rank != 0:
        return
    import multiprocessing as mp
    cores = os.cpu_count() or 1
    ctx = mp.get_context("fork")
    with ctx.Pool(cores) as pool:
        value, ms_per_step, sample = f_ransac_leg(args, pool, cores)
        extra = {}
        if not args.skip_ba:
            extra["ba"] = ba_leg(pool)
            extra["c4_batched_solvers"] = c4_leg(pool, cores)
    import platform
    import scipy
    line = {
        "impl": "reference", "metric": cfg.METRIC, "value": value, "unit": cfg.UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": cfg.WORKLOAD, "threshold": cfg.THRESH, "sample": sample, "same_step_size": False},
        "cpu_baseline": {"value": value, "unit": cfg.UNIT, "cores": cores, "kind": kind(), "sample": sample},
        "e2e": {"value": value, "unit": cfg.UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "host": {"cpu_count": cores, "machine": platform.machine(), "python": platform.python_version(),
                 "numpy": np.__version__, "scipy": scipy.__version__,
                 "threads_per_process": os.environ.get("OMP_NUM_THREADS", "unset (workers set 1)")}}
    line.update(extra)
    print(json.dumps(line), flush=True)
