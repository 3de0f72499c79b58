#!/usr/bin/env python
"""bench.py -- headline benchmark: RANSAC F-hypotheses/s on BASELINE.json configs[1]
(synthetic two-view F-RANSAC: 50k correspondences, 30% outliers, 100k hypotheses per GPU).

    python bench.py --gpus N --steps K --warmup W            (our arm; torchrun for N > 1)
    python bench.py --impl reference --gpus N --steps K ...  (reference arm: CPU oracle port)

One step = one full RANSAC over the per-GPU batch: K1 8-point hypotheses, K2 consensus
scoring, K3 first-max argmax, K4 winner mask (+ one 8-byte NCCL max-allreduce for N > 1).
Prints ONE JSON line (rank 0). See DESIGN.md "Measurement" for every field.
"""
from __future__ import annotations

import argparse
import json
import os
import random
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from bench_config import DATA_SEED, METRIC, N_HYP, N_PTS, RNG_SEED, THRESH, UNIT, WORKLOAD  # noqa: E402


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f), "measured (MEASURED_PEAKS.json)"
    return {"hbm_gbs": 6650.0}, "fallback (B200_PROFILING.md)"


def _ncu_traffic(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel` at the bench shape, from the
    committed ncu --set full capture (profiles/r02_ncu_traffic.json, written by tools/ncu_traffic.py)."""
    try:
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "r02_ncu_traffic.json")) as f:
            return float(json.load(f)[kernel]["dram_bytes"])
    except Exception:
        return None


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed regions (every ~10 ms through NVML; nvidia-smi
    every 0.2 s if the NVML binding is unavailable)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index=0):
        self.samples, self.stop_evt, self.index = [], threading.Event(), index
        self.sm_max, self.how = None, "nvidia-smi"
        self.thread = threading.Thread(target=self._run, daemon=True)

    def _run_nvml(self):
        import pynvml as nv
        nv.nvmlInit()
        h = nv.nvmlDeviceGetHandleByIndex(self._physical_index())
        self.sm_max = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
        bits = [nv.nvmlClocksThrottleReasonHwSlowdown, nv.nvmlClocksThrottleReasonHwThermalSlowdown,
                nv.nvmlClocksThrottleReasonSwThermalSlowdown, nv.nvmlClocksThrottleReasonSwPowerCap]
        self.how = "nvml"
        while not self.stop_evt.is_set():
            mhz = float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
            r = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(h))
            self.samples.append((mhz, [bool(r & b) for b in bits]))
            self.stop_evt.wait(0.01)

    def _physical_index(self):
        vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
        ids = [v for v in vis.split(",") if v.strip()]
        if ids and all(v.strip().isdigit() for v in ids) and self.index < len(ids):
            return int(ids[self.index])
        return self.index

    def _run(self):
        try:
            self._run_nvml()
            return
        except Exception:
            self.how = "nvidia-smi"
        while not self.stop_evt.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--id={self._physical_index()}", f"--query-gpu={self.Q}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True,
                                     timeout=5).stdout.strip()
                if out:
                    c = [v.strip() for v in out.splitlines()[0].split(",")]
                    if c[0].replace(".", "").isdigit():
                        if c[1].replace(".", "").isdigit():
                            self.sm_max = float(c[1])
                        self.samples.append((float(c[0]), [v.lower().startswith("active") for v in c[3:7]]))
            except Exception:
                pass
            self.stop_evt.wait(0.2)

    def __enter__(self):
        self.thread.start()
        return self

    def __exit__(self, *a):
        self.stop_evt.set()
        self.thread.join(timeout=6)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["clock query unavailable"]}
        sm = sorted(s[0] for s in self.samples)
        reasons = [n for i, n in enumerate(self.NAMES) if any(s[1][i] for s in self.samples)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_min_mhz": sm[0], "sm_max_mhz": self.sm_max,
                "reasons": reasons, "samples": len(self.samples), "how": self.how,
                "covers": "every timed region of this line (F-RANSAC steps, K2 alone, end-to-end, BA)"}


def make_inputs(n_hyp_total):
    """The C2 scene and the reference's sample stream (random.sample from RNG_SEED, replayed bit-exactly by the
    library's MT19937 sampler: tests/test_abi_and_sampler.py)."""
    import bench_config
    from sfm_b200 import sampler
    sc = bench_config.two_view_scene()
    random.seed(RNG_SEED)
    samples = sampler.sample_stream(N_PTS, 8, n_hyp_total)
    return sc, samples


# ------------------------------------------------------------------------------------------
# CPU side: the oracle port of GetInliersRANSAC.py:53-92 (per-hypothesis numpy, like the reference)
# ------------------------------------------------------------------------------------------
def _cpu_chunk(args):
    x1, x2, samples, thr = args
    os.environ.setdefault("OMP_NUM_THREADS", "1")
    from oracle import sfm_oracle as so
    t = time.perf_counter()
    d = so.f_ransac_detailed(x1, x2, thr, samples)
    return time.perf_counter() - t, int(d["best_count"])


def cpu_baseline_single(sc, samples, n_hyp=4000):
    """Rank 0, N=1 only: the oracle port on ONE core (the reference is single-threaded)."""
    dt, _ = _cpu_chunk((sc["x1"], sc["x2"], samples[:n_hyp], THRESH))
    return {"value": n_hyp / dt, "unit": UNIT, "cores": 1, "kind": "port",
            "sample": f"first {n_hyp} hypotheses of the same stream over all {N_PTS} correspondences "
                      f"({dt:.1f} s)"}


def _oracle_counts_chunk(args):
    """Checker leg: the oracle's per-hypothesis inlier counts and F for a chunk of the stream."""
    x1, x2, samples, thr = args
    os.environ.setdefault("OMP_NUM_THREADS", "1")
    from oracle import sfm_oracle as so
    d = so.f_ransac_detailed(x1, x2, thr, samples)
    return np.asarray(d["counts"], np.int64), np.asarray(d["F"], np.float64).reshape(len(samples), 9)


def parity_full_stream(pool, cores, sc, samples, gpu_counts, gpu_F, gpu_best, gpu_mask, gpu_Fbest):
    """Benchmark-size parity (N = 1): the oracle (checker, not the thing measured) over the WHOLE hypothesis
    stream the timed steps scored, on all host cores; compared with what the GPU produced for that stream."""
    from oracle import sfm_oracle as so
    t0 = time.perf_counter()
    h = len(samples)
    per = max(64, -(-h // (cores * 8)))
    jobs = [(sc["x1"], sc["x2"], samples[a:a + per], THRESH) for a in range(0, h, per)]
    res = pool.map(_oracle_counts_chunk, jobs, chunksize=1)
    counts = np.concatenate([r[0] for r in res])
    F = np.concatenate([r[1] for r in res])
    dt = time.perf_counter() - t0
    best = int(np.argmax(counts))                      # first maximum == the reference's strict '>' rule
    err = so.f_errors(F[best].reshape(3, 3), sc["x1"], sc["x2"])
    band = np.abs(err - THRESH) < 1e-9
    mask_ref = err < THRESH
    mism = np.nonzero(counts != gpu_counts)[0]
    explained = 0
    for k in mism[:200]:                               # a count may differ only through points in the 1e-9 band
        e = so.f_errors(F[k].reshape(3, 3), sc["x1"], sc["x2"])
        explained += int(abs(int(counts[k]) - int(gpu_counts[k])) <= int((np.abs(e - THRESH) < 1e-9).sum()))
    nrm = np.linalg.norm(F, axis=1)
    rel = np.linalg.norm(gpu_F - F, axis=1) / np.where(nrm > 0, nrm, 1.0)
    scoring = counts > 0                               # hypotheses that can win at all
    return {"checker": "oracle/sfm_oracle.f_ransac_detailed over the whole stream", "hypotheses_checked": int(h),
            "oracle_seconds": dt, "oracle_cores": cores, "oracle_hypotheses_s": h / dt,
            "winner_equal": bool(best == int(gpu_best[0]) and int(counts[best]) == int(gpu_best[1])),
            "winner": [best, int(counts[best])], "count_mismatches": int(len(mism)),
            "count_mismatches_explained_by_band": int(explained),
            "band_population": int(band.sum()),
            "mask_equal_outside_1e-9": bool(np.array_equal(gpu_mask[~band].astype(bool), mask_ref[~band])),
            "rel_F_winner": float(np.linalg.norm(gpu_Fbest.ravel() - F[best]) / np.linalg.norm(F[best])),
            "frac_hypotheses_rel_F_below_1e-6": float(np.mean(rel < 1e-6)),
            "max_rel_F_over_scoring_hypotheses": float(rel[scoring].max()) if scoring.any() else None,
            "note": "F of a (near-)degenerate 8-point sample is an ill-conditioned null vector: hypotheses whose F "
                    "differs beyond 1e-6 are listed by fraction; counts are compared for ALL hypotheses"}


def run_reference(args, rank, world):
    """--impl reference: see bench_ref.py (the reference's own CPU code from oracle/_ref, else the oracle port, on
    all host cores; this process never imports the product package)."""
    import bench_ref
    bench_ref.run(args, rank, world)


def run_ours(args, rank, world):
    import torch
    import torch.distributed as dist
    from sfm_b200 import _lib, device, ransac

    local = int(os.environ.get("LOCAL_RANK", 0))
    # the checker's process pool is forked BEFORE this process touches CUDA; it stays idle during every timed region
    cores = os.cpu_count() or 1
    pool = None
    if world == 1 and not args.skip_parity:
        import multiprocessing as mp
        pool = mp.get_context("fork").Pool(cores)
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    lib = _lib.load()
    sc, samples_all = make_inputs(N_HYP * world)
    x1x2 = np.ascontiguousarray(np.hstack([sc["x1"], sc["x2"]]))
    d_x = device.to_device(x1x2)
    d_s_all = device.to_device(samples_all, np.int32)
    d_s = d_s_all[rank * N_HYP:(rank + 1) * N_HYP]
    plan = ransac.FRansacPlan(N_PTS, N_HYP)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")     # > 126 MB L2

    def step():
        if world > 1:
            return plan.run_sharded(d_x, d_s_all, THRESH, rank, world)
        return plan.run(d_x, d_s, THRESH)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        flush.zero_()
        step()
    barrier()
    launches0 = _lib.launch_count()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    clocks = ClockSampler(local).__enter__()   # stopped right before the line is printed
    barrier()
    for a, b in ev:
        flush.zero_()                      # L2 flush between timed iterations (outside the events)
        a.record()
        r = step()
        b.record()
    barrier()
    launches = _lib.launch_count() - launches0
    ms = sum(a.elapsed_time(b) for a, b in ev)
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    value = N_HYP * world * args.steps / (ms_total * 1e-3)

    # dominant kernel (K2 consensus scoring) timed alone on the launching stream
    k2 = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    st = device.stream()
    for a, b in k2:
        flush.zero_()
        a.record()
        _lib.check(lib.sfm_f_score_count(device.ptr(d_x), N_PTS, device.ptr(plan.r.F), N_HYP, THRESH,
                                         device.ptr(plan.r.counts), st), "sfm_f_score_count")
        b.record()
    torch.cuda.synchronize()
    k2_ms = sum(a.elapsed_time(b) for a, b in k2) / args.steps
    peaks, peak_src = measured_peaks()
    alg_bytes = (32.0 * N_PTS + 76.0) * N_HYP
    achieved = alg_bytes / (k2_ms * 1e-3) / 1e9
    tests = float(N_PTS) * N_HYP
    fma_peak = 148 * 128 * 1.965e9          # FP32 FMA lanes x clock: 119-125 of 128 per clock and SM measured
    # executed packed-FP32 lane operations per test, counted by ncu (profiles/r02_k2_two_stage_ncu_full.txt): 13 when
    # the group's first-stage vote passes (86 % of the groups at C2), the repeated stage, F^T x2 and the second norm
    # (+20) for the groups that fail it, the square-root path for the ~6 % of pairs that fail the second stage
    ops_per_test = 14.16
    lane_ops = ops_per_test * tests / (k2_ms * 1e-3)
    roofline = {"bound": "fp32_pipe", "kernel": "f_score_packed_kernel (K2)", "achieved": lane_ops / 1e12,
                "peak": fma_peak / 1e12, "unit": "T FMA-lane-op/s", "frac": lane_ops / fma_peak,
                "traffic": _ncu_traffic("f_score_packed_kernel"), "kernel_ms": k2_ms,
                "peak_source": "148 SMs x 128 FP32 lanes x 1.965 GHz (max SM clock); profiles/r01_ubench_fp32_b200.txt "
                               "measures 119 (FFMA) and 125 (FFMA2, broadcast operands) of 128 lanes per clock and SM",
                "what": "certified FP32 filter on packed FFMA2/FMUL2/FADD2 with a certified, sqrt-free, TWO-STAGE early "
                        "reject (round 2): 13 FMA-lane operations per (hypothesis, correspondence) test when the warp "
                        "certifies a group of 4 correspondences as gross outliers from F x1 and one line norm alone (86 % "
                        "of the groups at 0.06 px), +6 for F^T x2 and the second norm when that vote fails, 26 + 2 "
                        "MUFU.SQRT on the full path (round 1: 23 + 2 MUFU for every test); 14.16 executed on average "
                        "(ncu); 2e-4 of the tests are redone in FP64; tests per launch = N x H. frac counts the "
                        "operations EXECUTED, so a kernel that does less work per test shows a similar fraction at a "
                        "higher hypotheses/s",
                "fma_lane_ops_per_test": ops_per_test,
                "tests_per_launch": tests,
                "why_not_hbm": "the 1.6 MB correspondence set is staged in shared memory and re-read by every "
                               "hypothesis tile: DRAM traffic per launch (`traffic`, ncu) is ~1e-4 of the algorithmic "
                               "bytes below, so HBM cannot be the bound",
                "hbm_algorithmic": {"bytes_per_launch": alg_bytes, "achieved_gbs": achieved, "peak_gbs": peaks["hbm_gbs"],
                                    "ratio": achieved / peaks["hbm_gbs"], "peak_source": peak_src,
                                    "note": "(32 B x N + 76 B) per hypothesis (SURVEY 8d): what a kernel without "
                                            "on-chip reuse would stream; ratio > 1 is the reuse factor, NOT a "
                                            "roofline fraction"}}

    # end to end through the drop-in signature: host numpy in, host numpy out, RNG replay, H2D, D2H
    index = np.arange(N_PTS)
    e2e_steps = max(1, min(args.steps, 5))
    random.seed(RNG_SEED)
    ransac.GetInliersRANSAC(sc["x1"], sc["x2"], index, THRESH, N_HYP)       # warm
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        inl, outl, F = ransac.GetInliersRANSAC(sc["x1"], sc["x2"], index, THRESH, N_HYP)
    torch.cuda.synchronize()
    te = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e = {"value": N_HYP * world * e2e_steps / float(te.item()), "unit": UNIT,
           "h2d_bytes_per_step": int(x1x2.nbytes + N_HYP * 8 * 4), "d2h_bytes_per_step": 80 + N_PTS,
           "api": "sfm_b200.ransac.GetInliersRANSAC(points1, points2, index, 0.06, 100000) per GPU "
                  "(host numpy in/out; includes the MT19937 sample replay)",
           "steps": e2e_steps}

    best = r.best.cpu().numpy()
    parity = None
    if world == 1 and pool is not None:
        parity = parity_full_stream(pool, cores, sc, samples_all, plan.r.counts.cpu().numpy().astype(np.int64),
                                    plan.r.F.cpu().numpy(), best, plan.r.mask.cpu().numpy(),
                                    plan.r.F_best.cpu().numpy())
        pool.close()
        pool.join()
    elif world > 1:
        # the sharded run against ONE GPU scoring the whole stream (rank 0, outside every timed region)
        if rank == 0:
            whole = ransac.FRansacPlan(N_PTS, N_HYP * world)
            r1 = whole.run(d_x, d_s_all, THRESH)
            b1 = r1.best.cpu().numpy()
            parity = {"checker": "the same stream scored by one GPU (sfm_f_ransac over all N x 100k hypotheses)",
                      "parity_vs_n1": {"winner_equal": bool(b1[0] == best[0] and b1[1] == best[1]),
                                       "winner_n1": [int(b1[0]), int(b1[1])],
                                       "mask_equal": bool(torch.equal(r1.mask, r.mask)),
                                       "F_best_bit_equal": bool(torch.equal(r1.F_best, r.F_best))}}
            del whole, r1
        dist.barrier()
    ba_mgpu = None
    if world > 1 and not args.skip_ba:
        import bench_ba
        ba_mgpu = bench_ba.run_sharded(args, rank, world, flush)
    if rank == 0:
        out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
               "warmup": max(args.warmup, 3), "ms_per_step": ms_total / args.steps,
               "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
               "data": "synthetic",
               "config": {"workload": WORKLOAD, "threshold": THRESH, "n_correspondences": N_PTS,
                          "hypotheses_per_gpu": N_HYP, "l2": "256 MiB flush write between timed steps",
                          "winner": [int(best[0]), int(best[1])]},
               "roofline": roofline, "e2e": e2e, "gpu_launches": int(launches),
               "clocks": clocks.summary()}
        if parity is not None:
            out["parity"] = parity
        if world == 1:
            out["cpu_baseline"] = cpu_baseline_single(sc, samples_all, n_hyp=args.cpu_hyp)
            if not args.skip_ba:
                import bench_ba
                del d_x, d_s_all, d_s, plan
                out["ba"] = bench_ba.run_single_gpu(args, peaks, peak_src, flush)
                out["ba"]["cpu_baseline"] = bench_ba.cpu_baseline()
                out["gpu_launches"] += out["ba"]["gpu_launches"]
                import bench_pipeline
                before = _lib.launch_count()
                out["c1_pipeline"] = bench_pipeline.run()
                out["c1_pipeline"]["gpu_launches"] = int(_lib.launch_count() - before)
                import bench_c4
                before = _lib.launch_count()
                out["c4_batched_solvers"] = bench_c4.run(flush)
                out["c4_batched_solvers"]["gpu_launches"] = int(_lib.launch_count() - before)
        if ba_mgpu is not None:
            out["ba"] = ba_mgpu
            out["gpu_launches"] += ba_mgpu["gpu_launches"]
        clocks.__exit__()
        out["clocks"] = clocks.summary()
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cpu-hyp", type=int, default=6000, help="hypotheses in the 1-core CPU baseline sample")
    ap.add_argument("--skip-ba", action="store_true", help="F-RANSAC part only")
    ap.add_argument("--skip-parity", action="store_true", help="no oracle pass over the whole hypothesis stream")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    # rank 0's stdout carries ONE JSON line: NCCL's "NCCL version ..." banner (NCCL_DEBUG=VERSION, set by
    # some launch environments) would otherwise land in front of it
    if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
        os.environ["NCCL_DEBUG"] = "WARN"
    # ... and anything else a library writes to file descriptor 1 goes to stderr; print() below still reaches
    # the real stdout through the saved descriptor
    sys.stdout.flush()
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    sys.stdout = real_stdout
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        if world != args.gpus and world == 1 and args.gpus > 1:
            raise SystemExit("launch with torch.distributed.run --nproc-per-node N for --gpus N > 1")
        run_ours(args, rank, world)


if __name__ == "__main__":
    main()
