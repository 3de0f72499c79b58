#!/usr/bin/env python
"""K2 (consensus scoring) at the C2 shape: the certified-FP32 kernel against the all-FP64 kernel.
Prints one JSON line: identical counts?, time of each (CUDA events, L2 flushed between launches),
and the fraction of point-hypothesis tests the filter handed to the FP64 path."""
import ctypes as C
import json
import os
import random
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from sfm_b200 import _lib, device, sampler
from oracle import scenes

lib = _lib.load()
n, h = 50_000, int(os.environ.get("K2_H", 100_000))
out = {}
for thr in (0.06, 1.0):
    sc = scenes.two_view()
    random.seed(12345)
    samples = sampler.sample_stream(n, 8, h)
    d_x = device.to_device(np.hstack([sc["x1"], sc["x2"]]))
    d_s = device.to_device(samples, np.int32)
    d_F = torch.empty((h, 9), dtype=torch.float64, device="cuda")
    assert lib.sfm_f8pt_batch(d_x.data_ptr(), n, d_s.data_ptr(), h, d_F.data_ptr(), None) == 0
    c_mixed = torch.empty(h, dtype=torch.int32, device="cuda")
    c_f64 = torch.empty(h, dtype=torch.int32, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def timed(fn, dst, reps=5):
        for _ in range(2):
            assert fn(d_x.data_ptr(), n, d_F.data_ptr(), h, thr, dst.data_ptr(), None) == 0
        torch.cuda.synchronize()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
        for a, b in ev:
            flush.zero_()
            a.record()
            assert fn(d_x.data_ptr(), n, d_F.data_ptr(), h, thr, dst.data_ptr(), None) == 0
            b.record()
        torch.cuda.synchronize()
        return sum(a.elapsed_time(b) for a, b in ev) / reps

    v = C.c_uint64(0)
    ms_f64 = timed(lib.sfm_f_score_count_fp64, c_f64)
    lib.sfm_f_score_fp64_tests(C.byref(v), 1)
    assert lib.sfm_f_score_count(d_x.data_ptr(), n, d_F.data_ptr(), h, thr, c_mixed.data_ptr(), None) == 0
    lib.sfm_f_score_fp64_tests(C.byref(v), 1)
    ms_mixed = timed(lib.sfm_f_score_count, c_mixed)
    out[f"thr_{thr}"] = {
        "identical_counts": bool(torch.equal(c_mixed, c_f64)),
        "mismatches": int((c_mixed != c_f64).sum().item()),
        "ms_fp64": ms_f64, "ms_mixed": ms_mixed,
        "hyp_per_s_fp64": h / (ms_f64 * 1e-3), "hyp_per_s_mixed": h / (ms_mixed * 1e-3),
        "fp64_test_fraction": v.value / (n * h),
        "max_count": int(c_f64.max().item()),
    }
print(json.dumps(out))
