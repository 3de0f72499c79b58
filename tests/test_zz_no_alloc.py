"""GPU test of the allocation-free entry point sfm_fundamental_npt_static (include/sfm_b200.h): no counterpart
in the reference, added for callers that cannot allocate device memory.

Collected LAST on purpose (file name): a failure here hides nothing.
The variant launches the same kernel as sfm_fundamental_npt with its scratch in a file-scope device array, so
its output is compared with the existing entry point's on the same inputs (tolerance of the other F tests,
1e-6 relative; the two are in fact expected to agree to the last bit) on two calls in a row of DIFFERENT sizes
-- the second, smaller call runs over the stale contents the first one left in the shared array."""
import numpy as np
import pytest

from conftest import golden, rel_err

pytestmark = pytest.mark.gpu

F_RTOL = 1e-6
MAX_N = 16384          # SFM_FUNDAMENTAL_NPT_STATIC_MAX_N


def _reference_call(lib, device, torch, d_x1, d_x2, n):
    from sfm_b200 import _lib
    d_F = device.empty((9,), torch.float64)
    nbytes = int(lib.sfm_fundamental_npt_workspace_bytes(n))
    ws = device.empty((nbytes,), torch.uint8)
    _lib.check(lib.sfm_fundamental_npt(device.ptr(d_x1), device.ptr(d_x2), n, device.ptr(d_F), device.ptr(ws),
                                       nbytes, device.stream()), "sfm_fundamental_npt")
    return d_F.cpu().numpy()


def test_static_scratch_variant_matches_workspace_variant(cuda_device):
    import torch
    from sfm_b200 import _lib, device
    from oracle import scenes
    lib = _lib.load()
    g = golden("f8pt.npz")
    sc = scenes.two_view(n=5000, outlier_frac=0.0, seed=31)
    cases = [(sc["x1"], sc["x2"]),                                   # 5000 correspondences first ...
             (np.asarray(g["x1"], float), np.asarray(g["x2"], float))]   # ... then the small golden set
    # everything the timed-style calls touch exists before them: inputs, outputs
    bufs = [(device.to_device(a), device.to_device(b), device.empty((9,), torch.float64)) for a, b in cases]
    want = [_reference_call(lib, device, torch, d1, d2, len(a)) for (a, _), (d1, d2, _) in zip(cases, bufs)]
    # first call on this device: looks the scratch address up (set-up), allocates nothing either
    d1, d2, dF = bufs[0]
    _lib.check(lib.sfm_fundamental_npt_static(device.ptr(d1), device.ptr(d2), len(cases[0][0]), device.ptr(dF),
                                              device.stream()), "sfm_fundamental_npt_static")
    torch.cuda.synchronize()
    torch_before = torch.cuda.memory_allocated()
    free_before, _ = torch.cuda.mem_get_info()
    got = []
    for (a, _), (d1, d2, dF) in zip(cases, bufs):                    # two calls in a row
        _lib.check(lib.sfm_fundamental_npt_static(device.ptr(d1), device.ptr(d2), len(a), device.ptr(dF),
                                                  device.stream()), "sfm_fundamental_npt_static")
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info()
    assert torch.cuda.memory_allocated() == torch_before
    assert free_before - free_after < (1 << 20), "the allocation-free entry point took device memory"
    for _, _, dF in bufs:
        got.append(dF.cpu().numpy())
    for k in range(2):
        assert np.all(np.isfinite(got[k]))
        assert rel_err(got[k], want[k]) < F_RTOL, (k, rel_err(got[k], want[k]))
    # and against the reference's own output for the golden set (EstimateFundamentalMatrix.py:3-83 on all points)
    assert rel_err(got[1].reshape(3, 3), g["F_all"]) < F_RTOL


def test_static_scratch_variant_rejects_what_does_not_fit(cuda_device):
    import torch
    from sfm_b200 import _lib, device
    lib = _lib.load()
    n = MAX_N + 1
    d1 = device.empty((n, 2), torch.float64)
    dF = device.empty((9,), torch.float64)
    rc = lib.sfm_fundamental_npt_static(device.ptr(d1), device.ptr(d1), n, device.ptr(dF), device.stream())
    assert rc == 3                                                   # SFM_E_WORKSPACE
    msg = lib.sfm_last_cuda_error().decode()
    assert "exceeds the static scratch maximum" in msg and str(MAX_N) in msg
    assert lib.sfm_fundamental_npt_static(None, None, 100, None, None) == 1       # SFM_E_INVALID
    assert lib.sfm_fundamental_npt_static(device.ptr(d1), device.ptr(d1), 7, device.ptr(dF), None) == 1
    # the limit itself is accepted
    from oracle import scenes
    sc = scenes.two_view(n=MAX_N, outlier_frac=0.0, seed=5)
    x1, x2 = sc["x1"], sc["x2"]
    a, b = device.to_device(x1), device.to_device(x2)
    _lib.check(lib.sfm_fundamental_npt_static(device.ptr(a), device.ptr(b), MAX_N, device.ptr(dF), device.stream()),
               "sfm_fundamental_npt_static")
    want = _reference_call(lib, device, torch, a, b, MAX_N)
    assert rel_err(dF.cpu().numpy(), want) < F_RTOL
