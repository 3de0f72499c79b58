"""Host side of bundle adjustment: mirrors BundleAdjustment.py (project_points :8-40,
bundle_adjustment_residuals :43-110, perform_bundle_adjustment :113-242; names, argument order,
return order, prints and the 'return the inputs' failure convention) and drives K8-K12 through
the C ABI (sfm_ba_solve / sfm_ba_residuals). The sparse entry points (`ba_solve_sparse`,
`BAProblem`) sit next to the reference signature for the C3/C5-scale workloads, where the
reference's dense (n_features x n_images) arrays do not scale."""
from __future__ import annotations

import collections
import ctypes as C
import zlib

import numpy as np
import torch
from scipy.spatial.transform import Rotation

from . import _lib, device

BA_MAX_NFEV = 100          # BundleAdjustment.py:211
REF_TOL = 1e-8             # scipy least_squares defaults (ftol = xtol = gtol)

STATUS_TEXT = {-1: "non-finite residuals", 0: "max evaluations", 1: "gtol", 2: "ftol", 3: "xtol",
               4: "ftol and xtol"}


def _hptr(a: np.ndarray) -> int:
    return a.ctypes.data


# ---------------------------------------------------------------------------------------------
# sparse device-level API
# ---------------------------------------------------------------------------------------------
class BAProblem:
    """Device-resident sparse BA problem: observations sorted by point (the reference's order,
    BundleAdjustment.py:164-169). Holds the workspace, so repeated solves do not allocate."""

    def __init__(self, n_cams: int, n_pts: int, cam_idx, pt_idx, obs, K):
        cam_idx = np.ascontiguousarray(cam_idx, np.int32)
        pt_idx = np.ascontiguousarray(pt_idx, np.int32)
        obs = np.ascontiguousarray(obs, np.float64).reshape(-1, 2)
        if not (len(cam_idx) == len(pt_idx) == len(obs)):
            raise ValueError("cam_idx, pt_idx and obs must have one entry per observation")
        self._order = None
        if len(pt_idx) > 1 and np.any(np.diff(pt_idx) < 0):
            order = np.argsort(pt_idx, kind="stable")       # point-major, original order within a point
            cam_idx, pt_idx, obs = cam_idx[order], pt_idx[order], obs[order]
            self._order = order
        self.n_cams, self.n_pts, self.n_obs = int(n_cams), int(n_pts), int(len(obs))
        self.K = np.ascontiguousarray(K, np.float64).reshape(3, 3)
        self.d_cam_idx = device.to_device(cam_idx, np.int32)
        self.d_pt_idx = device.to_device(pt_idx, np.int32)
        self.d_obs = device.to_device(obs)
        lib = _lib.load()
        self.ws_bytes = int(lib.sfm_ba_workspace_bytes(self.n_cams, self.n_pts, self.n_obs))
        if self.ws_bytes == 0:
            raise ValueError("empty bundle adjustment problem")
        self.ws = device.empty((self.ws_bytes,), torch.uint8)
        self._indexed = False          # the workspace holds this problem's camera-major index

    @classmethod
    def from_device(cls, n_cams, n_pts, d_cam_idx, d_pt_idx, d_obs, K):
        """The same from device tensors that are already point-sorted (sfm_ba_assemble's output): nothing goes
        through the host."""
        self = cls.__new__(cls)
        self._order = None
        self.n_cams, self.n_pts, self.n_obs = int(n_cams), int(n_pts), int(d_obs.shape[0])
        self.K = np.ascontiguousarray(K, np.float64).reshape(3, 3)
        self.d_cam_idx, self.d_pt_idx, self.d_obs = d_cam_idx.contiguous(), d_pt_idx.contiguous(), d_obs.contiguous()
        self.ws_bytes = int(_lib.load().sfm_ba_workspace_bytes(self.n_cams, self.n_pts, self.n_obs))
        if self.ws_bytes == 0:
            raise ValueError("empty bundle adjustment problem")
        self.ws = device.empty((self.ws_bytes,), torch.uint8)
        self._indexed = False
        return self

    def close(self):
        """Release what the library caches for this workspace (pair list, K11s plan) -- also run when the object is
        collected, so a long incremental run does not accumulate the plans of dead problems."""
        ws, self.ws = getattr(self, "ws", None), None
        if ws is not None and getattr(self, "_indexed", False):
            try:
                _lib.load().sfm_ba_release_workspace(device.ptr(ws), device.stream())
            except Exception:                       # interpreter shutdown: the process is going away anyway
                pass
        self._indexed = False

    def __del__(self):
        self.close()

    def set_observations(self, obs, K=None):
        """New measurements for the SAME index structure (cam_idx / pt_idx unchanged): the plan stays valid."""
        obs = np.ascontiguousarray(obs, np.float64).reshape(-1, 2)
        if len(obs) != self.n_obs:
            raise ValueError("set_observations: the number of observations changed")
        if self._order is not None:
            obs = obs[self._order]
        # through a pinned staging buffer owned by the problem: a pageable source makes the copy synchronous and
        # half as fast (3.2 MB at C3)
        if getattr(self, "_pin_obs", None) is None:
            self._pin_obs = torch.empty((self.n_obs, 2), dtype=torch.float64, pin_memory=True)
        device.upload_via_pinned(self.d_obs, obs, self._pin_obs)
        if K is not None:
            self.K = np.ascontiguousarray(K, np.float64).reshape(3, 3)

    def residuals_device(self, d_cam, d_pts, want_r=True):
        """K8: (d_r [n_obs,2] or None, d_cost [1]) at the given parameters."""
        d_r = device.empty((self.n_obs, 2), torch.float64) if want_r else None
        d_cost = device.empty((1,), torch.float64)
        rc = _lib.load().sfm_ba_residuals(
            device.ptr(d_cam), device.ptr(d_pts), device.ptr(self.d_obs), device.ptr(self.d_cam_idx),
            device.ptr(self.d_pt_idx), self.n_cams, self.n_pts, self.n_obs, _hptr(self.K),
            device.ptr(d_r), device.ptr(d_cost), device.ptr(self.ws), self.ws_bytes, device.stream())
        _lib.check(rc, "sfm_ba_residuals")
        return d_r, d_cost

    def linearise_device(self, d_cam, d_pts, which=3, reuse_table=False):
        """One linearisation pass in isolation (bench / profiling): which = 1 point pass K9p, 2 camera
        pass K9c, 4 cost pass K8 (bit mask)."""
        rc = _lib.load().sfm_ba_linearise(
            device.ptr(d_cam), device.ptr(d_pts), device.ptr(self.d_obs), device.ptr(self.d_cam_idx),
            device.ptr(self.d_pt_idx), self.n_cams, self.n_pts, self.n_obs, _hptr(self.K),
            int(which) | (8 if reuse_table else 0), int(self._indexed), device.ptr(self.ws), self.ws_bytes,
            device.stream())
        _lib.check(rc, "sfm_ba_linearise")
        self._indexed = True

    def solve_device(self, d_cam, d_pts, max_nfev=BA_MAX_NFEV, tol=REF_TOL, lambda0=1e-4, verbose=False, profile=False):
        """K8-K12 LM; d_cam [n_cams,6] and d_pts [n_pts,3] are refined in place. Returns a dict.
        profile: record the phase times of this solve (read them with last_profile())."""
        lib = _lib.load()
        opts = _lib.BAOptions()
        lib.sfm_ba_default_options(C.addressof(opts))
        opts.max_trials, opts.verbose = int(max_nfev), int(bool(verbose))
        opts.ftol = opts.xtol = opts.gtol = float(tol)
        opts.lambda0 = float(lambda0)
        opts.reuse_index = int(self._indexed)
        opts.reserved = 1 if profile else 0
        res = _lib.BAResult()
        rc = lib.sfm_ba_solve(
            device.ptr(d_cam), device.ptr(d_pts), device.ptr(self.d_obs), device.ptr(self.d_cam_idx),
            device.ptr(self.d_pt_idx), self.n_cams, self.n_pts, self.n_obs, _hptr(self.K),
            C.addressof(opts), C.addressof(res), device.ptr(self.ws), self.ws_bytes, device.stream())
        _lib.check(rc, "sfm_ba_solve")
        self._indexed = True
        return dict(cost0=res.cost0, cost=res.cost, gnorm=res.gnorm, lambda_=res.lambda_, status=res.status,
                    n_linearisations=res.n_linearisations, n_trials=res.n_trials, n_chol_fail=res.n_chol_fail)


def last_profile() -> dict:
    """Phase times [ms] of the last solve that ran with profile=True (sfm_ba_last_profile)."""
    out = np.zeros(8)
    _lib.check(_lib.load().sfm_ba_last_profile(out.ctypes.data), "sfm_ba_last_profile")
    return dict(schur_ms=out[0], s_allreduce_ms=out[1], camera_solve_k11_ms=out[2], trial_rest_ms=out[3],
                linearise_ms=out[4], trial_steps=int(out[5]), linearisations=int(out[6]),
                s_allreduce_bytes_per_trial=int(out[7]))


def release_cache() -> int:
    """Free the plan-time pair lists the library caches per workspace (sfm_ba_release_cache); returns the bytes."""
    import ctypes as C
    n = C.c_uint64(0)
    _lib.check(_lib.load().sfm_ba_release_cache(C.addressof(n), device.stream()), "sfm_ba_release_cache")
    return int(n.value)


# Problems recently solved through the host-array API, keyed by their index structure: everything the library
# prepares per problem (workspace, camera-major index, K10g pair list, K11s elimination plan) depends only on
# (cam_idx, pt_idx), so a caller that re-adjusts the same tracks -- every LM restart, every bench step -- pays for it
# once. Bounded by count and by device bytes; clear_problem_cache() empties it.
_PROBLEMS: "collections.OrderedDict" = collections.OrderedDict()
PROBLEM_CACHE_ENTRIES = 4
PROBLEM_CACHE_BYTES = 8 << 30


def clear_problem_cache():
    while _PROBLEMS:
        _PROBLEMS.popitem(last=False)[1].close()


def _cached_problem(n_cams, n_pts, cam_idx, pt_idx, obs, K):
    cam_in, pt_in = cam_idx, pt_idx
    cam_idx = np.ascontiguousarray(cam_idx, np.int32)
    pt_idx = np.ascontiguousarray(pt_idx, np.int32)
    dev = torch.cuda.current_device()
    # fast path: the very array objects of the most recent call (an LM restart, a bench step) -- found by identity,
    # still verified by content (the caller may have written into them), without the two checksum passes
    key = None
    if _PROBLEMS:
        k_last, p_last = next(reversed(_PROBLEMS.items()))
        if p_last._key_ids == (id(cam_in), id(pt_in)) and k_last[:3] == (int(n_cams), int(n_pts), len(cam_idx)) and \
                k_last[-1] == dev:
            key = k_last
    if key is None:
        key = (int(n_cams), int(n_pts), len(cam_idx), zlib.crc32(cam_idx), zlib.crc32(pt_idx), dev)
    prob = _PROBLEMS.pop(key, None)
    if prob is not None and not (device.arrays_equal(prob._key_cam, cam_idx) and device.arrays_equal(prob._key_pt, pt_idx)):
        prob.close()                                  # changed in place, or a checksum collision: not the same tracks
        prob = None
        key = (int(n_cams), int(n_pts), len(cam_idx), zlib.crc32(cam_idx), zlib.crc32(pt_idx), dev)
    if prob is None:
        prob = BAProblem(n_cams, n_pts, cam_idx, pt_idx, obs, K)
        prob._key_cam, prob._key_pt = cam_idx.copy(), pt_idx.copy()
    else:
        prob.set_observations(obs, K)
    prob._key_ids = (id(cam_in), id(pt_in))
    _PROBLEMS[key] = prob
    while len(_PROBLEMS) > PROBLEM_CACHE_ENTRIES or (len(_PROBLEMS) > 1 and
                                                      sum(p.ws_bytes for p in _PROBLEMS.values()) > PROBLEM_CACHE_BYTES):
        _PROBLEMS.popitem(last=False)[1].close()
    return prob


def ba_solve_sparse(cams, pts, cam_idx, pt_idx, obs, K, max_nfev=BA_MAX_NFEV, tol=REF_TOL, verbose=False):
    """Host arrays in/out: cams [Nc,6] = [rotvec, t], pts [Np,3] -> (cams, pts, info)."""
    cams = np.ascontiguousarray(cams, np.float64).reshape(-1, 6)
    pts = np.ascontiguousarray(pts, np.float64).reshape(-1, 3)
    prob = _cached_problem(len(cams), len(pts), cam_idx, pt_idx, obs, K)
    # parameters travel as ONE pinned buffer [cameras | points] each way: one asynchronous upload, one download and
    # one synchronisation per call (two pageable copies each way cost ~0.1 ms of a 1.5 ms call at C3)
    nc6, np3 = cams.size, pts.size
    if getattr(prob, "_pin_par", None) is None or prob._pin_par.numel() != nc6 + np3:
        prob._pin_par = torch.empty(nc6 + np3, dtype=torch.float64, pin_memory=True)
        prob._d_par = device.empty((nc6 + np3,), torch.float64)
    hp = prob._pin_par.numpy()
    d_par = prob._d_par
    np.copyto(hp[:nc6], cams.reshape(-1))
    if pts.nbytes >= device.PAR_MIN_BYTES:
        device.upload_via_pinned(d_par[nc6:], pts.reshape(-1), prob._pin_par[nc6:])
        d_par[:nc6].copy_(prob._pin_par[:nc6], non_blocking=True)
    else:
        np.copyto(hp[nc6:], pts.reshape(-1))
        d_par.copy_(prob._pin_par, non_blocking=True)
    d_cam, d_pts = d_par[:nc6].view(-1, 6), d_par[nc6:].view(-1, 3)
    info = prob.solve_device(d_cam, d_pts, max_nfev, tol, verbose=verbose)        # ends with a stream synchronisation
    prob._pin_par.copy_(d_par, non_blocking=True)
    torch.cuda.current_stream().synchronize()
    return hp[:nc6].reshape(-1, 6).copy(), hp[nc6:].reshape(-1, 3).copy(), info


# ---------------------------------------------------------------------------------------------
# BundleAdjustment.py names
# ---------------------------------------------------------------------------------------------
def project_points(K, C_, R, X):
    """BundleAdjustment.py:8-40 (exported helper; a handful of points, host arithmetic)."""
    K, C_, R, X = np.array(K, float), np.array(C_, float), np.array(R, float), np.array(X, float)
    P = K @ np.hstack([R, -R @ C_.reshape(3, 1)])
    xh = (P @ np.hstack([X, np.ones((X.shape[0], 1))]).T).T
    return xh[:, :2] / (xh[:, 2:3] + 1e-8)


def bundle_adjustment_residuals(params, n_cameras, n_points, camera_indices, point_indices, points_2d, K,
                                n_cam_params=6):
    """Drop-in for BundleAdjustment.py:43-110 -> residual vector (2*n_obs,), interleaved [du,dv],
    observed - projected, in the caller's observation order. Evaluated by K8 on the GPU."""
    if n_cam_params != 6:
        raise ValueError("only the reference's 6-parameter camera block [rotvec, t] is supported")
    params = np.asarray(params, np.float64)
    cams = params[:n_cameras * 6].reshape(n_cameras, 6)
    pts = params[n_cameras * 6:n_cameras * 6 + n_points * 3].reshape(n_points, 3)
    cam_idx = np.ascontiguousarray(camera_indices, np.int32)
    pt_idx = np.ascontiguousarray(point_indices, np.int32)
    if len(cam_idx) == 0:
        return np.array([])
    # K8 has no ordering requirement, but BAProblem sorts by point: undo that for the caller
    order = None
    if len(pt_idx) > 1 and np.any(np.diff(pt_idx) < 0):
        order = np.argsort(pt_idx, kind="stable")
    prob = BAProblem(n_cameras, n_points, cam_idx, pt_idx, points_2d, K)
    d_r, _ = prob.residuals_device(device.to_device(cams), device.to_device(pts))
    r = d_r.cpu().numpy()
    if order is not None:
        out = np.empty_like(r)
        out[order] = r
        r = out
    return r.reshape(-1)


def build_problem(all_world_coords, filtered_world_coords, feature_x, feature_y, filtered_feature_flags,
                  R_set, C_set):
    """Observation list and initial parameters of BundleAdjustment.py:148-199, vectorised (the
    reference's nested loops are O(Nobs * Np)). Returns None when there is nothing to optimise."""
    filtered_world_coords = np.asarray(filtered_world_coords)
    valid = np.where(filtered_world_coords.flatten() == 1)[0]                  # :149-150
    if len(valid) == 0:
        return None
    n_cameras = len(R_set)                                                     # :156 (cam_index unused)
    flags = np.asarray(filtered_feature_flags)[valid][:, :n_cameras] == 1
    pt_local, cam = np.nonzero(flags)                                          # point-major, camera ascending
    if len(cam) == 0:
        return None
    fx, fy = np.asarray(feature_x), np.asarray(feature_y)
    rows = valid[pt_local]
    obs = np.stack([fx[rows, cam], fy[rows, cam]], axis=1).astype(np.float64)
    cams = np.empty((n_cameras, 6))
    for i in range(n_cameras):
        R = np.array(R_set[i])
        Cc = np.array(C_set[i])
        cams[i, :3] = Rotation.from_matrix(R).as_rotvec()                      # :188 (raises for det <= 0)
        cams[i, 3:] = -R @ Cc                                                  # :191
    pts = np.asarray(all_world_coords, np.float64)[valid].reshape(-1, 3).copy()
    return dict(valid=valid, n_cameras=n_cameras, n_points=len(valid), cam_idx=cam.astype(np.int32),
                pt_idx=pt_local.astype(np.int32), obs=obs, cams=cams, pts=pts)


def build_problem_device(all_world_coords, filtered_world_coords, feature_x, feature_y, filtered_feature_flags,
                         n_cameras):
    """BundleAdjustment.py:148-176 on the device: the dense (n_features x n_images) arrays
    go up once, the point-sorted observation list is built there (sfm_ba_assemble) and stays there.
    Returns None when there is nothing to optimise, else dict of device tensors cam_idx, pt_idx, obs, pts,
    valid_rows (trimmed views) and the host counts n_points, n_obs."""
    lib = _lib.load()
    valid = np.ascontiguousarray(np.asarray(filtered_world_coords).reshape(-1) == 1, np.uint8)
    flags = np.asarray(filtered_feature_flags)
    F, stride = int(flags.shape[0]), int(flags.shape[1])
    if F == 0 or len(valid) != F:
        return None
    n_cameras = int(min(n_cameras, stride))
    d_valid = device.to_device(valid, np.uint8)
    d_flags = device.to_device(np.ascontiguousarray(flags == 1, np.uint8), np.uint8)
    d_fx, d_fy = device.to_device(np.asarray(feature_x, np.float64)), device.to_device(np.asarray(feature_y, np.float64))
    d_world = device.to_device(np.asarray(all_world_coords, np.float64).reshape(F, 3))
    cap = max(1, F * max(n_cameras, 1))
    d_cam, d_pt = device.empty((cap,), torch.int32), device.empty((cap,), torch.int32)
    d_obs, d_pts = device.empty((cap, 2), torch.float64), device.empty((F, 3), torch.float64)
    d_rows, d_tot = device.empty((F,), torch.int32), device.empty((2,), torch.int32)
    nbytes = int(lib.sfm_ba_assemble_workspace_bytes(F))
    ws = device.empty((nbytes,), torch.uint8)
    rc = lib.sfm_ba_assemble(device.ptr(d_valid), device.ptr(d_flags), device.ptr(d_fx), device.ptr(d_fy),
                             device.ptr(d_world), F, stride, n_cameras, device.ptr(d_cam), device.ptr(d_pt),
                             device.ptr(d_obs), device.ptr(d_pts), device.ptr(d_rows), device.ptr(d_tot), device.ptr(ws),
                             nbytes, device.stream())
    _lib.check(rc, "sfm_ba_assemble")
    n_points, n_obs = (int(v) for v in d_tot.cpu().numpy())
    if n_points == 0 or n_obs == 0:
        return None
    return dict(n_cameras=n_cameras, n_points=n_points, n_obs=n_obs, cam_idx=d_cam[:n_obs], pt_idx=d_pt[:n_obs],
                obs=d_obs[:n_obs], pts=d_pts[:n_points], valid_rows=d_rows[:n_points])


def perform_bundle_adjustment(all_world_coords, filtered_world_coords, feature_x, feature_y,
                              filtered_feature_flags, R_set, C_set, K, cam_index, tol=REF_TOL, info=None):
    """Drop-in for BundleAdjustment.py:113-242 -> (R_set_opt, C_set_opt, all_world_coords_opt).
    The observation list of :148-176 is assembled on the device (sfm_ba_assemble): the dense
    (n_features x n_images) arrays go up once and the point-sorted list never visits the host."""
    n_cameras = len(R_set)                                                     # :156 (cam_index unused)
    asm = build_problem_device(all_world_coords, filtered_world_coords, feature_x, feature_y,
                               filtered_feature_flags, n_cameras) if n_cameras > 0 else None
    if asm is None:                                                            # :152-153, :171-172
        return R_set, C_set, all_world_coords
    cams = np.empty((n_cameras, 6))
    for i in range(n_cameras):                                                 # :183-191, outside the try like the reference
        R = np.array(R_set[i])
        cams[i, :3] = Rotation.from_matrix(R).as_rotvec()                      # raises for det <= 0 (scipy), as there
        cams[i, 3:] = -R @ np.array(C_set[i])
    n_points, n_obs = asm["n_points"], asm["n_obs"]
    print(f"  Bundle adjustment: {n_cameras} cameras, {n_points} points, {n_obs} observations")
    try:
        m, n = 2 * n_obs, 6 * n_cameras + 3 * n_points
        if m < n:                                                              # scipy's check for 'lm'
            raise ValueError("Method 'lm' doesn't work when the number of residuals is less than the "
                             "number of variables.")
        prob = BAProblem.from_device(n_cameras, n_points, asm["cam_idx"], asm["pt_idx"], asm["obs"], np.array(K, float))
        d_cam, d_pts = device.to_device(cams), asm["pts"].contiguous()
        res = prob.solve_device(d_cam, d_pts, BA_MAX_NFEV, tol)
        if res["status"] < 0:
            raise ValueError("Residuals are not finite in the initial point.")
        cams_opt, pts = d_cam.cpu().numpy(), d_pts.cpu().numpy()
        rows = asm["valid_rows"].cpu().numpy()
        prob.close()
        Rm = Rotation.from_rotvec(cams_opt[:, :3]).as_matrix()
        R_set_opt = [Rm[i] for i in range(n_cameras)]                          # :220-228
        C_set_opt = [-Rm[i].T @ cams_opt[i, 3:] for i in range(n_cameras)]
        all_world_coords_opt = np.array(all_world_coords, copy=True)           # :231-234
        all_world_coords_opt[rows] = pts
        print(f"  Bundle adjustment completed. Final cost: {res['cost']:.6f}")
        if info is not None:
            info.update(res, cams=cams_opt, pts=pts)
        return R_set_opt, C_set_opt, all_world_coords_opt
    except _lib.SfmError:
        raise                                  # a broken GPU path must fail loudly, not look like a BA failure
    except Exception as e:                                                     # :240-242
        print(f"  Bundle adjustment failed: {e}")
        return R_set, C_set, all_world_coords
