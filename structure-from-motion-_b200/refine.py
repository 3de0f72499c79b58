"""Host side of the per-point / per-camera refiners: mirrors NonLinearTriangulation.py,
NonlinearPnP.py and PnPRANSAC.py (names, argument order, return order, error behaviour) and
drives K5 / K6 / K7 through the C ABI. Batched entry points (one call, many problems) sit next
to the reference signatures for the C4-scale workloads."""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch
from scipy.spatial.transform import Rotation

from . import _lib, device, sampler

# the reference's scipy defaults (README.md:203, least_squares ftol=xtol=gtol=1e-8)
REF_TOL = 1e-8
TRI_MAX_NFEV = 50      # NonLinearTriangulation.py:121
PNP_MAX_NFEV = 100     # NonlinearPnP.py:103


def _hptr(a: np.ndarray) -> int:
    return a.ctypes.data


def projection_matrix(K, C_, R):
    """P = K [R | -R C]  (NonLinearTriangulation.py:96-100)."""
    K, C_, R = np.array(K, float), np.array(C_, float), np.array(R, float)
    return np.ascontiguousarray(K @ np.hstack([R, -R @ C_.reshape(3, 1)]))


# ---------------------------------------------------------------------------------------------
# LinearTriangulation.py (it produces x0 for every NonLinearTriangulation call)
# ---------------------------------------------------------------------------------------------
def lintri_device(P1, P2, d_x1, d_x2):
    """N1 on device tensors -> d_X [n,3]."""
    n = int(d_x1.shape[0])
    d_X = device.empty((n, 3), torch.float64)
    P1 = np.ascontiguousarray(P1, np.float64)
    P2 = np.ascontiguousarray(P2, np.float64)
    rc = _lib.load().sfm_lintri(_hptr(P1), _hptr(P2), device.ptr(d_x1), device.ptr(d_x2), n, device.ptr(d_X),
                                device.stream())
    _lib.check(rc, "sfm_lintri")
    return d_X


def LinearTriangulation(K, C1, R1, C2, R2, x1, x2):
    """Drop-in for LinearTriangulation.py:3-92 -> points (N x 3)."""
    x1 = np.array(x1)
    x2 = np.array(x2)
    n_points = len(x1)
    if n_points == 0:
        return np.array([])
    P1, P2 = projection_matrix(K, C1, R1), projection_matrix(K, C2, R2)
    d_X = lintri_device(P1, P2, device.to_device(x1.reshape(n_points, 2)), device.to_device(x2.reshape(n_points, 2)))
    return d_X.cpu().numpy()


def linear_triangulation(K, C1, R1, C2, R2, x1, x2):
    """LinearTriangulation.py:95-99."""
    return LinearTriangulation(K, C1, R1, C2, R2, x1, x2)


# ---------------------------------------------------------------------------------------------
# NonLinearTriangulation.py
# ---------------------------------------------------------------------------------------------
def tri_residuals_device(P1, P2, d_x1, d_x2, d_X):
    """Loss() for n points on device tensors -> d_r [n,4] (sfm_tri_residuals)."""
    n = int(d_X.shape[0])
    d_r = device.empty((n, 4), torch.float64)
    P1, P2 = np.ascontiguousarray(P1, np.float64), np.ascontiguousarray(P2, np.float64)
    _lib.check(_lib.load().sfm_tri_residuals(_hptr(P1), _hptr(P2), device.ptr(d_x1), device.ptr(d_x2), device.ptr(d_X), n,
                                             device.ptr(d_r), device.stream()), "sfm_tri_residuals")
    return d_r


def Loss(X, x1, x2, P1, P2):
    """Drop-in for NonLinearTriangulation.py:5-50 -> residual (4,): one point through sfm_tri_residuals."""
    X, x1, x2 = np.asarray(X, np.float64), np.asarray(x1, np.float64), np.asarray(x2, np.float64)
    d_r = tri_residuals_device(P1, P2, device.to_device(x1.reshape(1, 2)), device.to_device(x2.reshape(1, 2)),
                               device.to_device(X.reshape(1, 3)))
    return d_r.cpu().numpy().reshape(4)


def nltri_device(P1, P2, d_x1, d_x2, d_X0, max_nfev=TRI_MAX_NFEV, tol=REF_TOL, want_info=False):
    """K5 on device tensors; returns d_X [n,3] (and cost/status tensors with want_info)."""
    n = int(d_x1.shape[0])
    d_X = device.empty((n, 3), torch.float64)
    d_cost = device.empty((n,), torch.float64) if want_info else None
    d_st = device.empty((n,), torch.int32) if want_info else None
    P1 = np.ascontiguousarray(P1, np.float64)
    P2 = np.ascontiguousarray(P2, np.float64)
    rc = _lib.load().sfm_nltri(_hptr(P1), _hptr(P2), device.ptr(d_x1), device.ptr(d_x2), device.ptr(d_X0),
                               n, int(max_nfev), float(tol), device.ptr(d_X), device.ptr(d_cost),
                               device.ptr(d_st), device.stream())
    _lib.check(rc, "sfm_nltri")
    return (d_X, d_cost, d_st) if want_info else d_X


def NonLinearTriangulation(K, C1, R1, C2, R2, x1, x2, x0, tol=REF_TOL, info=None):
    """Drop-in for NonLinearTriangulation.py:53-130 -> refined points (N x 3)."""
    x1 = np.array(x1)
    x2 = np.array(x2)
    x0 = np.array(x0)
    n_points = len(x1)
    if n_points == 0:
        return np.array([])
    P1, P2 = projection_matrix(K, C1, R1), projection_matrix(K, C2, R2)
    d_x1 = device.to_device(x1.reshape(n_points, 2))
    d_x2 = device.to_device(x2.reshape(n_points, 2))
    d_X0 = device.to_device(x0.reshape(n_points, 3))
    d_X, d_cost, d_st = nltri_device(P1, P2, d_x1, d_x2, d_X0, TRI_MAX_NFEV, tol, want_info=True)
    X = d_X.cpu().numpy()
    if info is not None:
        st = d_st.cpu().numpy()
        info.update(cost=d_cost.cpu().numpy(), status=st >> 16, nfev=st & 0xFFFF)
    return X


def nonlinear_triangulation(K, C1, R1, C2, R2, x1, x2, x0):
    """NonLinearTriangulation.py:134-138."""
    return NonLinearTriangulation(K, C1, R1, C2, R2, x1, x2, x0)


# ---------------------------------------------------------------------------------------------
# NonlinearPnP.py
# ---------------------------------------------------------------------------------------------
def pnp_residuals_device(K, rt, d_X, d_x):
    """NonLinearPnPLoss() on device tensors -> d_r [2n] interleaved (sfm_pnp_residuals)."""
    n = int(d_X.shape[0])
    d_r = device.empty((2 * n,), torch.float64)
    K, rt = np.ascontiguousarray(K, np.float64), np.ascontiguousarray(rt, np.float64)
    _lib.check(_lib.load().sfm_pnp_residuals(_hptr(K), _hptr(rt), device.ptr(d_X), device.ptr(d_x), n, device.ptr(d_r),
                                             device.stream()), "sfm_pnp_residuals")
    return d_r


def NonLinearPnPLoss(X0, X, x, K):
    """Drop-in for NonlinearPnP.py:5-44 -> residual (2N,), interleaved [du, dv]."""
    X, x = np.asarray(X, np.float64), np.asarray(x, np.float64)
    if len(X) == 0:
        return np.zeros(0)
    return pnp_residuals_device(K, np.asarray(X0, np.float64)[:6], device.to_device(X), device.to_device(x)).cpu().numpy()


def nlpnp_batch_device(K, d_X, d_x, d_offsets, d_rt0, max_nfev=PNP_MAX_NFEV, tol=REF_TOL):
    """K7 on device tensors: B independent registrations. Returns (d_rt [B,6], d_cost, d_status)."""
    B = int(d_rt0.shape[0])
    d_rt = device.empty((B, 6), torch.float64)
    d_cost = device.empty((B,), torch.float64)
    d_st = device.empty((B,), torch.int32)
    Kc = np.ascontiguousarray(K, np.float64)
    rc = _lib.load().sfm_nlpnp_batch(_hptr(Kc), device.ptr(d_X), device.ptr(d_x), device.ptr(d_offsets), B,
                                     device.ptr(d_rt0), int(max_nfev), float(tol), device.ptr(d_rt),
                                     device.ptr(d_cost), device.ptr(d_st), device.stream())
    _lib.check(rc, "sfm_nlpnp_batch")
    return d_rt, d_cost, d_st


def nonlinear_pnp_batch(X_list, x_list, K, C_list, R_list, tol=REF_TOL):
    """Many NonLinearPnP problems in one launch. Returns (C_list, R_list, info dict)."""
    sizes = [len(X) for X in X_list]
    offsets = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    rt0 = np.array([np.hstack([Rotation.from_matrix(np.array(R, float)).as_rotvec(),
                               -np.array(R, float) @ np.array(C_, float)])
                    for C_, R in zip(C_list, R_list)])
    d_X = device.to_device(np.concatenate([np.asarray(X, float).reshape(-1, 3) for X in X_list]))
    d_x = device.to_device(np.concatenate([np.asarray(x, float).reshape(-1, 2) for x in x_list]))
    d_rt, d_cost, d_st = nlpnp_batch_device(K, d_X, d_x, device.to_device(offsets, np.int32),
                                            device.to_device(rt0), PNP_MAX_NFEV, tol)
    rt = d_rt.cpu().numpy()
    st = d_st.cpu().numpy()
    Rn = Rotation.from_rotvec(rt[:, :3]).as_matrix()
    Cn = -np.einsum("nji,nj->ni", Rn, rt[:, 3:6])
    return list(Cn), list(Rn), dict(cost=d_cost.cpu().numpy(), status=st >> 16, nfev=st & 0xFFFF, rt=rt)


def NonLinearPnP(X, x, K, C_, R, tol=REF_TOL, info=None):
    """Drop-in for NonlinearPnP.py:47-123 -> (Cnew, Rnew)."""
    X = np.array(X)
    x = np.array(x)
    K = np.array(K)
    C_ = np.array(C_)
    R = np.array(R)
    if len(X) < 4:                                        # :82-84
        return C_, R
    rvec = Rotation.from_matrix(R).as_rotvec()            # :87, outside the try like the reference
    rt0 = np.hstack([rvec, -R @ C_]).reshape(1, 6)
    try:
        d_rt, d_cost, d_st = nlpnp_batch_device(
            K, device.to_device(X.reshape(-1, 3)), device.to_device(x.reshape(-1, 2)),
            device.to_device(np.array([0, len(X)], np.int32), np.int32), device.to_device(rt0),
            PNP_MAX_NFEV, tol)
        rt = d_rt.cpu().numpy()[0]
        st = int(d_st.cpu().numpy()[0])
        if (st >> 16) < 0:
            raise ValueError("Residuals are not finite in the initial point.")
        Rnew = Rotation.from_rotvec(rt[:3]).as_matrix()   # :113
        Cnew = -Rnew.T @ rt[3:6]                          # :116
        if info is not None:
            info.update(cost=float(d_cost.cpu().numpy()[0]), status=st >> 16, nfev=st & 0xFFFF, rt=rt)
        return Cnew, Rnew
    except _lib.SfmError:
        raise
    except Exception as e:                                # :120-123
        print(f"Non-linear PnP optimization failed: {e}")
        return C_, R


def nonlinear_PnP(K, C_, R, x, X):
    """NonlinearPnP.py:127-151 (argument order of the wrapper)."""
    return NonLinearPnP(X, x, K, C_, R)
