"""ctypes binding of libsfm_b200.so (the C ABI declared in include/sfm_b200.h).

There is deliberately no fallback: if the shared library is missing and cannot be built,
or an entry point fails, the caller gets an exception.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

from . import build as _build

_lock = threading.Lock()
_lib = None

c_f64p = C.c_void_p   # device or host pointers are passed as integers
c_i32p = C.c_void_p
c_u8p = C.c_void_p
c_vp = C.c_void_p

# name -> (restype, [argtypes]); must list EVERY symbol include/sfm_b200.h declares
# (tests/test_abi.py parses the header and checks both directions).
SIGNATURES = {
    "sfm_abi_version": (C.c_int, []),
    "sfm_last_cuda_error": (C.c_char_p, []),
    "sfm_launch_count": (C.c_int64, []),
    "sfm_mt_sample_stream": (C.c_int, [c_vp, c_vp, C.c_int32, C.c_int32, C.c_int64, c_vp]),
    "sfm_f8pt_batch": (C.c_int, [c_f64p, C.c_int32, c_i32p, C.c_int32, c_f64p, c_vp]),
    "sfm_f_score_count": (C.c_int, [c_f64p, C.c_int32, c_f64p, C.c_int32, C.c_double, c_i32p, c_vp]),
    "sfm_f_score_count_fp64": (C.c_int, [c_f64p, C.c_int32, c_f64p, C.c_int32, C.c_double, c_i32p, c_vp]),
    "sfm_f_score_fp64_tests": (C.c_int, [C.POINTER(C.c_uint64), C.c_int32]),
    "sfm_argmax_first": (C.c_int, [c_i32p, C.c_int32, C.c_int32, c_i32p, c_vp, c_vp]),
    "sfm_argmax_first_batch": (C.c_int, [c_i32p, C.c_int32, C.c_int32, c_i32p, c_vp]),
    "sfm_pnp_score_count_batch": (C.c_int, [c_f64p, c_f64p, c_f64p, c_i32p, C.c_int32, C.c_int32, C.c_int32, C.c_double,
                                            c_i32p, c_vp]),
    "sfm_f_inlier_mask": (C.c_int, [c_f64p, C.c_int32, c_f64p, C.c_double, c_u8p, c_f64p, c_vp]),
    "sfm_f_ransac": (C.c_int, [c_f64p, C.c_int32, c_i32p, C.c_int32, C.c_double, c_f64p, c_i32p,
                               c_i32p, c_f64p, c_u8p, c_f64p, c_vp]),
    "sfm_f_winner_from_key": (C.c_int, [c_f64p, C.c_int32, c_i32p, C.c_int32, c_vp, C.c_double, c_i32p,
                                        c_f64p, c_u8p, c_f64p, c_vp]),
    "sfm_nltri": (C.c_int, [c_vp, c_vp, c_f64p, c_f64p, c_f64p, C.c_int32, C.c_int32, C.c_double,
                            c_f64p, c_f64p, c_i32p, c_vp]),
    "sfm_lintri": (C.c_int, [c_vp, c_vp, c_f64p, c_f64p, C.c_int32, c_f64p, c_vp]),
    "sfm_h4pt_batch": (C.c_int, [c_f64p, C.c_int32, c_i32p, C.c_int32, c_f64p, c_vp]),
    "sfm_h_score_count": (C.c_int, [c_f64p, C.c_int32, c_f64p, C.c_int32, C.c_double, c_i32p, c_vp]),
    "sfm_h_inlier_mask": (C.c_int, [c_f64p, C.c_int32, c_f64p, C.c_double, c_vp, c_f64p, c_vp]),
    "sfm_h_ransac": (C.c_int, [c_f64p, C.c_int32, c_i32p, C.c_int32, C.c_double, c_f64p, c_i32p, c_i32p,
                               c_f64p, c_vp, c_f64p, c_vp]),
    "sfm_homography_npt_workspace_bytes": (C.c_size_t, [C.c_int32]),
    "sfm_homography_npt": (C.c_int, [c_f64p, c_f64p, C.c_int32, c_f64p, c_vp, C.c_size_t, c_vp]),
    "sfm_ba_assemble_workspace_bytes": (C.c_size_t, [C.c_int32]),
    "sfm_ba_assemble": (C.c_int, [c_vp, c_vp, c_f64p, c_f64p, c_f64p, C.c_int32, C.c_int32, C.c_int32, c_i32p, c_i32p,
                                  c_f64p, c_f64p, c_i32p, c_i32p, c_vp, C.c_size_t, c_vp]),
    "sfm_two_view_configs": (C.c_int, [c_f64p, c_vp, c_f64p, c_f64p, c_i32p, c_vp]),
    "sfm_cheirality_count": (C.c_int, [c_f64p, c_f64p, C.c_int32, c_i32p, c_vp]),
    "sfm_two_view_pose": (C.c_int, [c_f64p, c_vp, c_f64p, c_f64p, C.c_int32, c_f64p, c_f64p, c_f64p,
                                    c_i32p, c_i32p, c_f64p, c_vp]),
    "sfm_pnp_score_count": (C.c_int, [c_f64p, c_f64p, c_f64p, C.c_int32, C.c_int32, C.c_double,
                                      c_i32p, c_vp]),
    "sfm_nlpnp_batch": (C.c_int, [c_vp, c_f64p, c_f64p, c_i32p, C.c_int32, c_f64p, C.c_int32,
                                  C.c_double, c_f64p, c_f64p, c_i32p, c_vp]),
    "sfm_fundamental_npt_workspace_bytes": (C.c_size_t, [C.c_int32]),
    "sfm_fundamental_npt": (C.c_int, [c_f64p, c_f64p, C.c_int32, c_f64p, c_vp, C.c_size_t, c_vp]),
    "sfm_fundamental_npt_static": (C.c_int, [c_f64p, c_f64p, C.c_int32, c_f64p, c_vp]),
    "sfm_linear_pnp4_batch": (C.c_int, [c_f64p, c_f64p, c_i32p, C.c_int32, C.c_int32, c_i32p, C.c_int32, c_vp, c_vp,
                                        C.c_int32, c_f64p, c_f64p, c_f64p, c_u8p, c_vp]),
    "sfm_linear_pnp_npt_workspace_bytes": (C.c_size_t, [C.c_int32]),
    "sfm_linear_pnp_npt": (C.c_int, [c_f64p, c_f64p, C.c_int32, c_vp, c_vp, c_f64p, c_f64p, c_u8p, c_vp, C.c_size_t,
                                     c_vp]),
    "sfm_tri_residuals": (C.c_int, [c_vp, c_vp, c_f64p, c_f64p, c_f64p, C.c_int32, c_f64p, c_vp]),
    "sfm_pnp_residuals": (C.c_int, [c_vp, c_vp, c_f64p, c_f64p, C.c_int32, c_f64p, c_vp]),
    "sfm_ba_default_options": (None, [c_vp]),
    "sfm_ba_workspace_bytes": (C.c_size_t, [C.c_int32, C.c_int32, C.c_int32]),
    "sfm_ba_last_profile": (C.c_int, [c_vp]),
    "sfm_ba_release_workspace": (C.c_int, [c_vp, c_vp]),
    "sfm_ba_release_cache": (C.c_int, [C.c_void_p, C.c_void_p]),
    "sfm_ba_residuals": (C.c_int, [c_f64p, c_f64p, c_f64p, c_i32p, c_i32p, C.c_int32, C.c_int32, C.c_int32,
                                   c_vp, c_f64p, c_f64p, c_vp, C.c_size_t, c_vp]),
    "sfm_ba_solve": (C.c_int, [c_f64p, c_f64p, c_f64p, c_i32p, c_i32p, C.c_int32, C.c_int32, C.c_int32,
                               c_vp, c_vp, c_vp, c_vp, C.c_size_t, c_vp]),
    "sfm_ba_solve_mgpu": (C.c_int, [c_f64p, c_f64p, c_f64p, c_i32p, c_i32p, C.c_int32, C.c_int32, C.c_int32,
                                    c_vp, c_vp, c_vp, c_vp, C.c_size_t, c_vp, C.c_int32, C.c_int32, c_vp]),
    "sfm_nccl_last_error": (C.c_char_p, []),
    "sfm_nccl_unique_id": (C.c_int, [c_vp]),
    "sfm_nccl_comm_create": (C.c_int, [c_vp, C.c_int32, C.c_int32, c_vp]),
    "sfm_nccl_comm_destroy": (C.c_int, [c_vp]),
    "sfm_allreduce_sum_f64": (C.c_int, [c_vp, c_f64p, C.c_int64, c_vp]),
    "sfm_allreduce_max_u64": (C.c_int, [c_vp, c_vp, C.c_int64, c_vp]),
    "sfm_ba_linearise": (C.c_int, [c_f64p, c_f64p, c_f64p, c_i32p, c_i32p, C.c_int32, C.c_int32, C.c_int32,
                                   c_vp, C.c_int32, C.c_int32, c_vp, C.c_size_t, c_vp]),
    "sfm_chol_solve_workspace_bytes": (C.c_size_t, [C.c_int32]),
    "sfm_chol_solve": (C.c_int, [c_f64p, c_f64p, C.c_int32, c_f64p, c_i32p, c_vp, C.c_size_t, c_vp]),
    "sfm_chol_solve_structured": (C.c_int, [c_f64p, c_f64p, C.c_int32, C.c_int32, C.c_int32, c_f64p, c_i32p, c_vp,
                                            C.c_size_t, c_vp, c_vp]),
    "sfm_chol_plan_export": (C.c_int, [c_vp, C.c_int32, C.c_int32, C.c_int32, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp,
                                       c_vp, c_vp, c_vp, c_vp, c_vp]),
    "sfm_chol_diag_bench": (C.c_int, [c_f64p, C.c_int32, c_vp, c_vp]),
    "sfm_chol_solve_profiled": (C.c_int, [c_f64p, c_f64p, C.c_int32, c_f64p, c_i32p, c_vp, C.c_size_t, c_vp,
                                          c_vp]),
}


class BAOptions(C.Structure):
    """sfm_ba_options (include/sfm_b200.h)."""
    _fields_ = [("max_trials", C.c_int32), ("verbose", C.c_int32), ("ftol", C.c_double),
                ("xtol", C.c_double), ("gtol", C.c_double), ("lambda0", C.c_double), ("reuse_index", C.c_int32),
                ("reserved", C.c_int32)]


class BAResult(C.Structure):
    """sfm_ba_result (include/sfm_b200.h)."""
    _fields_ = [("cost0", C.c_double), ("cost", C.c_double), ("gnorm", C.c_double), ("lambda_", C.c_double),
                ("status", C.c_int32), ("n_linearisations", C.c_int32), ("n_trials", C.c_int32),
                ("n_chol_fail", C.c_int32)]

_ERR = {1: "SFM_E_INVALID", 2: "SFM_E_CUDA", 3: "SFM_E_WORKSPACE", 4: "SFM_E_NCCL",
        5: "SFM_E_NOT_CONVERGED"}


class SfmError(RuntimeError):
    pass


def lib_path() -> str:
    return _build.LIB_PATH


def load(build_if_missing: bool = True):
    """Load (building first when stale and nvcc is available) and type the library."""
    global _lib
    with _lock:
        if _lib is not None:
            return _lib
        path = _build.LIB_PATH
        if build_if_missing and _build.is_stale():
            try:
                _build.build()
            except Exception as e:  # no nvcc on the box: use the shipped .so if there is one
                if not os.path.exists(path):
                    raise SfmError(
                        f"libsfm_b200.so is missing and could not be built ({e}); "
                        "there is no CPU fallback") from e
        if not os.path.exists(path):
            raise SfmError(f"{path} not found; run __graft_entry__.build() (no CPU fallback)")
        lib = C.CDLL(path)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
        return lib


def check(rc: int, what: str):
    if rc != 0:
        msg = load().sfm_last_cuda_error().decode(errors="replace") if rc == 2 else ""
        if rc == 4:
            msg = load().sfm_nccl_last_error().decode(errors="replace")
        raise SfmError(f"{what} failed: {_ERR.get(rc, rc)} {msg}")


def launch_count() -> int:
    return int(load().sfm_launch_count())
