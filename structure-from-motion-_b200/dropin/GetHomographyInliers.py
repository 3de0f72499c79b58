"""Drop-in replacement for the reference module of the same name (Wrapper_dev.py:6 imports it with `*`).
Put this directory ahead of the reference's `Phase 1/` on sys.path; exports the same names as
GetHomographyInliers.py:4,85."""
import os as _os
import sys as _sys

_sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__)))))
from sfm_b200.homography import find_homography, get_homography_inliers  # noqa: E402,F401

_sys.path.pop(0)
