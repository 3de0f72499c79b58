"""Drop-in replacement for the reference module of the same name (Wrapper_dev.py:7 does
`from GetInliersRANSAC import get_inliers_ransac`). Put this directory ahead of the reference's
`Phase 1/` on sys.path. Exports the same names as GetInliersRANSAC.py:5,109."""
import os as _os
import sys as _sys

_sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__)))))
from sfm_b200.ransac import GetInliersRANSAC, get_inliers_ransac  # noqa: E402,F401

_sys.path.pop(0)
