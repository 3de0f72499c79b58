"""Drop-in replacement for the reference module of the same name (imported by PnPRANSAC.py:3). Put this directory
ahead of the reference's `Phase 1/` on sys.path; exports the same names."""
import os as _os
import sys as _sys

_sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__)))))
from sfm_b200.pnp_ransac import LinearPnP  # noqa: E402,F401

_sys.path.pop(0)
