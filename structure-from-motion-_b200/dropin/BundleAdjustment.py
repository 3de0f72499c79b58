"""Drop-in replacement for the reference module of the same name (Wrapper_dev.py:14 does
`from BundleAdjustment import perform_bundle_adjustment`). Put this directory ahead of the
reference's `Phase 1/` on sys.path; exports the same names as BundleAdjustment.py:8,43,113."""
import os as _os
import sys as _sys

_sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__)))))
from sfm_b200.bundle import (project_points, bundle_adjustment_residuals,  # noqa: E402,F401
                             perform_bundle_adjustment)

_sys.path.pop(0)
