"""Drop-in replacement for the reference module of the same name (GetInliersRANSAC.py:3 does
`from EstimateFundamentalMatrix import EstimateFundamentalMatrix`)."""
import os as _os
import sys as _sys

_sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__)))))
from sfm_b200.ransac import EstimateFundamentalMatrix  # noqa: E402,F401

_sys.path.pop(0)
