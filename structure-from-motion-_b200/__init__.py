"""structure-from-motion-_b200: B200-native numerical core for the Phase-1 SfM hot path.

The directory name follows the repository contract and is not a Python identifier; import
it as ``sfm_b200`` (the thin alias package at the repository root).

Layout:
  csrc/      hand-written sm_100a kernels + the C ABI (include/sfm_b200.h)
  _lib.py    ctypes binding (no fallback)
  device.py  torch-backed device buffers / streams (plumbing only)
  ransac.py  GetInliersRANSAC / EstimateFundamentalMatrix host logic
  dropin/    same-named modules Wrapper_dev.py imports (the drop-in boundary)
"""
__version__ = "0.1.0"
