"""Recipe that stages the UNMODIFIED reference (its Phase 1 Python modules and the P3Data inputs the wrapper reads)
under oracle/_ref/ -- TEST / BENCH INFRASTRUCTURE.

The reference is pure Python, so "building" it is a file copy: run in the build container, where /root/reference
exists (`__graft_entry__.build()` calls stage()), outputs only into oracle/_ref/, which is git-ignored (no reference
source enters the history) but NOT gpurun-ignored, so it travels to the GPU box like the built .so files do. There
`bench.py --impl reference` times the reference ITSELF (cpu_baseline.kind = "reference") and the `-m gpu` tests run
the unmodified Wrapper_dev.py on P3Data with the drop-in modules ahead of it on sys.path. Nothing under the product
package imports this; when oracle/_ref/ is absent the callers fall back to the oracle port and say so.
"""
from __future__ import annotations

import glob
import importlib
import os
import shutil
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = os.environ.get("SFM_REF_DIR", "/root/reference/Phase 1")
DST = os.path.join(HERE, "_ref", "phase1")
DATA = os.path.join(DST, "P3Data")

MODULES = ("BundleAdjustment", "DisambiguateCameraPose", "EssentialMatrixFromFundamentalMatrix",
           "EstimateFundamentalMatrix", "ExtractCameraPose", "GetHomographyInliers", "GetInliersRANSAC", "LinearPnP",
           "LinearTriangulation", "NonLinearTriangulation", "NonlinearPnP", "PnPRANSAC", "Utils", "Wrapper_dev")


def stage(verbose: bool = False) -> bool:
    """Copy the reference files when the reference tree is present. Returns True when oracle/_ref is usable."""
    if os.path.isdir(REF_SRC):
        os.makedirs(DATA, exist_ok=True)
        for m in MODULES:
            shutil.copyfile(os.path.join(REF_SRC, m + ".py"), os.path.join(DST, m + ".py"))
        for pat in ("*.png", "matching*.txt", "calibration.txt"):
            for f in glob.glob(os.path.join(REF_SRC, "P3Data", pat)):
                shutil.copyfile(f, os.path.join(DATA, os.path.basename(f)))
        if verbose:
            print(f"staged {len(MODULES)} reference modules + P3Data into {DST}")
    return available()


def available() -> bool:
    return all(os.path.exists(os.path.join(DST, m + ".py")) for m in MODULES)


def data_path() -> str:
    """P3Data directory with the trailing separator Wrapper_dev.py:28 needs."""
    return DATA + os.sep


def stub_matplotlib():
    """matplotlib is not installed; Utils.py imports it at module level (Utils.py:4-5) for drawing only."""
    for name in ("matplotlib", "matplotlib.pyplot", "mpl_toolkits", "mpl_toolkits.mplot3d"):
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    plt = sys.modules["matplotlib.pyplot"]
    if not hasattr(plt, "figure"):
        class _Anything:
            def __getattr__(self, k):
                return lambda *a, **kw: _Anything()

            def __call__(self, *a, **kw):
                return _Anything()
        for fn in ("figure", "savefig", "close", "show", "imshow", "axis", "title", "plot", "scatter", "subplot",
                   "subplots", "legend", "xlabel", "ylabel", "clf", "gca", "gcf", "tight_layout"):
            setattr(plt, fn, _Anything())
    sys.modules["matplotlib"].pyplot = plt
    if not hasattr(sys.modules["mpl_toolkits.mplot3d"], "Axes3D"):
        sys.modules["mpl_toolkits.mplot3d"].Axes3D = object


def load_modules(names=MODULES, first_on_path=()):
    """Import staged reference modules by bare name, in isolation: sys.path is [*first_on_path, oracle/_ref/phase1]
    for the duration, previously imported modules of the same names are dropped first and restored afterwards.
    Returns {name: module}. `first_on_path` is how the drop-in directory is put AHEAD of the reference."""
    if not available():
        raise RuntimeError("oracle/_ref is not staged (run __graft_entry__.build() where /root/reference exists)")
    stub_matplotlib()
    saved = {m: sys.modules.pop(m) for m in MODULES if m in sys.modules}
    saved_path = list(sys.path)
    sys.path[:0] = [*first_on_path, DST]
    try:
        out = {n: importlib.import_module(n) for n in names}
    finally:
        sys.path[:] = saved_path
        mine = {m: sys.modules.pop(m) for m in MODULES if m in sys.modules}
        sys.modules.update(saved)
    out["_all"] = mine
    return out


if __name__ == "__main__":
    print("available" if stage(verbose=True) else "reference tree not found; nothing staged")
