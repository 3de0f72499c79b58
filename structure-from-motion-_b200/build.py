"""In-tree build of libsfm_b200.so (nvcc, sm_100a only).

The library is built next to this file so that it travels with the repository snapshot to
the GPU box (a JIT cache under ~/.cache would not). nvcc cross-compiles without a GPU.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ_DIR = os.path.join(HERE, "build")
LIB_PATH = os.path.join(HERE, "libsfm_b200.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC,-O3,-fno-math-errno",
    "-Xptxas", "-v",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: libsfm_b200.so cannot be built (there is no CPU fallback)")


def sources() -> list[str]:
    return sorted(
        os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cpp"))
    )


def _deps_mtime() -> float:
    m = 0.0
    for root in (CSRC, os.path.join(HERE, "..", "include")):
        for f in os.listdir(root):
            if f.endswith((".cu", ".cuh", ".cpp", ".h")):
                m = max(m, os.path.getmtime(os.path.join(root, f)))
    return max(m, os.path.getmtime(os.path.abspath(__file__)))


def is_stale() -> bool:
    return (not os.path.exists(LIB_PATH)) or os.path.getmtime(LIB_PATH) < _deps_mtime()


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile every csrc/*.cu|*.cpp for sm_100a and link libsfm_b200.so. Returns its path."""
    if not force and not is_stale():
        return LIB_PATH
    nvcc = _nvcc()
    os.makedirs(OBJ_DIR, exist_ok=True)
    hdr_m = _deps_mtime()
    srcs = sources()
    objs = [os.path.join(OBJ_DIR, os.path.basename(s) + ".o") for s in srcs]

    def compile_one(pair):
        src, obj = pair
        if not force and os.path.exists(obj) and os.path.getmtime(obj) >= hdr_m:
            return ""
        # .cpp sources are host-only (the sample replay, the K11s planner): nvcc hands them to g++ as they are, so
        # GCC attributes (function multiversioning in mt_sample.cpp) are seen by the compiler that implements them
        lang = [] if src.endswith(".cpp") else ["-x", "cu"]
        cmd = [nvcc, *NVCC_FLAGS, *lang, "-c", src, "-o", obj]
        p = subprocess.run(cmd, capture_output=True, text=True)
        if p.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src}:\n{p.stdout}\n{p.stderr}")
        return p.stderr

    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        logs = list(ex.map(compile_one, zip(srcs, objs)))
    with open(os.path.join(OBJ_DIR, "ptxas.log"), "w") as f:
        f.write("\n".join(logs))
    if verbose:
        sys.stderr.write("\n".join(logs))
    link = [nvcc, "-shared", "-o", LIB_PATH + ".tmp", *objs, "-ldl",
            "-gencode", "arch=compute_100a,code=sm_100a"]
    p = subprocess.run(link, capture_output=True, text=True)
    if p.returncode != 0:
        raise RuntimeError(f"link failed:\n{p.stdout}\n{p.stderr}")
    os.replace(LIB_PATH + ".tmp", LIB_PATH)
    return LIB_PATH


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
