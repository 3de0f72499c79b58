"""Test-only host build of the kernels' per-thread arithmetic (see hostsim.cpp)."""
import ctypes as C
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "_hostsim.so")
CSRC = os.path.join(HERE, "..", "..", "structure-from-motion-_b200", "csrc")


def load():
    srcs = [os.path.join(HERE, "hostsim.cpp")] + [os.path.join(CSRC, f) for f in os.listdir(CSRC)
                                                  if f.endswith((".cuh", ".inc", ".cu"))]
    if not os.path.exists(SO) or os.path.getmtime(SO) < max(os.path.getmtime(s) for s in srcs):
        subprocess.run(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-ffp-contract=off", "-o", SO,
                        os.path.join(HERE, "hostsim.cpp")], check=True)
    return C.CDLL(SO)
