"""Workload constants and input generators shared by both arms of bench.py. Imports numpy and oracle/scenes.py
(data generation) only -- in particular not the product package, so the reference arm never maps libsfm_b200.so."""
from __future__ import annotations

import os
import random
import sys

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

N_PTS = 50_000
N_HYP = 100_000          # per GPU (weak scaling: the hypothesis stream grows with N)
THRESH = 0.06            # the reference default (GetInliersRANSAC.py:5, Wrapper_dev.py:105)
DATA_SEED = 20260102
RNG_SEED = 12345
METRIC = "RANSAC F-hypotheses/sec"
UNIT = "hypotheses/s"
WORKLOAD = "C2 two-view F-RANSAC: 50k correspondences, 30% outliers, 100k hypotheses/GPU"

C3 = dict(n_cams=50, n_pts=20_000, obs_per_pt=10, seed=20260103, layout="arc", window=None)
C5 = dict(n_cams=1000, n_pts=500_000, obs_per_pt=8, seed=20260106, layout="ring", window=40)


def two_view_scene():
    from oracle import scenes
    return scenes.two_view(n=N_PTS, seed=DATA_SEED)


def python_sample_stream(n_hyp_total):
    """The reference's own draw (GetInliersRANSAC.py:55), one random.sample per hypothesis, from RNG_SEED."""
    random.seed(RNG_SEED)
    return np.array([random.sample(range(N_PTS), 8) for _ in range(n_hyp_total)], dtype=np.int32)


def chunk_states(n_chunks, per_chunk):
    """random.getstate() at the start of every chunk of `per_chunk` hypotheses of the RNG_SEED stream, so that
    independent worker processes replay consecutive pieces of the ONE stream the GPU arm scores."""
    random.seed(RNG_SEED)
    states = []
    for _ in range(n_chunks):
        states.append(random.getstate())
        for _ in range(per_chunk):
            random.sample(range(N_PTS), 8)
    return states
