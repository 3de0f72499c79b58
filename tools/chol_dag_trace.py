#!/usr/bin/env python
"""Reads a K11s task trace (SFM_CHOL_DAG_TRACE=<file>, written by sfm_chol_solve_structured) and prints where
the time of the factorisation goes: the chain of diagonal tasks, how long each waits for its last operand, and
per-type busy time."""
import sys

import numpy as np

raw = open(sys.argv[1], "rb").read()
nt = int(np.frombuffer(raw[:8], np.int64)[0])
tasks = np.frombuffer(raw[8:8 + 48 * nt], np.int32).reshape(nt, 12)
tr = np.frombuffer(raw[8 + 48 * nt:], np.uint64).reshape(nt, 4).astype(np.int64)
ran = tr[:, 2] > 0
t0 = tr[ran, 0].min()
start, ready, end, cta = (tr[:, 0] - t0) / 1e3, (tr[:, 1] - t0) / 1e3, (tr[:, 2] - t0) / 1e3, tr[:, 3]
ty = tasks[:, 0]
names = {0: "diag", 1: "off", 2: "back", 3: "part"}
fwd = ran & (ty != 2)
print(f"tasks {nt}: forward ends at {end[fwd].max():.1f} us, backward from {start[ran & (ty == 2)].min():.1f} to {end[ran & (ty == 2)].max():.1f} us")
for k, nm in names.items():
    m = ran & (ty == k)
    if not m.any():
        continue
    n_ent = tasks[m, 8] - tasks[m, 7]
    print(f"  {nm:5s} n={m.sum():5d} entries/task {n_ent.mean():5.1f}  wait+accumulate {np.mean(ready[m] - start[m]):6.2f} us  "
          f"tail {np.mean(end[m] - ready[m]):6.2f} us  (tail median {np.median(end[m] - ready[m]):.2f})")
# the chain of diagonal tasks ordered by end time: gaps between consecutive ends on the critical path
d = np.where(ran & (ty == 0))[0]
d = d[np.argsort(end[d])]
print("last 40 diagonal tasks by end time: column, entries, start, operands-complete, end, gap to previous end")
prev = None
for k in d[-40:]:
    gap = end[k] - prev if prev is not None else 0.0
    print(f"  col {tasks[k, 1]:4d} ent {tasks[k, 8] - tasks[k, 7]:3d} cta {cta[k]:3d}  {start[k]:8.1f} {ready[k]:8.1f} {end[k]:8.1f}  gap {gap:6.2f}  factor {end[k] - ready[k]:5.2f}")
    prev = end[k]
busy = np.zeros(int(cta[ran].max()) + 1)
for k in np.where(fwd)[0]:
    busy[cta[k]] += end[k] - start[k]
print(f"forward occupancy: mean CTA busy {busy.mean():.1f} us, max {busy.max():.1f} us of {end[fwd].max():.1f}")
