"""TEST INFRASTRUCTURE: the oracle's F-RANSAC over a long hypothesis stream on all host cores, as a process of its
own (so that a pytest process holding a CUDA context never forks):  python tests/oracle_pool.py in.npz out.npz
in.npz: x1, x2, samples, threshold.  out.npz: counts [H], F [H,9], best_idx, best_count, mask, err (of the winner)."""
import multiprocessing as mp
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _chunk(args):
    x1, x2, samples, thr = args
    os.environ.setdefault("OMP_NUM_THREADS", "1")
    from oracle import sfm_oracle as so
    d = so.f_ransac_detailed(x1, x2, thr, samples)
    return np.asarray(d["counts"], np.int64), np.asarray(d["F"], np.float64).reshape(len(samples), 9)


def main(inp, outp):
    from oracle import sfm_oracle as so
    d = np.load(inp)
    x1, x2, samples, thr = d["x1"], d["x2"], d["samples"], float(d["threshold"])
    cores = os.cpu_count() or 1
    per = max(32, -(-len(samples) // (cores * 8)))
    jobs = [(x1, x2, samples[a:a + per], thr) for a in range(0, len(samples), per)]
    with mp.get_context("fork").Pool(cores) as pool:
        res = pool.map(_chunk, jobs, chunksize=1)
    counts = np.concatenate([r[0] for r in res])
    F = np.concatenate([r[1] for r in res])
    best = int(np.argmax(counts))
    err = so.f_errors(F[best].reshape(3, 3), x1, x2)
    np.savez(outp, counts=counts, F=F, best_idx=best, best_count=int(counts[best]), mask=err < thr, err=err)


def run(x1, x2, samples, threshold, tmpdir):
    """Helper for tests: returns the dict of arrays."""
    import subprocess
    inp, outp = os.path.join(tmpdir, "pool_in.npz"), os.path.join(tmpdir, "pool_out.npz")
    np.savez(inp, x1=x1, x2=x2, samples=samples, threshold=threshold)
    subprocess.run([sys.executable, os.path.abspath(__file__), inp, outp], check=True)
    return dict(np.load(outp))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
