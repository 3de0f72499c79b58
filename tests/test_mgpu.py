"""Multi-GPU tests: the host-side sharding / plumbing on CPU with gloo (world size 2, runs
everywhere) and, marked gpu, the torchrun parity check when the box has at least two GPUs."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_points_partitions_and_balances():
    from sfm_b200.bundle_mgpu import shard_points
    rng = np.random.default_rng(0)
    counts = rng.integers(0, 9, 1000)                   # some points have no observation at all
    pt_idx = np.repeat(np.arange(1000), counts)
    for world in (1, 2, 3, 8):
        sh = shard_points(pt_idx, 1000, world)
        assert sh[0][0] == 0 and sh[-1][1] == 1000 and sh[0][2] == 0 and sh[-1][3] == len(pt_idx)
        for a, b in zip(sh[:-1], sh[1:]):
            assert a[1] == b[0] and a[3] == b[2]           # contiguous in points and observations
        for p_lo, p_hi, o_lo, o_hi in sh:
            assert np.all((pt_idx[o_lo:o_hi] >= p_lo) & (pt_idx[o_lo:o_hi] < p_hi))
            assert abs((o_hi - o_lo) - len(pt_idx) / world) <= 8          # balanced to one point
    with pytest.raises(ValueError):
        shard_points(np.array([1, 0, 2]), 3, 2)
    # fewer points than ranks: empty tail ranges, still a partition
    sh = shard_points(np.array([0, 0, 1]), 2, 4)
    assert sum(s[3] - s[2] for s in sh) == 3


def _gloo_worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    from sfm_b200.bundle_mgpu import exchange_unique_id, shard_points
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # 1) the 128-byte unique id travels from rank 0 to every rank
        uid = exchange_unique_id(rank, lambda: bytes(range(128)))
        assert uid == bytes(range(128))
        # 2) the packed (count << 32 | ~idx) key: a MAX all-reduce picks the highest count and, among
        #    equal counts, the LOWEST global hypothesis index (GetInliersRANSAC.py:85 strict '>')
        counts = np.array([[3, 7, 7, 1], [7, 2, 7, 7]])[rank]
        h = counts.shape[0]
        keys = (counts.astype(np.int64) << 32) | (0xFFFFFFFF - (np.arange(h) + rank * h))
        key = torch.tensor([keys.max()], dtype=torch.int64)
        dist.all_reduce(key, op=dist.ReduceOp.MAX)
        k = int(key.item())
        assert (k >> 32) == 7 and 0xFFFFFFFF - (k & 0xFFFFFFFF) == 1
        # 3) point shards: summing per-rank partial normal-equation terms reproduces the global sum
        rng = np.random.default_rng(5)
        pt_idx = np.sort(rng.integers(0, 50, 400))
        vals = rng.normal(size=400)
        p_lo, p_hi, o_lo, o_hi = shard_points(pt_idx, 50, world)[rank]
        part = torch.tensor([vals[o_lo:o_hi].sum()], dtype=torch.float64)
        dist.all_reduce(part)
        assert abs(float(part) - vals.sum()) < 1e-12
        q.put((rank, "ok"))
    except Exception as e:                                   # surface the failure in the parent
        q.put((rank, repr(e)))
    finally:
        dist.destroy_process_group()


def test_shard_range_is_a_balanced_partition():
    from sfm_b200.shard import shard_range
    for n in (0, 1, 2, 7, 8, 1000, 1_000_003):
        for world in (1, 2, 3, 8):
            r = [shard_range(n, k, world) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(r[:-1], r[1:]))
            sizes = [b - a for a, b in r]
            assert max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)
    with pytest.raises(ValueError):
        shard_range(5, 2, 2)


def _gloo_map_worker(rank, world, port, q):
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    from sfm_b200.shard import map_sharded, shard_range
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(3)
        for n in (11, 2, 1, 0, 1000):                                    # odd split, fewer problems than ranks, none
            a, b = rng.normal(size=(n, 2)), rng.normal(size=(n, 3))
            seen = []

            def fn(x, y):
                seen.append(len(x))
                return np.hstack([2.0 * x, y - 1.0, (x[:, :1] * y[:, :1])])
            out = map_sharded(fn, n, (a, b))
            lo, hi = shard_range(n, rank, world)
            assert seen == ([hi - lo] if hi > lo else [])                 # only this rank's chunk was computed here
            if n:
                assert out.shape == (n, 6) and np.array_equal(out, np.hstack([2.0 * a, b - 1.0, a[:, :1] * b[:, :1]]))
            else:
                assert out.shape[0] == 0
        out = map_sharded(lambda x: np.einsum("ni,nj->nij", x, x), 5, (np.arange(15.0).reshape(5, 3),))   # trailing shape (3,3)
        assert out.shape == (5, 3, 3) and out[4, 2, 2] == 14.0 ** 2
        q.put((rank, "ok"))
    except Exception as e:                                                # noqa: BLE001
        q.put((rank, repr(e)))
    finally:
        dist.destroy_process_group()


def test_gloo_world2_independent_problems_are_chunked_and_gathered():
    """SURVEY 8e, second row: per-point / per-registration solvers shard as contiguous chunks with an all-gather of
    the results -- the plumbing (sfm_b200.shard.map_sharded) on two gloo ranks with a stand-in for the kernel."""
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + os.getpid() % 200
    procs = [ctx.Process(target=_gloo_map_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(0, "ok"), (1, "ok")], res


def test_gloo_world2_plumbing():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 400)
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(0, "ok"), (1, "ok")], res


@pytest.mark.gpu
def test_mgpu_parity_two_gpus(cuda_device):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29533", os.path.join(ROOT, "tests", "mgpu_check.py")]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-3000:]
