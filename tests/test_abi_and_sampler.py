"""CPU-only: the C-ABI library loads and exports every symbol include/sfm_b200.h declares, and
the MT19937 replay reproduces random.sample's stream bit for bit."""
import os
import random
import re

import numpy as np
import pytest

from conftest import ROOT
from sfm_b200 import _lib, sampler


def _declared_symbols():
    txt = open(os.path.join(ROOT, "include", "sfm_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(sfm_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol():
    lib = _lib.load()
    declared = _declared_symbols()
    assert len(declared) >= 10
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in sfm_b200.h but not exported"
    assert sorted(_lib.SIGNATURES) == declared
    assert lib.sfm_abi_version() >= 1


def test_invalid_arguments_are_rejected_without_a_gpu():
    lib = _lib.load()
    assert lib.sfm_f8pt_batch(None, 8, None, 1, None, None) == 1
    assert lib.sfm_f_ransac(None, 8, None, 1, 0.1, None, None, None, None, None, None, None) == 1
    assert lib.sfm_fundamental_npt_workspace_bytes(100) == 7200


def test_sample_stream_matches_cpython():
    for n, k in [(50000, 8), (558, 8), (85, 8), (86, 8), (20, 8), (8, 8), (2000, 4), (21, 4), (22, 4),
                 (4, 4), (5, 4), (3, 4), (6, 8)]:
        random.seed(9000 + n)
        kk = min(k, n)
        a = np.array([random.sample(range(n), kk) for _ in range(300)])
        after_a = random.random()
        random.seed(9000 + n)
        b = sampler.sample_stream(n, k, 300)
        after_b = random.random()
        assert np.array_equal(a, b), (n, k)
        assert after_a == after_b, "global generator state must advance exactly like the reference"


def test_sample_stream_two_pass_replay_edges():
    """The bulk replay (csrc/mt_sample.cpp: compaction of accepted words, rows of k, slow path for rows
    with a repeated value, rows carried across 624-word generations): n a power of two and one below,
    n barely above CPython's set threshold (many repeats), k other than 4 / 8, every starting offset
    class inside a generation, counts that end a call in the middle of a generation."""
    for n, k in [(65536, 8), (65535, 8), (86, 8), (100, 8), (22, 4), (23, 3), (40, 5), (200, 6), (1000, 20),
                 (500, 64), (64, 64)]:
        random.seed(n * 131 + k)
        a = np.array([random.sample(range(n), k) for _ in range(1500)])
        after_a = random.random()
        random.seed(n * 131 + k)
        b = sampler.sample_stream(n, k, 1500)
        assert np.array_equal(a, b), (n, k)
        assert after_a == random.random(), (n, k)
    for off in range(0, 700, 37):
        for cnt in (16, 17, 77, 500):
            random.seed(5)
            [random.getrandbits(32) for _ in range(off)]
            a = np.array([random.sample(range(50000), 8) for _ in range(cnt)])
            after_a = random.random()
            random.seed(5)
            [random.getrandbits(32) for _ in range(off)]
            b = sampler.sample_stream(50000, 8, cnt)
            assert np.array_equal(a, b) and after_a == random.random(), (off, cnt)


def test_sample_stream_short_runs_use_python_path():
    random.seed(5)
    a = [random.sample(range(100), 8) for _ in range(3)]
    random.seed(5)
    b = sampler.sample_stream(100, 8, 3)
    assert np.array_equal(np.array(a), b)


def test_stream_replayer_chunks_equal_the_reference_stream():
    """Chunked replay with one getstate/setstate: same index sets, same final generator state as
    random.sample called once per hypothesis."""
    from sfm_b200 import sampler
    n, k, total = 977, 8, 5000
    random.seed(99)
    ref = np.array([random.sample(range(n), k) for _ in range(total)], dtype=np.int32)
    state_ref = random.getstate()
    random.seed(99)
    rep = sampler.StreamReplayer(n, k)
    out = np.empty((total, k), dtype=np.int32)
    off = 0
    for c in (7, 100, 1, 2048, total - 2156):
        rep.draw(c, out[off:off + c])
        off += c
    rep.finish()
    assert off == total and np.array_equal(out, ref)
    assert random.getstate() == state_ref


def test_sample_stream_property_random_shapes():
    """Property test of the bulk replay against CPython itself: random population sizes (both branches of
    random.sample), sample sizes, counts and starting offsets inside a 624-word generation."""
    from hypothesis import given, settings, strategies as st

    @settings(max_examples=60, deadline=None)
    @given(n=st.integers(1, 70000), k=st.integers(1, 12), count=st.integers(16, 400), off=st.integers(0, 1300),
           seed=st.integers(0, 2 ** 31 - 1))
    def check(n, k, count, off, seed):
        k = min(k, n)
        random.seed(seed)
        [random.getrandbits(32) for _ in range(off)]
        a = np.array([random.sample(range(n), k) for _ in range(count)])
        after_a = random.random()
        random.seed(seed)
        [random.getrandbits(32) for _ in range(off)]
        b = sampler.sample_stream(n, k, count)
        assert np.array_equal(a, b), (n, k, count, off)
        assert after_a == random.random(), (n, k, count, off)

    check()


def test_stream_chunk_schedule_covers_the_stream():
    """ransac.f_ransac_stream scores the hypotheses chunk by chunk: the chunk sizes must be positive, sum to the
    stream length, never end on a sliver, and the first one stays small (the only one the GPU waits for)."""
    from sfm_b200 import ransac
    for h in (1, 2, 4095, 4096, 4097, 16384, 20000, 99999, 100000, 123457, 1000000):
        for chunks in (1, 2, 3, 8):
            s = ransac._chunk_schedule(h, chunks)
            assert sum(s) == h and min(s) > 0, (h, chunks, s)
            assert s[0] <= ransac.STREAM_FIRST_CHUNK + ransac.STREAM_FIRST_CHUNK // 3
            if len(s) > 1:
                assert s[-1] >= s[-2] // 4, (h, chunks, s)


def test_stream_replayer_with_a_population_per_draw():
    """pnp_ransac_batch replays the samples of consecutive registrations of DIFFERENT sizes with one replayer
    (StreamReplayer.draw(..., n=...)): the draws and the generator state afterwards must equal those of the
    reference's consecutive random.sample calls, including the pool branch of random.sample (n <= 21)."""
    import random
    from sfm_b200 import sampler
    sizes = [300, 2000, 97, 50000, 5, 4, 21, 22]
    random.seed(9)
    ref = [np.array([random.sample(range(n), 4) for _ in range(60)]) for n in sizes]
    after = random.random()
    random.seed(9)
    rep = sampler.StreamReplayer(min(sizes), 4)
    outs = []
    for n in sizes:
        o = np.empty((60, 4), np.int32)
        rep.draw(60, o, n=n)
        outs.append(o)
    rep.finish()
    assert random.random() == after
    assert all(np.array_equal(a, b) for a, b in zip(ref, outs))
    rep = sampler.StreamReplayer(10, 4)
    try:
        with pytest.raises(ValueError):
            rep.draw(1, np.empty((1, 4), np.int32), n=3)
    finally:
        rep.finish()
