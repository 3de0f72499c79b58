"""CPU-only: host helpers of the Python layer that need no device (sfm_b200.device)."""
import numpy as np

from sfm_b200 import device


def test_arrays_equal_matches_numpy_on_small_and_sliced_arrays():
    rng = np.random.default_rng(0)
    small = rng.integers(0, 1000, 1000).astype(np.int32)
    assert device.arrays_equal(small, small.copy()) and not device.arrays_equal(small, small[:-1])
    big = rng.integers(0, 1 << 30, 3_000_000).astype(np.int32)          # above PAR_MIN_BYTES: the threaded path
    assert big.nbytes >= device.PAR_MIN_BYTES
    assert device.arrays_equal(big, big.copy())
    for pos in (0, 1, len(big) // 4, len(big) // 2 + 1, len(big) - 1):   # a difference in every slice, and at the ends
        other = big.copy()
        other[pos] ^= 1
        assert not device.arrays_equal(big, other), pos
    assert not device.arrays_equal(big, big[:-1])


def test_slices_cover_the_range_exactly():
    for n in (1, 2, 7, 8, 9, 1000, 4_000_000):
        for parts in (1, 3, 4, 8):
            sl = device._slices(n, parts)
            assert sl[0][0] == 0 and sl[-1][1] == n and len(sl) <= parts
            assert all(a < b for a, b in sl) and all(sl[k][1] == sl[k + 1][0] for k in range(len(sl) - 1))
