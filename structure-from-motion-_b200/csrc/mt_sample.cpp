// mt_sample.cpp -- host-side replay of CPython's random.sample(range(n), k) stream.
//
// The reference draws its RANSAC samples from the interpreter-global Mersenne Twister, one
// random.sample(range(N), min(8,N)) per F hypothesis (GetInliersRANSAC.py:55) and one
// random.sample(range(N), min(4,N)) per PnP hypothesis (PnPRANSAC.py:49); the homography
// pre-filter of the wrapper interleaves on the same stream (GetHomographyInliers.py:126).
// Parity therefore needs the *same* draws. Calling random.sample 100k times costs ~0.9 s of
// interpreter time, 30x the GPU time of the whole RANSAC, so the drop-in copies the MT19937
// state out of random.getstate(), advances it here exactly as CPython 3.x would
// (Lib/random.py: sample() -> _randbelow_with_getrandbits() -> getrandbits(k) =
// genrand_uint32() >> (32-k)), and writes the state back with random.setstate().
// tests/test_sampler.py checks the stream against random.sample itself.
#include "../../include/sfm_b200.h"
#include <math.h>
#include <vector>

namespace {

struct MT {
    uint32_t *mt;
    int idx;
    inline uint32_t next() {
        const int N = 624, M = 397;
        if (idx >= N) {
            int kk;
            uint32_t y;
            for (kk = 0; kk < N - M; kk++) {
                y = (mt[kk] & 0x80000000u) | (mt[kk + 1] & 0x7fffffffu);
                mt[kk] = mt[kk + M] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
            }
            for (; kk < N - 1; kk++) {
                y = (mt[kk] & 0x80000000u) | (mt[kk + 1] & 0x7fffffffu);
                mt[kk] = mt[kk + (M - N)] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
            }
            y = (mt[N - 1] & 0x80000000u) | (mt[0] & 0x7fffffffu);
            mt[N - 1] = mt[M - 1] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
            idx = 0;
        }
        uint32_t y = mt[idx++];
        y ^= (y >> 11);
        y ^= (y << 7) & 0x9d2c5680u;
        y ^= (y << 15) & 0xefc60000u;
        y ^= (y >> 18);
        return y;
    }
    // _randbelow_with_getrandbits(n), n > 0, n < 2^31
    inline int32_t randbelow(uint32_t n) {
        int bits = 32 - __builtin_clz(n);  // n.bit_length()
        uint32_t r = next() >> (32 - bits);
        while (r >= n) r = next() >> (32 - bits);
        return (int32_t)r;
    }
};

}  // namespace

extern "C" int sfm_mt_sample_stream(uint32_t *mt_state, int32_t *mt_index, int32_t n, int32_t k,
                                    int64_t count, int32_t *out) {
    if (!mt_state || !mt_index || !out || n <= 0 || k < 0 || k > n || count < 0 || k > 64)
        return SFM_E_INVALID;
    if (*mt_index < 0 || *mt_index > 624) return SFM_E_INVALID;
    MT g{mt_state, *mt_index};
    // random.sample: setsize = 21; if k > 5: setsize += 4 ** ceil(log(k*3, 4))
    long setsize = 21;
    if (k > 5) {
        int e = 0;  // smallest e with 4^e >= 3k  (== ceil(log(3k, 4)) for the k <= 64 accepted here;
        long p = 1;  // 3k is never an exact power of 4, so float rounding in CPython cannot matter)
        while (p < 3L * k) { p *= 4; ++e; }
        (void)e;
        setsize += p;
    }
    if (k == 0) { *mt_index = g.idx; return SFM_OK; }
    if (n <= setsize) {
        std::vector<int32_t> pool((size_t)n);
        for (int64_t c = 0; c < count; ++c) {
            for (int32_t i = 0; i < n; ++i) pool[i] = i;
            int32_t *res = out + c * k;
            for (int32_t i = 0; i < k; ++i) {
                int32_t j = g.randbelow((uint32_t)(n - i));
                res[i] = pool[j];
                pool[j] = pool[n - i - 1];
            }
        }
    } else {
        for (int64_t c = 0; c < count; ++c) {
            int32_t *res = out + c * k;
            for (int32_t i = 0; i < k; ++i) {
                int32_t j;
                bool dup;
                do {
                    j = g.randbelow((uint32_t)n);
                    dup = false;
                    for (int32_t t = 0; t < i; ++t) dup |= (res[t] == j);
                } while (dup);
                res[i] = j;
            }
        }
    }
    *mt_index = g.idx;
    return SFM_OK;
}
