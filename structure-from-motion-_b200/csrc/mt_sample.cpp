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
// tests/test_abi_and_sampler.py checks the stream against random.sample itself.
//
// Since the scoring kernel got faster than this replay, the replay is the end-to-end limiter of
// GetInliersRANSAC, so it is written for throughput: the generator is advanced a whole 624-word
// generation at a time (three dependence-free loops the compiler vectorises) with the tempering
// applied in bulk, and the set-based branch of random.sample (n > setsize: every RANSAC call of
// the reference on real data) consumes that buffer without data-dependent branches (see
// sample_rows_set), so the 24 % rejection rate of getrandbits at n = 50,000 of 65,536 costs no
// branch mispredictions and nothing waits on a loop-carried chain.
#include "../../include/sfm_b200.h"
#include <math.h>
#include <string.h>
#include <vector>

// The entry point is compiled twice, for x86-64-v3 (AVX2, BMI2) and for the baseline ISA, and the loader picks by
// CPUID (GCC function multiversioning): the generator and compaction loops vectorise 8 wide instead of 4, about a
// quarter off the replay. Everything hot is forced inline so that it is compiled with the clone's ISA.
#if defined(__GNUC__) && defined(__x86_64__) && !defined(__CUDACC__)
#define SFM_MT_CLONES __attribute__((target_clones("arch=x86-64-v3", "avx2", "default")))
#define SFM_MT_INLINE __attribute__((always_inline)) inline
#else
#define SFM_MT_CLONES
#define SFM_MT_INLINE inline
#endif

namespace {

constexpr int MT_N = 624, MT_M = 397;

struct MT {
    uint32_t *mt;          // CPython's state words (current generation, untempered)
    int idx;               // next word of the generation to hand out (624 = regenerate first)
    uint32_t tb[MT_N];     // tempered outputs of the current generation (valid from idx0 on)

    static inline uint32_t mix(uint32_t a, uint32_t b, uint32_t far) {
        uint32_t y = (a & 0x80000000u) | (b & 0x7fffffffu);
        return far ^ (y >> 1) ^ ((0u - (y & 1u)) & 0x9908b0dfu);
    }
    SFM_MT_INLINE void temper_from(int from) {
        for (int i = from; i < MT_N; ++i) {
            uint32_t y = mt[i];
            y ^= (y >> 11);
            y ^= (y << 7) & 0x9d2c5680u;
            y ^= (y << 15) & 0xefc60000u;
            y ^= (y >> 18);
            tb[i] = y;
        }
    }
    SFM_MT_INLINE void regenerate() {
        // in place, as genrand_uint32 does; each loop reads only words the loop itself does not
        // write ahead of the read (distances +1 / +397 forward, -227 backward in blocks <= 227)
        uint32_t *s = mt;
        for (int k = 0; k < MT_N - MT_M; ++k) s[k] = mix(s[k], s[k + 1], s[k + MT_M]);
        for (int k = MT_N - MT_M; k < 2 * (MT_N - MT_M); ++k)
            s[k] = mix(s[k], s[k + 1], s[k - (MT_N - MT_M)]);
        for (int k = 2 * (MT_N - MT_M); k < MT_N - 1; ++k)
            s[k] = mix(s[k], s[k + 1], s[k - (MT_N - MT_M)]);
        s[MT_N - 1] = mix(s[MT_N - 1], s[0], s[MT_M - 1]);
        idx = 0;
        temper_from(0);
    }
    void init(uint32_t *state, int index) {
        mt = state;
        idx = index;
        if (idx < MT_N) temper_from(idx);
    }
    inline uint32_t next() {
        if (idx >= MT_N) regenerate();
        return tb[idx++];
    }
    // _randbelow_with_getrandbits(n), n > 0, n < 2^31
    inline int32_t randbelow(uint32_t n) {
        int bits = 32 - __builtin_clz(n);  // n.bit_length()
        uint32_t r = next() >> (32 - bits);
        while (r >= n) r = next() >> (32 - bits);
        return (int32_t)r;
    }
};

// Set-based branch of random.sample for small k: selected = set(); for each of the k slots draw
// j = randbelow(n) until j is not in the set. Both rejections (r >= n inside randbelow, j already
// selected) just skip a generator word, so the stream of ACCEPTED words (r < n), cut into rows
// of k distinct values, is the result. Two passes per 624-word generation, neither with a
// loop-carried dependence longer than one add:
//   1. branch-free compaction of the words with r < n (24 % are rejected at n = 50,000) together
//      with their position in the generation;
//   2. rows of k: all-pairs distinctness test; a row with a repeated value (a few in 10^4) is
//      redone slot by slot exactly as CPython would.
// A row that cannot be completed from the words at hand is carried into the next generation.
template <int K>
SFM_MT_INLINE bool row_distinct(const int32_t *v, int k) {
    const int kk = K ? K : k;
    int dup = 0;
    for (int a = 1; a < kk; ++a)
        for (int b = 0; b < a; ++b) dup |= (v[a] == v[b]);
    return dup == 0;
}

template <int K>
SFM_MT_INLINE void sample_rows_set(MT &g, uint32_t n, int k, int64_t count, int32_t *out) {
    const int kk = K ? K : k;
    const int shift = __builtin_clz(n);  // 32 - n.bit_length()
    // accepted values and their word positions: carry (< 2 rows) + one generation
    std::vector<int32_t> val((size_t)MT_N + 4 * 64 + 8);
    std::vector<int32_t> pos(val.size());
    int have = 0;  // valid entries in val/pos
    int64_t c = 0;
    while (c < count) {
        if (g.idx >= MT_N) g.regenerate();
        if ((size_t)have + MT_N > val.size()) {  // a row with very many repeats (n barely above k)
            val.resize((size_t)have + MT_N + 64);
            pos.resize(val.size());
        }
        {
            int w = have;
            int32_t *vp = val.data(), *pp = pos.data();
            const uint32_t *tb = g.tb;
            for (int p = g.idx; p < MT_N; ++p) {
                uint32_t r = tb[p] >> shift;
                vp[w] = (int32_t)r;
                pp[w] = p;
                w += (r < n);
            }
            have = w;
        }
        int e = 0;           // first unconsumed entry
        int last = -1;       // word position of the last consumed entry
        while (c < count && have - e >= kk) {
            const int32_t *v = val.data() + e;
            int32_t *res = out + c * kk;
            if (row_distinct<K>(v, kk)) {
                for (int t = 0; t < kk; ++t) res[t] = v[t];
                e += kk;
            } else {
                int i = 0, q = e;
                for (; q < have && i < kk; ++q) {
                    int32_t j = val[q];
                    bool dup = false;
                    for (int t = 0; t < i; ++t) dup |= (res[t] == j);
                    if (!dup) res[i++] = j;
                }
                if (i < kk) break;  // needs words of the next generation: carry the whole row
                e = q;
            }
            last = pos[e - 1];
            ++c;
        }
        if (c == count) {
            // the last row always ends on a word of this generation (it would have been emitted
            // a generation earlier otherwise)
            g.idx = last + 1;
            return;
        }
        // carry the unconsumed tail; their positions are never needed again
        int rest = have - e;
        for (int t = 0; t < rest; ++t) val[t] = val[e + t];
        have = rest;
        g.idx = MT_N;
    }
}

}  // namespace

extern "C" SFM_MT_CLONES int sfm_mt_sample_stream(uint32_t *mt_state, int32_t *mt_index, int32_t n, int32_t k,
                                    int64_t count, int32_t *out) {
    if (!mt_state || !mt_index || !out || n <= 0 || k < 0 || k > n || count < 0 || k > 64)
        return SFM_E_INVALID;
    if (*mt_index < 0 || *mt_index > 624) return SFM_E_INVALID;
    MT g;
    g.init(mt_state, *mt_index);
    // random.sample: setsize = 21; if k > 5: setsize += 4 ** ceil(log(k*3, 4))
    long setsize = 21;
    if (k > 5) {
        long p = 1;  // smallest power of 4 >= 3k (== 4**ceil(log(3k, 4)) for the k <= 64 accepted
        while (p < 3L * k) p *= 4;  // here; 3k is never an exact power of 4, so CPython's float
        setsize += p;               // rounding cannot matter)
    }
    if (k == 0 || count == 0) { *mt_index = g.idx; return SFM_OK; }
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
    } else if (k == 8) {
        sample_rows_set<8>(g, (uint32_t)n, k, count, out);
    } else if (k == 4) {
        sample_rows_set<4>(g, (uint32_t)n, k, count, out);
    } else if (k <= 64) {
        sample_rows_set<0>(g, (uint32_t)n, k, count, out);
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
