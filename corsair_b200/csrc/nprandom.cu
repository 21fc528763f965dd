// Host-side replay of the NumPy draws symmetric_cut4 makes (utils/symmetry.py:193):
//     np.random.seed(s); np.random.choice(n, size, replace=False)
// i.e. RandomState's legacy path: MT19937 seeded by init_genrand(s), permutation(n) = Fisher-Yates
// from the top with the masked-rejection bounded draw (32-bit outputs for n <= 2^32), first `size`
// entries.  Lets cb_sym_pose draw its seed rows without the interpreter; pinned against NumPy in
// tests/test_abi.py.  No device code.
#include "common.cuh"

#include <vector>

namespace {

struct Mt19937 {
    uint32_t key[624];
    int pos;
    explicit Mt19937(uint32_t seed) {
        for (int i = 0; i < 624; ++i) {
            key[i] = seed;
            seed = 1812433253u * (seed ^ (seed >> 30)) + (uint32_t)i + 1u;
        }
        pos = 624;
    }
    void refill() {
        constexpr uint32_t kUpper = 0x80000000u, kLower = 0x7fffffffu, kMatrix = 0x9908b0dfu;
        int i = 0;
        for (; i < 624 - 397; ++i) {
            const uint32_t y = (key[i] & kUpper) | (key[i + 1] & kLower);
            key[i] = key[i + 397] ^ (y >> 1) ^ ((y & 1u) ? kMatrix : 0u);
        }
        for (; i < 623; ++i) {
            const uint32_t y = (key[i] & kUpper) | (key[i + 1] & kLower);
            key[i] = key[i + (397 - 624)] ^ (y >> 1) ^ ((y & 1u) ? kMatrix : 0u);
        }
        const uint32_t y = (key[623] & kUpper) | (key[0] & kLower);
        key[623] = key[396] ^ (y >> 1) ^ ((y & 1u) ? kMatrix : 0u);
        pos = 0;
    }
    uint32_t next() {
        if (pos == 624) refill();
        uint32_t y = key[pos++];
        y ^= y >> 11;
        y ^= (y << 7) & 0x9d2c5680u;
        y ^= (y << 15) & 0xefc60000u;
        y ^= y >> 18;
        return y;
    }
    // uniform integer in [0, max]: smallest bit mask >= max, redraw until the masked value fits
    uint32_t interval(uint32_t max) {
        if (max == 0) return 0;
        uint32_t mask = max;
        mask |= mask >> 1;
        mask |= mask >> 2;
        mask |= mask >> 4;
        mask |= mask >> 8;
        mask |= mask >> 16;
        uint32_t v;
        while ((v = next() & mask) > max) {}
        return v;
    }
};

}  // namespace

extern "C" int cb_np_seeded_choices(uint32_t seed, const int32_t* n, int count, int size, int32_t* out) {
    if (!n || !out || count < 0 || size < 0) return CB_EINVAL;
    Mt19937 rng(seed);
    std::vector<int32_t> perm;
    for (int c = 0; c < count; ++c) {
        if (n[c] < size) return CB_EINVAL;  // "Cannot take a larger sample than population when replace=False"
        perm.resize(n[c]);
        for (int i = 0; i < n[c]; ++i) perm[i] = i;
        for (int i = n[c] - 1; i >= 1; --i) {
            const uint32_t j = rng.interval((uint32_t)i);
            const int32_t t = perm[i];
            perm[i] = perm[j];
            perm[j] = t;
        }
        for (int i = 0; i < size; ++i) out[(size_t)c * size + i] = perm[i];
    }
    return CB_OK;
}
