// acx_ms.cuh -- enumeration step of generate_miller_schupp_presentations
// (ac_solver/search/miller_schupp/miller_schupp.py:20-83): one element per word w of
// itertools.product([1, 2, -1, -2], repeat=k) (:53).  The functor applies the exponent-sum filter (:56-57),
// builds x^-1 w, reduces it freely and cyclically (simplify_relator(cyclical=True), utils.py:178-243), drops
// x^-1 itself (:66-67) and emits the reduced relator together with the key of its class under cyclic
// permutation (:69-75) -- the minimum rotation.  Words are 2 bits per letter (code 0..3 = 1, 2, -1, -2, the
// inverse of code c is c ^ 2), k <= 15 so that x^-1 w fits 32 bits.  __host__ __device__ like the other row
// code: tests/host_harness runs the same functor on the CPU.
#pragma once
#include <stdint.h>

#include "acx_planes.cuh"

namespace acx {

static constexpr uint64_t MS_INVALID = ~0ull;

struct MsEnumerate {
    int k;           // len(w)
    int64_t first;   // index of the first word of this launch in product order (last letter varies fastest)
    uint64_t* key;   // [count] class key: min rotation | length << 32, MS_INVALID = filtered out
    uint64_t* rel;   // [count] reduced x^-1 w: letters | length << 32
    ACX_HD void operator()(int64_t j) const {
        const uint64_t idx = (uint64_t)(first + j);
        uint64_t s = 2ull;  // x^-1
        int sp = 1, ex = 0;
        for (int t = 0; t < k; ++t) {
            const uint64_t c = (idx >> (2 * (k - 1 - t))) & 3ull;
            ex += (c == 0ull) - (c == 2ull);
            if (sp > 0 && ((s >> (2 * (sp - 1))) & 3ull) == (c ^ 2ull)) {  // utils.py:211-220: the reduced word is unique
                --sp;
                s &= ~(3ull << (2 * sp));
            } else {
                s |= c << (2 * sp);
                ++sp;
            }
        }
        key[j] = MS_INVALID;
        rel[j] = 0ull;
        if (ex != 0 || sp == 0) return;  // :56-57
        int lo = 0, hi = sp - 1;
        while (hi > lo && ((s >> (2 * lo)) & 3ull) == (((s >> (2 * hi)) & 3ull) ^ 2ull)) {  // utils.py:223-232
            ++lo;
            --hi;
        }
        const int len = hi - lo + 1;
        const uint64_t mask = (1ull << (2 * len)) - 1ull;
        const uint64_t word = (s >> (2 * lo)) & mask;
        if (len == 1 && word == 2ull) return;  // :66-67
        uint64_t best = word;
        for (int i = 1; i < len; ++i) {
            const uint64_t rot = ((word >> (2 * i)) | (word << (2 * (len - i)))) & mask;
            best = rot < best ? rot : best;
        }
        key[j] = best | ((uint64_t)len << 32);
        rel[j] = word | ((uint64_t)len << 32);
    }
};

}  // namespace acx
