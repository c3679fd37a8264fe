// acx_rel64.cuh -- the plane arithmetic of acx_planes.cuh specialised to relators of at most 64
// letters: each plane is ONE 64-bit register pair, so a variable shift is two SHF instructions
// and a prefix mask is N itself (the non-zero plane of a valid relator is always 2^n - 1).
// This is the path every m <= 64 takes (BASELINE configs 1-4 use m <= 36).
#pragma once
#include "acx_planes.cuh"

namespace acx {

#ifdef __CUDA_ARCH__
ACX_HD uint64_t shr64(uint64_t x, int s) {  // 0 <= s <= 64 (PTX shifts clamp, C++ shifts do not)
    uint64_t r;
    asm("shr.b64 %0, %1, %2;" : "=l"(r) : "l"(x), "r"(s));
    return r;
}
ACX_HD uint64_t shl64(uint64_t x, int s) {
    uint64_t r;
    asm("shl.b64 %0, %1, %2;" : "=l"(r) : "l"(x), "r"(s));
    return r;
}
ACX_HD uint64_t brev64(uint64_t x) { return __brevll(x); }
ACX_HD int ctz64(uint64_t x) { return __clzll(__brevll(x)); }  // 64 when x == 0
ACX_HD int popc64(uint64_t x) { return __popcll(x); }
#else
ACX_HD uint64_t shr64(uint64_t x, int s) { return s >= 64 ? 0ull : (x >> s); }
ACX_HD uint64_t shl64(uint64_t x, int s) { return s >= 64 ? 0ull : (x << s); }
ACX_HD uint64_t brev64(uint64_t x) {
    uint64_t r = 0;
    for (int i = 0; i < 64; ++i) r |= ((x >> i) & 1ull) << (63 - i);
    return r;
}
ACX_HD int ctz64(uint64_t x) { return x ? __builtin_ctzll(x) : 64; }
ACX_HD int popc64(uint64_t x) { return __builtin_popcountll(x); }
#endif

ACX_HD uint64_t mask64(int n) { return ~shl64(~0ull, n); }  // low n bits, 0 <= n <= 64

struct R64 {
    uint64_t S, T;  // sign / type planes, zero at positions >= n
    int n;
};

// r^-1 (ac_moves.py:41-48)
ACX_HD R64 r64_inverse(const R64& r) {
    R64 o;
    const int sh = 64 - r.n;
    o.S = shr64(brev64(r.S), sh) ^ mask64(r.n);
    o.T = shr64(brev64(r.T), sh);
    o.n = r.n;
    return o;
}

ACX_HD int r64_common_prefix(const R64& a, const R64& b, int cap) {
    return min(ctz64((a.S ^ b.S) | (a.T ^ b.T)), cap);
}

// concatenate_relators (ac_moves.py:4-76); b is r_j or r_j^-1.  false = guard failed, a untouched.
ACX_HD bool r64_concat(R64& a, const R64& b, int m) {
    const R64 ia = r64_inverse(a);
    const int acc = r64_common_prefix(ia, b, min(a.n, b.n));  // :56-60
    const int nn = a.n + b.n - 2 * acc;                       // :62
    if (nn > m) return false;                                 // :64
    const int keep = a.n - acc;
    const uint64_t km = mask64(keep);
    a.S = (a.S & km) | shl64(shr64(b.S, acc), keep);
    a.T = (a.T & km) | shl64(shr64(b.T, acc), keep);
    a.n = nn;
    return true;
}

// conjugate (ac_moves.py:79-156): a <- g a g^-1, g sign gs (1 = inverse letter), type gt (1 = y)
ACX_HD bool r64_conjugate(R64& a, uint32_t gs, uint32_t gt, int m) {
    const uint32_t s0 = (uint32_t)a.S & 1u, t0 = (uint32_t)a.T & 1u;
    const uint32_t sl = (uint32_t)shr64(a.S, a.n - 1) & 1u, tl = (uint32_t)shr64(a.T, a.n - 1) & 1u;
    const int sc = (t0 == gt) && (s0 != gs);  // first letter == g^-1 (:119)
    const int ec = (tl == gt) && (sl == gs);  // last letter == g     (:120)
    const int nn = a.n + 2 - 2 * (sc + ec);   // :123
    if (nn > m) return false;                 // :126
    const int core = a.n - sc - ec;
    const uint64_t cm = mask64(core);
    // core letters moved to position 1 - sc... i.e. shifted right by sc, then left by (1 - sc)
    uint64_t S = shr64(a.S, sc) & cm, T = shr64(a.T, sc) & cm;
    const int lead = 1 - sc;  // room for g in front
    S = shl64(S, lead) | (uint64_t)(lead ? gs : 0u);
    T = shl64(T, lead) | (uint64_t)(lead ? gt : 0u);
    const int pos = core + lead;
    if (!ec) {  // append g^-1
        S |= shl64((uint64_t)(gs ^ 1u), pos);
        T |= shl64((uint64_t)gt, pos);
    }
    a.S = S;
    a.T = T;
    a.n = nn;
    return true;
}

// cyclic reduction (utils.py:223-232); returns the depth k
ACX_HD int r64_cyclic_reduce(R64& r) {
    const R64 ir = r64_inverse(r);
    const int k = r64_common_prefix(ir, r, r.n >> 1);
    const int nn = r.n - 2 * k;
    const uint64_t mk = mask64(nn);
    r.S = shr64(r.S, k) & mk;
    r.T = shr64(r.T, k) & mk;
    r.n = nn;
    return k;
}

// does cyclic reduction have anything to do?  (first letter is the inverse of the last one)
ACX_HD bool r64_needs_cyclic(const R64& r) {
    const uint32_t s0 = (uint32_t)r.S & 1u, t0 = (uint32_t)r.T & 1u;
    const uint32_t sl = (uint32_t)shr64(r.S, r.n - 1) & 1u, tl = (uint32_t)shr64(r.T, r.n - 1) & 1u;
    return (r.n >= 2) && (t0 == tl) && (s0 != sl);
}

struct Raw64 {
    uint64_t B0, B1, S7;
};

// classification of a packed row; fills r[] and the non-zero counts
ACX_HD int classify_row64(const Raw64 (&raw)[2], uint32_t chk, R64 (&r)[2], int (&cnt)[2]) {
    uint64_t bad_alpha = 0, bad_red = 0;
    bool valid = true;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const uint64_t B0 = raw[h].B0, B1 = raw[h].B1, S = raw[h].S7;
        bad_alpha |= (~S & B1 & B0) | (S & ~B1);  // bytes 0x03 / 0xFC, 0xFD
        const uint64_t N = B0 | B1;
        const int n = popc64(N);
        cnt[h] = n;
        valid = valid && (N != 0ull) && ((N & (N + 1ull)) == 0ull);  // non-empty prefix mask
        const uint64_t T = B1 & ~B0;
        r[h].S = S;
        r[h].T = T;
        r[h].n = n;
        bad_red |= (S ^ (S >> 1)) & ~(T ^ (T >> 1)) & (N >> 1);  // adjacent inverse pair
    }
    if ((chk & 0x7C7C7C7Cu) != 0u || bad_alpha != 0ull) return ROW_SLOW;
    if (!valid) return ROW_INVALID;
    return bad_red == 0ull ? ROW_FAST : ROW_SLOW;
}

}  // namespace acx
