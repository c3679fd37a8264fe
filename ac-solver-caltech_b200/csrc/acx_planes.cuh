// acx_planes.cuh -- bit-plane word arithmetic for Andrews-Curtis moves (sm_100a).
//
// A relator of letters in {x, x^-1, y, y^-1} = {1,-1,2,-2} is held by ONE thread as
// two bit planes of 32*W bits (letter k = bit k) plus its length:
//     S (sign plane)  bit k = 1  <=>  letter k is an inverse (-1 or -2)
//     T (type plane)  bit k = 1  <=>  letter k is y or y^-1
// With that layout every step of the reference's ACMove is O(W) ALU work:
//   * r^-1 (reverse + negate, ac_moves.py:41-48)  = bit-reverse, shift, xor with mask
//   * junction cancellation depth (ac_moves.py:56-60) = ctz of a plane difference
//   * splice / conjugate / cyclic reduction (utils.py:223-232) = shifts and masks
// so the branchy per-letter loops of the reference never appear on the fast path.
//
// Rows whose letters are outside {0,+-1,+-2}, or which are not freely reduced on
// input, take a byte-wise slow path (acx_slow.cuh) that follows the reference
// literally (stack free reduction, utils.py:211-220).
#pragma once
#include <stdint.h>

// The row arithmetic is __host__ __device__ so that tests/ can run the very same code on the
// CPU against the oracle (tests/host_harness); the product only ever launches it on the GPU.
#define ACX_HD __host__ __device__ __forceinline__

namespace acx {

#ifdef __CUDA_ARCH__
ACX_HD uint32_t fshl(uint32_t lo, uint32_t hi, int s) { return __funnelshift_l(lo, hi, s); }
ACX_HD uint32_t fshr(uint32_t lo, uint32_t hi, int s) { return __funnelshift_r(lo, hi, s); }
ACX_HD uint32_t fshl_c(uint32_t lo, uint32_t hi, int s) { return __funnelshift_lc(lo, hi, s); }
ACX_HD uint32_t fshr_c(uint32_t lo, uint32_t hi, int s) { return __funnelshift_rc(lo, hi, s); }
ACX_HD uint32_t brev32(uint32_t x) { return __brev(x); }
ACX_HD int clz32(uint32_t x) { return __clz(x); }
ACX_HD int popc32(uint32_t x) { return __popc(x); }
ACX_HD uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) { return __byte_perm(a, b, sel); }
#else
// byte permute: result byte i = byte (selector nibble i & 7) of the 8-byte pool {b, a} (no sign replication used)
ACX_HD uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {
    const uint64_t pool = ((uint64_t)b << 32) | a;
    uint32_t r = 0;
    for (int i = 0; i < 4; ++i) r |= (uint32_t)((pool >> (8 * ((sel >> (4 * i)) & 7u))) & 0xFFu) << (8 * i);
    return r;
}
ACX_HD uint32_t fshl(uint32_t lo, uint32_t hi, int s) {
    s &= 31;
    return (uint32_t)(((((uint64_t)hi) << 32 | lo) << s) >> 32);
}
ACX_HD uint32_t fshr(uint32_t lo, uint32_t hi, int s) {
    s &= 31;
    return (uint32_t)((((uint64_t)hi) << 32 | lo) >> s);
}
ACX_HD uint32_t fshl_c(uint32_t lo, uint32_t hi, int s) {
    if (s >= 32) return lo;
    return (uint32_t)(((((uint64_t)hi) << 32 | lo) << s) >> 32);
}
ACX_HD uint32_t fshr_c(uint32_t lo, uint32_t hi, int s) {
    if (s >= 32) return hi;
    return (uint32_t)((((uint64_t)hi) << 32 | lo) >> s);
}
ACX_HD uint32_t brev32(uint32_t x) {
    uint32_t r = 0;
    for (int i = 0; i < 32; ++i) r |= ((x >> i) & 1u) << (31 - i);
    return r;
}
ACX_HD int clz32(uint32_t x) { return x ? __builtin_clz(x) : 32; }
ACX_HD int popc32(uint32_t x) { return __builtin_popcount(x); }
#endif

template <int W>
struct BV {
    uint32_t w[W];
};

template <int W>
ACX_HD BV<W> bv_zero() {
    BV<W> r;
#pragma unroll
    for (int i = 0; i < W; ++i) r.w[i] = 0u;
    return r;
}

#define ACX_BV_BINOP(name, expr)                                            \
    template <int W>                                                        \
    ACX_HD BV<W> name(const BV<W>& a, const BV<W>& b) { \
        BV<W> r;                                                            \
        _Pragma("unroll") for (int i = 0; i < W; ++i) {                     \
            const uint32_t x = a.w[i], y = b.w[i];                          \
            r.w[i] = (expr);                                                \
        }                                                                   \
        return r;                                                           \
    }
ACX_BV_BINOP(bv_and, x& y)
ACX_BV_BINOP(bv_or, x | y)
ACX_BV_BINOP(bv_xor, x ^ y)
ACX_BV_BINOP(bv_andn, x & ~y)
#undef ACX_BV_BINOP

template <int W>
ACX_HD bool bv_any(const BV<W>& a) {
    uint32_t o = 0;
#pragma unroll
    for (int i = 0; i < W; ++i) o |= a.w[i];
    return o != 0u;
}

template <int W>
ACX_HD bool bv_eq(const BV<W>& a, const BV<W>& b) {
    uint32_t o = 0;
#pragma unroll
    for (int i = 0; i < W; ++i) o |= a.w[i] ^ b.w[i];
    return o == 0u;
}

template <int W>
ACX_HD int bv_popc(const BV<W>& a) {
    int n = 0;
#pragma unroll
    for (int i = 0; i < W; ++i) n += popc32(a.w[i]);
    return n;
}

// index of the lowest set bit; 32*W when a == 0
template <int W>
ACX_HD int bv_ctz(const BV<W>& a) {
    int res = 32 * W;
#pragma unroll
    for (int i = W - 1; i >= 0; --i) {
        const int c = clz32(brev32(a.w[i]));  // 32 when the word is zero
        res = (c < 32) ? (32 * i + c) : res;
    }
    return res;
}

// logical right shift by 0 <= s <= 32*W
template <int W>
ACX_HD BV<W> bv_shr(const BV<W>& a, int s) {
    BV<W> t;
    if (W == 1) {
        t.w[0] = fshr_c(a.w[0], 0u, s);
        return t;
    }
    const int r = s & 31, q = s >> 5;
#pragma unroll
    for (int i = 0; i < W; ++i) t.w[i] = fshr(a.w[i], (i + 1 < W) ? a.w[i + 1] : 0u, r);
#pragma unroll
    for (int st = 1; st <= W; st <<= 1) {
        const bool on = (q & st) != 0;
#pragma unroll
        for (int i = 0; i < W; ++i) {
            const uint32_t moved = (i + st < W) ? t.w[(i + st < W) ? i + st : 0] : 0u;
            t.w[i] = on ? moved : t.w[i];
        }
    }
    return t;
}

// logical left shift by 0 <= s <= 32*W
template <int W>
ACX_HD BV<W> bv_shl(const BV<W>& a, int s) {
    BV<W> t;
    if (W == 1) {
        t.w[0] = fshl_c(0u, a.w[0], s);
        return t;
    }
    const int r = s & 31, q = s >> 5;
#pragma unroll
    for (int i = W - 1; i >= 0; --i) t.w[i] = fshl((i > 0) ? a.w[i - 1] : 0u, a.w[i], r);
#pragma unroll
    for (int st = 1; st <= W; st <<= 1) {
        const bool on = (q & st) != 0;
#pragma unroll
        for (int i = W - 1; i >= 0; --i) {
            const uint32_t moved = (i - st >= 0) ? t.w[(i - st >= 0) ? i - st : 0] : 0u;
            t.w[i] = on ? moved : t.w[i];
        }
    }
    return t;
}

// static shift right by one bit
template <int W>
ACX_HD BV<W> bv_shr1(const BV<W>& a) {
    BV<W> t;
#pragma unroll
    for (int i = 0; i < W; ++i) t.w[i] = fshr(a.w[i], (i + 1 < W) ? a.w[i + 1] : 0u, 1);
    return t;
}

// low n bits set, 0 <= n <= 32*W
template <int W>
ACX_HD BV<W> bv_mask(int n) {
    BV<W> r;
#pragma unroll
    for (int i = 0; i < W; ++i) {
        const int k = n - 32 * i;  // bits wanted in this word
        r.w[i] = (k <= 0) ? 0u : (0xffffffffu >> (32 - min(k, 32)));
    }
    return r;
}

// bit reversal over the whole 32*W-bit vector
template <int W>
ACX_HD BV<W> bv_rev(const BV<W>& a) {
    BV<W> r;
#pragma unroll
    for (int i = 0; i < W; ++i) r.w[i] = brev32(a.w[W - 1 - i]);
    return r;
}

template <int W>
ACX_HD uint32_t bv_bit(const BV<W>& a, int k) {
    return bv_shr<W>(a, k).w[0] & 1u;
}

// ---------------------------------------------------------------------------
// relator = (S, T, n); bits at positions >= n are zero in both planes
template <int W>
struct Rel {
    BV<W> S, T;
    int n;
};

// r^-1: reversed and negated word (ac_moves.py:41-48)
template <int W>
ACX_HD Rel<W> rel_inverse(const Rel<W>& r) {
    Rel<W> o;
    const int sh = 32 * W - r.n;
    o.S = bv_xor<W>(bv_shr<W>(bv_rev<W>(r.S), sh), bv_mask<W>(r.n));
    o.T = bv_shr<W>(bv_rev<W>(r.T), sh);
    o.n = r.n;
    return o;
}

// number of leading letters on which two words agree, capped at `cap`
template <int W>
ACX_HD int rel_common_prefix(const Rel<W>& a, const Rel<W>& b, int cap) {
    const BV<W> d = bv_or<W>(bv_xor<W>(a.S, b.S), bv_xor<W>(a.T, b.T));
    return min(bv_ctz<W>(d), cap);
}

// true when some adjacent pair of letters cancels (word is not freely reduced)
template <int W>
ACX_HD bool rel_has_adjacent_inverse(const Rel<W>& r) {
    const BV<W> ds = bv_xor<W>(r.S, bv_shr1<W>(r.S));  // sign differs from the next letter
    const BV<W> dt = bv_xor<W>(r.T, bv_shr1<W>(r.T));  // type differs from the next letter
    const BV<W> bad = bv_and<W>(bv_andn<W>(ds, dt), bv_mask<W>(max(r.n - 1, 0)));
    return bv_any<W>(bad);
}

// concatenate_relators (ac_moves.py:4-76): a <- a * b with junction-only cancellation
// (b already holds r_j or r_j^-1).  Returns false when the guard new_size <= m fails
// (row unchanged, ac_moves.py:64).  new length 0 is returned as a.n == 0 (caller reports it).
template <int W>
ACX_HD bool rel_concat(Rel<W>& a, const Rel<W>& b, int m) {
    const Rel<W> ia = rel_inverse<W>(a);
    const int acc = rel_common_prefix<W>(ia, b, min(a.n, b.n));  // ac_moves.py:56-60
    const int nn = a.n + b.n - 2 * acc;                          // ac_moves.py:62
    if (nn > m) return false;
    const int keep = a.n - acc;
    const BV<W> km = bv_mask<W>(keep);
    a.S = bv_or<W>(bv_and<W>(a.S, km), bv_shl<W>(bv_shr<W>(b.S, acc), keep));
    a.T = bv_or<W>(bv_and<W>(a.T, km), bv_shl<W>(bv_shr<W>(b.T, acc), keep));
    a.n = nn;
    return true;
}

// conjugate (ac_moves.py:79-156): a <- g a g^-1 with g = (gs ? inverse : plain) of (gt ? y : x).
// One letter is tested at each end (:119-120).  Requires a.n >= 1.
template <int W>
ACX_HD bool rel_conjugate(Rel<W>& a, uint32_t gs, uint32_t gt, int m) {
    const uint32_t s0 = a.S.w[0] & 1u, t0 = a.T.w[0] & 1u;
    const uint32_t sl = bv_bit<W>(a.S, a.n - 1), tl = bv_bit<W>(a.T, a.n - 1);
    const int sc = (t0 == gt) && (s0 != gs);  // first letter == g^-1
    const int ec = (tl == gt) && (sl == gs);  // last letter == g
    const int nn = a.n + 2 - 2 * (sc + ec);   // ac_moves.py:123
    if (nn > m) return false;
    const int core = a.n - sc - ec;
    const BV<W> cm = bv_mask<W>(core);
    BV<W> S = bv_and<W>(bv_shr<W>(a.S, sc), cm);
    BV<W> T = bv_and<W>(bv_shr<W>(a.T, sc), cm);
    int n = core;
    if (!sc) {  // prepend g
        S = bv_shl<W>(S, 1);
        T = bv_shl<W>(T, 1);
        S.w[0] |= gs;
        T.w[0] |= gt;
        n += 1;
    }
    if (!ec) {  // append g^-1
        BV<W> one = bv_zero<W>();
        one.w[0] = 1u;
        const BV<W> at = bv_shl<W>(one, n);
        if (!gs) S = bv_or<W>(S, at);
        if (gt) T = bv_or<W>(T, at);
        n += 1;
    }
    a.S = S;
    a.T = T;
    a.n = n;
    return true;
}

// cyclic reduction (utils.py:223-232): drop k letters from both ends while r[p] == -r[n-1-p]
template <int W>
ACX_HD void rel_cyclic_reduce(Rel<W>& r) {
    const Rel<W> ir = rel_inverse<W>(r);
    const int k = rel_common_prefix<W>(ir, r, r.n >> 1);
    const int nn = r.n - 2 * k;
    const BV<W> mk = bv_mask<W>(nn);
    r.S = bv_and<W>(bv_shr<W>(r.S, k), mk);
    r.T = bv_and<W>(bv_shr<W>(r.T, k), mk);
    r.n = nn;
}

// 12-entry decode of move_id (ac_moves.py:192-206):
//   ids 0..3  concatenation  r_i <- r_i r_j^{sign}
//   ids 4..11 conjugation    r_i <- g r_i g^-1
// i = (id+1) & 1.  Encoded as: bit0 = i, bit1 = is_conj, bit2 = inverse flag
// (concat: use r_j^-1; conj: g is an inverse letter), bit3 = g is y.
ACX_HD uint32_t decode_move(int id) {
    // id : kind i sign/g     -> bits
    // 0  : cat  1 +             1
    // 1  : cat  0 -             4
    // 2  : cat  1 -             5
    // 3  : cat  0 +             0
    // 4  : conj 1 g=-1 (x^-1)   1|2|4
    // 5  : conj 0 g=-2 (y^-1)   2|4|8
    // 6  : conj 1 g=-2          1|2|4|8
    // 7  : conj 0 g=+1          2
    // 8  : conj 1 g=+1          1|2
    // 9  : conj 0 g=+2          2|8
    // 10 : conj 1 g=+2          1|2|8
    // 11 : conj 0 g=-1          2|4
    const uint64_t table = 0x1ull | (0x4ull << 4) | (0x5ull << 8) | (0x0ull << 12) | (0x7ull << 16) |
                           (0xEull << 20) | (0xFull << 24) | (0x2ull << 28) | (0x3ull << 32) |
                           (0xAull << 36) | (0xBull << 40) | (0x6ull << 44);
    return (uint32_t)(table >> (4 * id)) & 0xFu;
}

}  // namespace acx
