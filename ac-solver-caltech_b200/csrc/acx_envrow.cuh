// acx_envrow.cuh -- one ACEnv.step move on a CERTIFIED row (see acx_rows.cuh: letters in {0,+-1,+-2},
// zero padded, both relators freely and cyclically reduced, lengths known), split by move class:
//
//   conjugation (move ids 4..11, two thirds of uniform moves; ac_moves.py:79-156 + utils.py:223-232):
//     g r g^-1 of a cyclically reduced r is the left rotation (r[0] == g^-1), the right rotation
//     (r[-1] == g) or r itself, so it is done ON THE BYTES: two byte tests and a one-byte rotation of the
//     9 row words with PRMT -- no bit planes, about 50 instructions instead of 240.
//   concatenation (ids 0..3; ac_moves.py:4-76): needs the junction cancellation depth and the cyclic
//     reduction depth, which are one ctz each on bit planes: pack two planes per relator, splice, reduce,
//     unpack the touched relator (acx_rel64.cuh).
//
// The env kernel (acx_envtile.cuh) sorts the rows of a tile by class so that a warp runs ONE of the two.
// Results are identical to acmove_row64_certified / acmove_row64 (tests/test_host_rows.py, host harness).
#pragma once
#include "acx_rows.cuh"

namespace acx {

ACX_HD uint32_t* words32(uint8_t* p) { return static_cast<uint32_t*>(__builtin_assume_aligned(p, 4)); }
// keeps the compiler from re-deriving the (selected) base per load: ten loads off ONE register
ACX_HD const uint32_t* one_base(const uint32_t* p) {
#ifdef __CUDA_ARCH__
    asm("" : "+l"(p));
#endif
    return p;
}

// rel: 4-byte aligned relator of MS slots with n >= 1 letters; readable from rel - 4 to rel + MS + 3.
// d = decode_move() bits (bit2: g is an inverse letter, bit3: g is y).  Returns true when the word changed.
template <int MS>
ACX_HD bool cert_conj_bytes(uint8_t* rel, int n, uint32_t d) {
    constexpr int NW = MS / 4;
    const uint32_t mag = 1u + ((d >> 3) & 1u);
    const uint32_t neg = 0x100u - mag;
    const uint32_t g = (d & 4u) ? neg : mag, gi = (d & 4u) ? mag : neg;
    const uint32_t first = rel[0], last = rel[n - 1];
    const bool sc = first == gi;  // first letter == g^-1: r[1:] + [r[0]]   (ac_moves.py:119)
    const bool ec = last == g;    // last letter == g:      [r[-1]] + r[:-1] (ac_moves.py:120)
    if (!(sc || ec) || n <= 1) return false;
    // word k of the result = bytes 4k+1 .. 4k+4 (left) or 4k-1 .. 4k+2 (right) of the old word
    const uint32_t* p = one_base(words32(rel) - (sc ? 0 : 1));
    uint32_t u[NW + 1];
#pragma unroll
    for (int k = 0; k <= NW; ++k) u[k] = p[k];
    if (sc) u[NW] = 0u;
    else u[0] = last << 24;
    const uint32_t sel = sc ? 0x4321u : 0x6543u;
    uint32_t* q = words32(rel);
#pragma unroll
    for (int k = 0; k < NW; ++k) q[k] = prmt(u[k], u[k + 1], sel);
    if (sc) rel[n - 1] = (uint8_t)first;
    else if (n < MS) rel[n] = 0;
    return true;
}

// The conjugation of relator i of a row HELD IN REGISTERS (w: the 2 * MS / 4 row words), for the frontier expansion
// where one parent gives eight conjugation children: no shared-memory loads at all.  g r g^-1 of a freely reduced r is
// r shifted one byte to the left (sc: r[0] == g^-1 cancels g) or to the right (g goes in front), with the last letter
// dropped (ec: r[-1] == g) or g^-1 appended (ac_moves.py:79-156).  CERT: r is also cyclically reduced and the result
// is reduced cyclically (utils.py:223-232), which leaves a rotation or r itself.  The shifted row comes back in o
// except for ONE byte, o_bytes[fix_pos] = fix_val (fix_pos < 0: none), which the caller stores after the words.
// Returns the new length of the relator; the word is unchanged when the result would not fit (ac_moves.py:126).
template <int MS, bool CERT>
ACX_HD int conj_row_words(const uint32_t (&w)[2 * (MS / 4)], uint32_t (&o)[2 * (MS / 4)], bool i, int n, bool sc, bool ec,
                          uint32_t d, int& fix_pos, uint32_t& fix_val) {
    constexpr int NW = MS / 4;
    const uint32_t mag = 1u + ((d >> 3) & 1u);
    const uint32_t neg = 0x100u - mag;
    const uint32_t g = (d & 4u) ? neg : mag, gi = (d & 4u) ? mag : neg;
    const int nn = CERT ? n : n + 2 - 2 * ((int)sc + (int)ec);
    const bool move = CERT ? (sc || ec) : (nn <= MS);
    const bool left = move && sc;
    const uint32_t lead = g << 24;
    // v[k] = word k (left) or word k - 1 (right, and when nothing moves) of the row, g in front, zero behind
    uint32_t v[2 * NW + 1];
    v[0] = left ? w[0] : lead;
#pragma unroll
    for (int k = 1; k < 2 * NW; ++k) v[k] = left ? w[k] : w[k - 1];
    v[2 * NW] = left ? 0u : w[2 * NW - 1];
    const uint32_t keep = left ? 0x3210u : 0x7654u, shift = left ? 0x4321u : 0x6543u;
    const uint32_t sel0 = (move && !i) ? shift : keep, sel1 = (move && i) ? shift : keep;
    const uint32_t a_mid = (move && i && !left) ? lead : v[NW];  // relator 1 shifted right: g, not the end of relator 0
    const uint32_t b_mid = (left && !i) ? 0u : v[NW];            // relator 0 shifted left: zero, not the start of relator 1
#pragma unroll
    for (int k = 0; k < NW; ++k) o[k] = prmt(v[k], k == NW - 1 ? b_mid : v[k + 1], sel0);
#pragma unroll
    for (int k = NW; k < 2 * NW; ++k) o[k] = prmt(k == NW ? a_mid : v[k], v[k + 1], sel1);
    // left: r[1:-1] (the old last letter now sits at n - 2) or r[1:] + [g^-1]; right: [g] + r + [g^-1] or [g] + r[:-1]
    const int pos = n - 1 + (left ? (ec ? -1 : 0) : (ec ? 1 : 2));
    fix_val = ec ? 0u : gi;
    fix_pos = (move && pos < MS) ? pos + (i ? MS : 0) : -1;
    return move ? nn : n;
}

template <int MS>
ACX_HD void unpack_words64(uint32_t (&u)[MS / 4], const R64& r) {
    const uint64_t N = mask64(r.n), X = r.S ^ r.T;
    const uint32_t n2[2] = {(uint32_t)N, (uint32_t)(N >> 32)};
    const uint32_t s2[2] = {(uint32_t)r.S, (uint32_t)(r.S >> 32)};
    const uint32_t x2[2] = {(uint32_t)X, (uint32_t)(X >> 32)};
#pragma unroll
    for (int k = 0; k < MS / 4; ++k) {
        const int q = k / 8, sh = 4 * (k % 8);
        u[k] = spread4((n2[q] >> sh) & 0xFu) + spread4((x2[q] >> sh) & 0xFu) + 253u * spread4((s2[q] >> sh) & 0xFu);
    }
}

// row: 8-byte aligned.  d = decode_move() bits of a concatenation (bit0: relator touched, bit2: use r_j^-1).
template <int MS>
ACX_HD uint32_t cert_concat_planes(uint8_t* row, uint32_t d, int& l0, int& l1) {
    constexpr int NW = MS / 4;
    uint32_t w[2 * NW];
    const uint2* p = reinterpret_cast<const uint2*>(row);
#pragma unroll
    for (int k = 0; k < NW; ++k) {
        const uint2 v = p[k];
        w[2 * k] = v.x;
        w[2 * k + 1] = v.y;
    }
    R64 r[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        uint32_t b0[2] = {0u, 0u}, s7[2] = {0u, 0u};
#pragma unroll
        for (int k = NW - 1; k >= 0; --k) pack_word2(w[h * NW + k], b0[k / 8], s7[k / 8]);
        const int n = h ? l1 : l0;
        const uint64_t B0 = ((uint64_t)b0[1] << 32) | b0[0];
        r[h].S = ((uint64_t)s7[1] << 32) | s7[0];
        r[h].T = ~B0 & mask64(n);  // x / x^-1 have bit 0 set, y / y^-1 do not
        r[h].n = n;
    }
    const bool i = d & 1u;
    R64 a, b;
    a.S = i ? r[1].S : r[0].S;
    a.T = i ? r[1].T : r[0].T;
    a.n = i ? r[1].n : r[0].n;
    b.S = i ? r[0].S : r[1].S;
    b.T = i ? r[0].T : r[1].T;
    b.n = i ? r[0].n : r[1].n;
    if (d & 4u) b = r64_inverse(b);
    if (!r64_concat(a, b, MS)) return ACX_ROW_OK;      // guard failed: row unchanged (ac_moves.py:64)
    if (a.n == 0) return ACX_ROW_EMPTY_RELATOR;        // utils.py:264-266 raises; row stays as it was
    r64_cyclic_reduce(a);
    unpack_rel_static64<MS>(row + (d & 1u) * MS, a);
    if (i) l1 = a.n;
    else l0 = a.n;
    return ACX_ROW_OK;
}

template <int MS>
ACX_HD uint32_t cert_row_step(uint8_t* row, int move, int& l0, int& l1) {
    const uint32_t d = decode_move(move);
    if (d & 2u) {  // the lengths never change
        cert_conj_bytes<MS>(row + (d & 1u) * MS, (d & 1u) ? l1 : l0, d);
        return ACX_ROW_OK;
    }
    return cert_concat_planes<MS>(row, d, l0, l1);
}

}  // namespace acx
