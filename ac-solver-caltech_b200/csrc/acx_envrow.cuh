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
    const uint32_t* p = reinterpret_cast<const uint32_t*>(rel) - (sc ? 0 : 1);
    uint32_t u[NW + 1];
#pragma unroll
    for (int k = 0; k <= NW; ++k) u[k] = p[k];
    if (sc) u[NW] = 0u;
    else u[0] = last << 24;
    const uint32_t sel = sc ? 0x4321u : 0x6543u;
    uint32_t* q = reinterpret_cast<uint32_t*>(rel);
#pragma unroll
    for (int k = 0; k < NW; ++k) q[k] = prmt(u[k], u[k + 1], sel);
    if (sc) rel[n - 1] = (uint8_t)first;
    else if (n < MS) rel[n] = 0;
    return true;
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
