// acx_rows.cuh -- one presentation row (int8 [r_0 | r_1], 2*m bytes, staged in shared
// memory) <-> bit planes, and the row-level ACMove built on acx_planes.cuh.
//
// pack:   4 letters per 32-bit word -> one nibble of each plane with ONE multiply
//         (a multiply-gather: the wanted bit of each byte lands in bits 28..31, the
//         cross terms land below bit 28 or above bit 31 and never collide), inserted
//         into the plane with a funnel shift.  3 instructions per plane per word.
// unpack: one nibble of a plane -> 4 bytes with one multiply (bit j -> byte j), and
//         byte = n + 253*s + (s^t) (mod 256) turns the planes back into {1,2,-1,-2,0}
//         without any per-byte branch; sums never carry between bytes (max 255).
#pragma once
#include "acx_planes.cuh"

namespace acx {

enum : uint32_t {
    ACX_ROW_OK = 0,
    ACX_ROW_EMPTY_RELATOR = 1,  // the reference raises AssertionError (utils.py:264-266)
    ACX_ROW_INVALID_INPUT = 2   // not a valid presentation on input (utils.py:13-54)
};

// Result of classifying a packed row
enum : int { ROW_FAST = 0, ROW_SLOW = 1, ROW_INVALID = 2 };

}  // namespace acx
#include "acx_rel64.cuh"
namespace acx {

// raw planes of one relator as read from memory
template <int W>
struct RawRel {
    BV<W> B0, B1, S7;  // bit 0, bit 1 and bit 7 of every letter byte
};

// insert the top nibble of x below the bits already in plane word p
ACX_HD uint32_t nib_push(uint32_t x, uint32_t p) { return fshl(x, p, 4); }

ACX_HD void pack_word(uint32_t x, uint32_t& b0, uint32_t& b1, uint32_t& s7, uint32_t& chk) {
    chk |= x ^ (x >> 1);
    b0 = nib_push((x & 0x01010101u) * 0x10204080u, b0);
    b1 = nib_push((x & 0x02020202u) * 0x08102040u, b1);
    s7 = nib_push((x & 0x80808080u) * 0x00204081u, s7);
}

// ---- static width (MS % 4 == 0): row words come from 8-byte aligned shared memory -------
template <int W, int MS>
ACX_HD void pack_row_static(const uint8_t* row, RawRel<W> (&raw)[2], uint32_t& chk) {
    constexpr int NW = MS / 4;  // words per relator
    uint32_t w[2 * NW];
    const uint2* p = reinterpret_cast<const uint2*>(row);
#pragma unroll
    for (int k = 0; k < NW; ++k) {
        const uint2 v = p[k];
        w[2 * k] = v.x;
        w[2 * k + 1] = v.y;
    }
    chk = 0u;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        raw[h].B0 = bv_zero<W>();
        raw[h].B1 = bv_zero<W>();
        raw[h].S7 = bv_zero<W>();
#pragma unroll
        for (int k = NW - 1; k >= 0; --k)
            pack_word(w[h * NW + k], raw[h].B0.w[k / 8], raw[h].B1.w[k / 8], raw[h].S7.w[k / 8], chk);
    }
}

ACX_HD uint32_t spread4(uint32_t nib) { return (nib * 0x00204081u) & 0x01010101u; }

template <int W, int MS>
ACX_HD void unpack_row_static(uint8_t* row, const Rel<W> (&r)[2]) {
    constexpr int NW = MS / 4;
    uint32_t w[2 * NW];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const BV<W> N = bv_mask<W>(r[h].n);
        const BV<W> X = bv_xor<W>(r[h].S, r[h].T);
#pragma unroll
        for (int k = 0; k < NW; ++k) {
            const int g = k / 8, sh = 4 * (k % 8);
            const uint32_t nn = (N.w[g] >> sh) & 0xFu;
            const uint32_t ss = (r[h].S.w[g] >> sh) & 0xFu;
            const uint32_t xx = (X.w[g] >> sh) & 0xFu;
            w[h * NW + k] = spread4(nn) + spread4(xx) + 253u * spread4(ss);
        }
    }
    uint2* p = reinterpret_cast<uint2*>(row);
#pragma unroll
    for (int k = 0; k < NW; ++k) p[k] = make_uint2(w[2 * k], w[2 * k + 1]);
}

// ---- dynamic width: any 1 <= m <= 32*W, relators may start at any byte ------------------
template <int W>
ACX_HD void pack_rel_dynamic(const uint8_t* rel, int m, RawRel<W>& raw, uint32_t& chk) {
    raw.B0 = bv_zero<W>();
    raw.B1 = bv_zero<W>();
    raw.S7 = bv_zero<W>();
    const int nw = (m + 3) >> 2;
    for (int k = nw - 1; k >= 0; --k) {
        uint32_t x = 0u;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int idx = 4 * k + j;
            const uint32_t b = (idx < m) ? (uint32_t)rel[idx] : 0u;
            x |= b << (8 * j);
        }
        // make room: shift every plane left by one nibble, then insert at the bottom
#pragma unroll
        for (int i = W - 1; i > 0; --i) {
            raw.B0.w[i] = fshl(raw.B0.w[i - 1], raw.B0.w[i], 4);
            raw.B1.w[i] = fshl(raw.B1.w[i - 1], raw.B1.w[i], 4);
            raw.S7.w[i] = fshl(raw.S7.w[i - 1], raw.S7.w[i], 4);
        }
        pack_word(x, raw.B0.w[0], raw.B1.w[0], raw.S7.w[0], chk);
    }
}

template <int W>
ACX_HD void unpack_rel_dynamic(uint8_t* rel, int m, const Rel<W>& r) {
    BV<W> N = bv_mask<W>(r.n), S = r.S, X = bv_xor<W>(r.S, r.T);
    const int nw = (m + 3) >> 2;
    for (int k = 0; k < nw; ++k) {
        const uint32_t x = spread4(N.w[0] & 0xFu) + spread4(X.w[0] & 0xFu) + 253u * spread4(S.w[0] & 0xFu);
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (4 * k + j < m) rel[4 * k + j] = (uint8_t)(x >> (8 * j));
#pragma unroll
        for (int i = 0; i < W; ++i) {
            N.w[i] = fshr(N.w[i], (i + 1 < W) ? N.w[i + 1] : 0u, 4);
            S.w[i] = fshr(S.w[i], (i + 1 < W) ? S.w[i + 1] : 0u, 4);
            X.w[i] = fshr(X.w[i], (i + 1 < W) ? X.w[i + 1] : 0u, 4);
        }
    }
}

// ---- byte-wise slow path: literal restatement of the reference on one row ---------------
// Taken only for rows with letters outside {0,+-1,+-2} or with a non-reduced relator.
// `in` and `out` may be the same buffer.
// mode: bit 0 = cyclic reduction; bit 1 = no move (simplify_presentation only, utils.py:246-283);
//       bit 2 = move only (concatenate_relators / conjugate as the reference exposes them:
//       junction cancellation + guard, no validity assert, no reduction).
// Follows ac_moves.py:4-156 (move, junction cancel, guard), utils.py:264-266 (validity),
// utils.py:211-232 (stack free reduction, cyclic reduction).
static __host__ __device__ __noinline__ void slow_acmove(const int8_t* in, int8_t* out, int m, int move, int cyc,
                                         int& l0, int& l1, uint32_t& status) {
    int cnt[2];
    bool valid = true;
    for (int h = 0; h < 2; ++h) {
        int c = 0;
        for (int k = 0; k < m; ++k) {
            const int8_t v = in[h * m + k];
            out[h * m + k] = v;
            c += (v != 0);
        }
        cnt[h] = c;
        if (c == 0) valid = false;
        for (int k = c; k < m; ++k)
            if (in[h * m + k] != 0) valid = false;
    }
    l0 = cnt[0];
    l1 = cnt[1];
    if (!valid) {
        status = ACX_ROW_INVALID_INPUT;
        return;
    }
    const int raw = (cyc >> 2) & 1, no_move = (cyc >> 1) & 1;
    cyc &= 1;
    const uint32_t d = decode_move(no_move ? 0 : move);
    const int i = d & 1, is_conj = (d >> 1) & 1, inv = (d >> 2) & 1, gy = (d >> 3) & 1;
    int8_t* a = out + i * m;
    const int8_t* b = in + (1 - i) * m;
    int la = cnt[i];
    const int lb = cnt[1 - i];
    if (no_move) {
    } else if (!is_conj) {
        int acc = 0;
        const int lim = la < lb ? la : lb;
        while (acc < lim) {
            // r_j^-1 is materialised as int8 (so -(-128) wraps), the comparison is done in int
            const int bv = inv ? (int)(int8_t)(-b[lb - 1 - acc]) : (int)b[acc];
            if ((int)a[la - 1 - acc] != -bv) break;
            ++acc;
        }
        const int nn = la + lb - 2 * acc;
        if (nn == 0 && !raw) {  // utils.py:264-266 would raise; row is handed back unchanged
            status = ACX_ROW_EMPTY_RELATOR;
            return;
        }
        if (nn <= m) {
            for (int t = 0; t < lb - acc; ++t)
                a[la - acc + t] = inv ? (int8_t)(-b[lb - 1 - (acc + t)]) : b[acc + t];
            for (int k = nn; k < m; ++k) a[k] = 0;
            la = nn;
        }
    } else {
        const int g = (gy ? 2 : 1) * (inv ? -1 : 1);
        const int sc = ((int)a[0] == -g), ec = ((int)a[la - 1] == g);
        const int nn = la + 2 - 2 * (sc + ec);
        if (nn == 0 && !raw) {
            status = ACX_ROW_EMPTY_RELATOR;
            return;
        }
        if (nn <= m) {
            const int core = la - sc - ec;
            int pos;
            if (!sc) {
                for (int k = core - 1; k >= 0; --k) a[k + 1] = a[k];
                a[0] = (int8_t)g;
                pos = core + 1;
            } else {
                for (int k = 0; k < core; ++k) a[k] = a[k + 1];
                pos = core;
            }
            if (!ec) a[pos++] = (int8_t)(-g);
            for (int k = pos; k < m; ++k) a[k] = 0;
            la = nn;
        }
    }
    if (raw) {
        l0 = i ? cnt[0] : la;
        l1 = i ? la : cnt[1];
        status = ACX_ROW_OK;
        return;
    }
    int len[2];
    for (int h = 0; h < 2; ++h) {
        int8_t* r = out + h * m;
        const int n0 = (h == i) ? la : cnt[h];
        int top = 0;
        for (int k = 0; k < n0; ++k) {  // utils.py:211-220
            const int8_t v = r[k];
            if (top > 0 && (int)r[top - 1] == -(int)v) --top;
            else r[top++] = v;
        }
        int n = top;
        if (cyc && n > 0) {  // utils.py:223-232
            int k = 0;
            while (2 * k + 1 < n && (int)r[k] == -(int)r[n - 1 - k]) ++k;
            if (k) {
                for (int t = 0; t < n - 2 * k; ++t) r[t] = r[t + k];
                n -= 2 * k;
            }
        }
        for (int k = n; k < n0; ++k) r[k] = 0;
        len[h] = n;
    }
    l0 = len[0];
    l1 = len[1];
    status = ACX_ROW_OK;
}

template <int W>
ACX_HD Rel<W> rel_select(bool second, const Rel<W>& r0, const Rel<W>& r1) {
    Rel<W> o;
#pragma unroll
    for (int i = 0; i < W; ++i) {
        o.S.w[i] = second ? r1.S.w[i] : r0.S.w[i];
        o.T.w[i] = second ? r1.T.w[i] : r0.T.w[i];
    }
    o.n = second ? r1.n : r0.n;
    return o;
}

template <int W>
ACX_HD int classify_row(const RawRel<W> (&raw)[2], uint32_t chk, Rel<W> (&r)[2], int (&cnt)[2]) {
    bool alpha_ok = (chk & 0x7C7C7C7Cu) == 0u;
    bool valid = true, reduced = true;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const BV<W> bad3 = bv_and<W>(bv_andn<W>(raw[h].B1, raw[h].S7), raw[h].B0);  // byte 0x03
        const BV<W> badn = bv_andn<W>(raw[h].S7, raw[h].B1);                        // 0xFC, 0xFD
        alpha_ok = alpha_ok && !bv_any<W>(bv_or<W>(bad3, badn));
        const BV<W> N = bv_or<W>(raw[h].B0, raw[h].B1);
        const int n = bv_popc<W>(N);
        cnt[h] = n;
        valid = valid && (n > 0) && bv_eq<W>(N, bv_mask<W>(n));
        r[h].S = raw[h].S7;
        r[h].T = bv_andn<W>(raw[h].B1, raw[h].B0);
        r[h].n = n;
        reduced = reduced && !rel_has_adjacent_inverse<W>(r[h]);
    }
    if (!alpha_ok) return ROW_SLOW;  // the byte path also does the validity test
    if (!valid) return ROW_INVALID;
    return reduced ? ROW_FAST : ROW_SLOW;
}

// Apply move `move` to the planes (fast path).  Returns the row status.
template <int W>
ACX_HD uint32_t apply_move_planes(Rel<W> (&r)[2], int move, int m, bool cyc) {
    const uint32_t d = decode_move(move);
    const bool i = d & 1u, is_conj = d & 2u, inv = d & 4u;
    const uint32_t gy = (d >> 3) & 1u;
    Rel<W> a = rel_select<W>(i, r[0], r[1]);
    if (!is_conj) {
        Rel<W> b = rel_select<W>(!i, r[0], r[1]);
        if (inv) b = rel_inverse<W>(b);
        rel_concat<W>(a, b, m);
    } else {
        rel_conjugate<W>(a, inv ? 1u : 0u, gy, m);
    }
    if (a.n == 0) return ACX_ROW_EMPTY_RELATOR;
    if (i) r[1] = a;
    else r[0] = a;
    if (cyc) {
        rel_cyclic_reduce<W>(r[0]);
        rel_cyclic_reduce<W>(r[1]);
    }
    return ACX_ROW_OK;
}

// One ACMove on one row in shared memory: in_row -> out_row (may be the same buffer: the row is
// read completely before anything is written on the plane path).
// MS > 0: static width (MS % 4 == 0, rows 8-byte aligned); MS == 0: dynamic width m.
template <int W, int MS>
ACX_HD void acmove_row(const uint8_t* in_row, uint8_t* out_row, int m_rt, int move, bool cyc,
                                           int& l0, int& l1, uint32_t& status) {
    const int m = MS ? MS : m_rt;
    RawRel<W> raw[2];
    uint32_t chk = 0u;
    if (MS) {
        pack_row_static<W, (MS ? MS : 4)>(in_row, raw, chk);
    } else {
        pack_rel_dynamic<W>(in_row, m, raw[0], chk);
        pack_rel_dynamic<W>(in_row + m, m, raw[1], chk);
    }
    Rel<W> r[2];
    int cnt[2];
    const int cls = classify_row<W>(raw, chk, r, cnt);
    if (cls == ROW_SLOW) {
        slow_acmove(reinterpret_cast<const int8_t*>(in_row), reinterpret_cast<int8_t*>(out_row), m, move, cyc ? 1 : 0,
                    l0, l1, status);
        return;
    }
    status = (cls == ROW_INVALID) ? (uint32_t)ACX_ROW_INVALID_INPUT : apply_move_planes<W>(r, move, m, cyc);
    if (status != ACX_ROW_OK) {  // row handed back unchanged, lengths = non-zero counts of the input
        if (out_row != in_row)
            for (int k = 0; k < 2 * m; ++k) out_row[k] = in_row[k];
        l0 = cnt[0];
        l1 = cnt[1];
        return;
    }
    if (MS) {
        unpack_row_static<W, (MS ? MS : 4)>(out_row, r);
    } else {
        unpack_rel_dynamic<W>(out_row, m, r[0]);
        unpack_rel_dynamic<W>(out_row + m, m, r[1]);
    }
    l0 = r[0].n;
    l1 = r[1].n;
}

// =============================================================================================
// m <= 64: the 64-bit plane path (acx_rel64.cuh).  Works IN PLACE on the row: only a relator whose
// letters changed is written back (the other bytes already hold the right answer).
// =============================================================================================
template <int MS>
ACX_HD void pack_row_static64(const uint8_t* row, Raw64 (&raw)[2], uint32_t& chk) {
    constexpr int NW = MS / 4;
    uint32_t w[2 * NW];
    const uint2* p = reinterpret_cast<const uint2*>(row);
#pragma unroll
    for (int k = 0; k < NW; ++k) {
        const uint2 v = p[k];
        w[2 * k] = v.x;
        w[2 * k + 1] = v.y;
    }
    chk = 0u;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        uint32_t b0[2] = {0u, 0u}, b1[2] = {0u, 0u}, s7[2] = {0u, 0u};
#pragma unroll
        for (int k = NW - 1; k >= 0; --k) pack_word(w[h * NW + k], b0[k / 8], b1[k / 8], s7[k / 8], chk);
        raw[h].B0 = ((uint64_t)b0[1] << 32) | b0[0];
        raw[h].B1 = ((uint64_t)b1[1] << 32) | b1[0];
        raw[h].S7 = ((uint64_t)s7[1] << 32) | s7[0];
    }
}

// planes -> letters of ONE relator, 32-bit stores (rel is 4-byte aligned)
template <int MS>
ACX_HD void unpack_rel_static64(uint8_t* rel, const R64& r) {
    constexpr int NW = MS / 4;
    const uint64_t N = mask64(r.n), X = r.S ^ r.T;
    const uint32_t n2[2] = {(uint32_t)N, (uint32_t)(N >> 32)};
    const uint32_t s2[2] = {(uint32_t)r.S, (uint32_t)(r.S >> 32)};
    const uint32_t x2[2] = {(uint32_t)X, (uint32_t)(X >> 32)};
    uint32_t* p = reinterpret_cast<uint32_t*>(rel);
#pragma unroll
    for (int k = 0; k < NW; ++k) {
        const int g = k / 8, sh = 4 * (k % 8);
        const uint32_t nn = (n2[g] >> sh) & 0xFu, ss = (s2[g] >> sh) & 0xFu, xx = (x2[g] >> sh) & 0xFu;
        p[k] = spread4(nn) + spread4(xx) + 253u * spread4(ss);
    }
}

ACX_HD void pack_rel_dynamic64(const uint8_t* rel, int m, Raw64& raw, uint32_t& chk) {
    uint32_t b0[2] = {0u, 0u}, b1[2] = {0u, 0u}, s7[2] = {0u, 0u};
    const int nw = (m + 3) >> 2;
    for (int k = nw - 1; k >= 0; --k) {
        uint32_t x = 0u;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int idx = 4 * k + j;
            const uint32_t b = (idx < m) ? (uint32_t)rel[idx] : 0u;
            x |= b << (8 * j);
        }
        b0[1] = fshl(b0[0], b0[1], 4);
        b1[1] = fshl(b1[0], b1[1], 4);
        s7[1] = fshl(s7[0], s7[1], 4);
        pack_word(x, b0[0], b1[0], s7[0], chk);
    }
    raw.B0 = ((uint64_t)b0[1] << 32) | b0[0];
    raw.B1 = ((uint64_t)b1[1] << 32) | b1[0];
    raw.S7 = ((uint64_t)s7[1] << 32) | s7[0];
}

ACX_HD void unpack_rel_dynamic64(uint8_t* rel, int m, const R64& r) {
    uint64_t N = mask64(r.n), S = r.S, X = r.S ^ r.T;
    const int nw = (m + 3) >> 2;
    for (int k = 0; k < nw; ++k) {
        const uint32_t x = spread4((uint32_t)N & 0xFu) + spread4((uint32_t)X & 0xFu) + 253u * spread4((uint32_t)S & 0xFu);
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (4 * k + j < m) rel[4 * k + j] = (uint8_t)(x >> (8 * j));
        N >>= 4;
        S >>= 4;
        X >>= 4;
    }
}

// One ACMove on one row, in place.  MS > 0: static width (MS % 4 == 0, MS <= 64, row 8-byte
// aligned); MS == 0: any 1 <= m <= 64.
template <int MS>
ACX_HD void acmove_row64(uint8_t* row, int m_rt, int move, bool cyc, int& l0, int& l1, uint32_t& status) {
    const int m = MS ? MS : m_rt;
    Raw64 raw[2];
    uint32_t chk = 0u;
    if (MS) {
        pack_row_static64<(MS ? MS : 4)>(row, raw, chk);
    } else {
        pack_rel_dynamic64(row, m, raw[0], chk);
        pack_rel_dynamic64(row + m, m, raw[1], chk);
    }
    R64 r[2];
    int cnt[2];
    const int cls = classify_row64(raw, chk, r, cnt);
    if (cls == ROW_SLOW) {
        slow_acmove(reinterpret_cast<const int8_t*>(row), reinterpret_cast<int8_t*>(row), m, move, cyc ? 1 : 0, l0, l1,
                    status);
        return;
    }
    l0 = cnt[0];
    l1 = cnt[1];
    if (cls == ROW_INVALID) {
        status = ACX_ROW_INVALID_INPUT;
        return;
    }
    const uint32_t d = decode_move(move);
    const bool i = d & 1u, is_conj = d & 2u, inv = d & 4u;
    R64 a, o;  // touched relator, other relator
    a.S = i ? r[1].S : r[0].S;
    a.T = i ? r[1].T : r[0].T;
    a.n = i ? r[1].n : r[0].n;
    o.S = i ? r[0].S : r[1].S;
    o.T = i ? r[0].T : r[1].T;
    o.n = i ? r[0].n : r[1].n;
    bool a_changed, o_changed = false;
    if (!is_conj) {
        R64 b = o;
        if (inv) b = r64_inverse(b);
        a_changed = r64_concat(a, b, m);
    } else {
        a_changed = r64_conjugate(a, inv ? 1u : 0u, (d >> 3) & 1u, m);
    }
    if (a.n == 0) {  // utils.py:264-266 raises; row stays as it was
        status = ACX_ROW_EMPTY_RELATOR;
        return;
    }
    if (cyc) {
        if (r64_cyclic_reduce(a) > 0) a_changed = true;
        if (r64_needs_cyclic(o)) {
            r64_cyclic_reduce(o);
            o_changed = true;
        }
    }
    uint8_t* pa = row + (d & 1u) * m;
    uint8_t* po = row + ((d & 1u) ^ 1u) * m;
    if (a_changed) {
        if (MS) unpack_rel_static64<(MS ? MS : 4)>(pa, a);
        else unpack_rel_dynamic64(pa, m, a);
    }
    if (o_changed) {
        if (MS) unpack_rel_static64<(MS ? MS : 4)>(po, o);
        else unpack_rel_dynamic64(po, m, o);
    }
    l0 = i ? o.n : a.n;
    l1 = i ? a.n : o.n;
    status = ACX_ROW_OK;
}

// =============================================================================================
// CERTIFIED rows (env state owned by the engine).  A row is certified when its letters are in
// {0,+-1,+-2}, zeros only pad on the right, and both relators are freely AND cyclically reduced;
// every ACMove(cyclical=True) output has these properties, so an env whose initial states are
// certified (acx_rows_certify) stays certified for ever and its lengths are known.  Then:
//   * the non-zero plane is mask(len): two planes per word instead of three, no classification;
//   * conjugation of a cyclically reduced word by g is a rotation (first letter == g^-1: rotate
//     left; last letter == g: rotate right) or the identity -- g a g^-1 reduces straight back to a,
//     and the length guard cannot reject a rotation (ac_moves.py:119-126 + utils.py:223-232);
//   * the untouched relator never changes.
// Results are identical to acmove_row64 on certified rows (tests/test_gpu_moves.py, host harness).
// =============================================================================================
ACX_HD void pack_word2(uint32_t x, uint32_t& b0, uint32_t& s7) {
    b0 = nib_push((x & 0x01010101u) * 0x10204080u, b0);
    s7 = nib_push((x & 0x80808080u) * 0x00204081u, s7);
}

template <int MS>
ACX_HD uint32_t acmove_row64_certified(uint8_t* row, int move, int& l0, int& l1) {
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
    const uint32_t d = decode_move(move);
    const bool i = d & 1u, is_conj = d & 2u, inv = d & 4u;
    R64 a, o;
    a.S = i ? r[1].S : r[0].S;
    a.T = i ? r[1].T : r[0].T;
    a.n = i ? r[1].n : r[0].n;
    o.S = i ? r[0].S : r[1].S;
    o.T = i ? r[0].T : r[1].T;
    o.n = i ? r[0].n : r[1].n;
    bool changed;
    if (!is_conj) {
        R64 b = o;
        if (inv) b = r64_inverse(b);
        changed = r64_concat(a, b, MS);
        if (a.n == 0) return ACX_ROW_EMPTY_RELATOR;  // utils.py:264-266 raises; row stays as it was
        if (changed) r64_cyclic_reduce(a);
    } else {
        const uint32_t gs = inv ? 1u : 0u, gt = (d >> 3) & 1u;
        const int n = a.n;
        const uint32_t s0 = (uint32_t)a.S & 1u, t0 = (uint32_t)a.T & 1u;
        const uint32_t sl = (uint32_t)shr64(a.S, n - 1) & 1u, tl = (uint32_t)shr64(a.T, n - 1) & 1u;
        const bool sc = (t0 == gt) && (s0 != gs);  // first letter == g^-1: a[1:] + [g^-1]
        const bool ec = (tl == gt) && (sl == gs);  // last letter == g:     [g] + a[:-1]
        changed = (sc || ec) && n > 1;
        if (sc) {
            a.S = (a.S >> 1) | shl64((uint64_t)s0, n - 1);
            a.T = (a.T >> 1) | shl64((uint64_t)t0, n - 1);
        } else if (ec) {
            const uint64_t mk = mask64(n);
            a.S = ((a.S << 1) & mk) | (uint64_t)sl;
            a.T = ((a.T << 1) & mk) | (uint64_t)tl;
        }
    }
    if (changed) unpack_rel_static64<MS>(row + (d & 1u) * MS, a);  // one copy of the code, address by arithmetic
    l0 = i ? o.n : a.n;
    l1 = i ? a.n : o.n;
    return ACX_ROW_OK;
}

// is this row certified (see above)?  Also returns the relator lengths (non-zero counts).
ACX_HD bool row_is_certified(const uint8_t* row, int m, int& l0, int& l1) {
    Raw64 raw[2];
    uint32_t chk = 0u;
    pack_rel_dynamic64(row, m, raw[0], chk);
    pack_rel_dynamic64(row + m, m, raw[1], chk);
    R64 r[2];
    int cnt[2];
    const int cls = classify_row64(raw, chk, r, cnt);
    l0 = cnt[0];
    l1 = cnt[1];
    if (cls != ROW_FAST) return false;
    return !r64_needs_cyclic(r[0]) && !r64_needs_cyclic(r[1]);
}

}  // namespace acx
