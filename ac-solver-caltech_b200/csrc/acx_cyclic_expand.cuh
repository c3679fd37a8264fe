// acx_cyclic_expand.cuh -- expand_path_with_cyclic_reductions (value_search/value_guided_search.py:581-635;
// SURVEY 8(f) rank 2): replay a compact path found with cyclical=True using free reduction only, and after every
// move emit the explicit conjugation moves that perform the cyclic reductions the search applied implicitly
// (removing g ... g^-1 from the ends of r_i is the conjugation of r_i by g^-1, :563-578).  One call = one path;
// the row arithmetic is the lockstep kernels' (acx_rows.cuh), so the function also runs on the host for
// tests/host_harness.
#pragma once
#include "acx_rows.cuh"

namespace acx {

static constexpr uint32_t ACX_PATH_OVERFLOW = 3;  // the output arrays are too short (caller retries with more room)

// _CYCLIC_REDUCE_MOVE (:566-578): (relator index, first letter) -> conjugation move; -1 = no such letter
ACX_HD int cyclic_reduce_move(int rel, int first) {
    switch (first) {
        case 1: return rel ? 4 : 11;
        case -1: return rel ? 8 : 7;
        case 2: return rel ? 6 : 5;
        case -2: return rel ? 10 : 9;
        default: return -1;
    }
}

// row: 2m bytes of 8-byte aligned scratch holding the initial presentation (rewritten in place).
// Returns the number of (action, total_length) pairs written; status / fail_step as in acx_path_replay.
template <int W, int MS>
ACX_HD int expand_cyclic_path(uint8_t* row, int m_rt, const uint8_t* actions, int n, uint8_t* out_actions, uint16_t* out_totals,
                              int cap, uint32_t& status, int& fail_step) {
    const int m = MS ? MS : m_rt;
    int l0 = 0, l1 = 0;
    for (int k = 0; k < m; ++k) {
        l0 += row[k] != 0;
        l1 += row[m + k] != 0;
    }
    status = ACX_ROW_OK;
    fail_step = -1;
    int o = 0;
    for (int t = 0; t < n; ++t) {
        int move = actions[t];
        while (move >= 0) {
            uint32_t s1 = move < 12 ? (uint32_t)ACX_ROW_OK : (uint32_t)ACX_ROW_INVALID_INPUT;
            if (s1 == ACX_ROW_OK) {
                if constexpr (W <= 2) acmove_row64<MS>(row, m, move, false, l0, l1, s1);  // :612, :628 (cyclical=False)
                else acmove_row<W, MS>(row, row, m, move, false, l0, l1, s1);
            }
            if (s1 == ACX_ROW_OK && o >= cap) s1 = ACX_PATH_OVERFLOW;
            if (s1 != ACX_ROW_OK) {
                status = s1;
                fail_step = t;
                return o;
            }
            out_actions[o] = (uint8_t)move;
            out_totals[o] = (uint16_t)(l0 + l1);
            ++o;
            // :616-634: the first relator whose ends cancel gets one conjugation; then both are re-checked
            move = -1;
            for (int rel = 0; rel < 2 && move < 0; ++rel) {
                const int len = rel ? l1 : l0;
                if (len <= 1) continue;
                const int first = (int8_t)row[rel * m], last = (int8_t)row[rel * m + len - 1];
                if (first != -last) continue;
                move = cyclic_reduce_move(rel, first);
                if (move < 0) {  // the reference's table has no entry (KeyError)
                    status = ACX_ROW_INVALID_INPUT;
                    fail_step = t;
                    return o;
                }
            }
        }
    }
    return o;
}

}  // namespace acx
