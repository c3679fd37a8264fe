// tests/host_harness/acx_host.cu -- TEST INFRASTRUCTURE.
// Runs the product's row arithmetic (csrc/acx_rows.cuh, __host__ __device__) on the CPU so that
// `pytest -m "not gpu"` can compare it with the oracle before any GPU time is spent.  Nothing in
// the product loads this library.
#include <stdint.h>
#include <string.h>

#include <vector>

#include "../../ac-solver-caltech_b200/csrc/acx_envrow.cuh"

using namespace acx;

template <int W, int MS>
static void run(const int8_t* in, const uint8_t* moves, int8_t* out, uint8_t* lens, uint8_t* status, int64_t n, int m,
                int cyc, int inplace) {
    const int rowb = 2 * m;
    std::vector<uint64_t> a(rowb / 8 + 2), b(rowb / 8 + 2);  // 8-byte aligned scratch rows
    for (int64_t r = 0; r < n; ++r) {
        uint8_t* ia = reinterpret_cast<uint8_t*>(a.data());
        uint8_t* ob = inplace ? ia : reinterpret_cast<uint8_t*>(b.data());
        memcpy(ia, in + r * rowb, rowb);
        int l0 = -1, l1 = -1;
        uint32_t st = 99;
        acmove_row<W, MS>(ia, ob, m, moves[r], cyc != 0, l0, l1, st);
        memcpy(out + r * rowb, ob, rowb);
        lens[2 * r] = (uint8_t)l0;
        lens[2 * r + 1] = (uint8_t)l1;
        status[r] = (uint8_t)st;
    }
}

template <int MS>
static void run64(const int8_t* in, const uint8_t* moves, int8_t* out, uint8_t* lens, uint8_t* status, int64_t n, int m,
                  int cyc) {
    const int rowb = 2 * m;
    std::vector<uint64_t> a(rowb / 8 + 2);
    for (int64_t r = 0; r < n; ++r) {
        uint8_t* ia = reinterpret_cast<uint8_t*>(a.data());
        memcpy(ia, in + r * rowb, rowb);
        int l0 = -1, l1 = -1;
        uint32_t st = 99;
        acmove_row64<MS>(ia, m, moves[r], cyc != 0, l0, l1, st);
        memcpy(out + r * rowb, ia, rowb);
        lens[2 * r] = (uint8_t)l0;
        lens[2 * r + 1] = (uint8_t)l1;
        status[r] = (uint8_t)st;
    }
}

// certified-row path (m == 36): rows that are not certified go through the checked path, as the
// env kernel does per launch
static void run_cert36(const int8_t* in, const uint8_t* moves, int8_t* out, uint8_t* lens, uint8_t* status, int64_t n) {
    const int m = 36, rowb = 72;
    std::vector<uint64_t> a(rowb / 8 + 2);
    for (int64_t r = 0; r < n; ++r) {
        uint8_t* ia = reinterpret_cast<uint8_t*>(a.data());
        memcpy(ia, in + r * rowb, rowb);
        int l0 = -1, l1 = -1;
        uint32_t st = 99;
        if (row_is_certified(ia, m, l0, l1)) st = acmove_row64_certified<36>(ia, moves[r], l0, l1);
        else acmove_row64<36>(ia, m, moves[r], true, l0, l1, st);
        memcpy(out + r * rowb, ia, rowb);
        lens[2 * r] = (uint8_t)l0;
        lens[2 * r + 1] = (uint8_t)l1;
        status[r] = (uint8_t)st;
    }
}

// class-split certified step of the env tile kernel (csrc/acx_envrow.cuh): conjugations on the bytes,
// concatenations on planes.  The scratch row is padded like the kernel's stage (the byte path reads the
// word in front of a relator); `junk` fills the padding so that nothing depends on it.
static void run_cert36_split(const int8_t* in, const uint8_t* moves, int8_t* out, uint8_t* lens, uint8_t* status, int64_t n,
                             uint8_t junk) {
    const int m = 36, rowb = 72;
    std::vector<uint64_t> a(rowb / 8 + 4);
    for (int64_t r = 0; r < n; ++r) {
        memset(a.data(), junk, a.size() * 8);
        uint8_t* ia = reinterpret_cast<uint8_t*>(a.data()) + 8;
        memcpy(ia, in + r * rowb, rowb);
        int l0 = -1, l1 = -1;
        uint32_t st = 99;
        if (row_is_certified(ia, m, l0, l1)) st = cert_row_step<36>(ia, moves[r], l0, l1);
        else acmove_row64<36>(ia, m, moves[r], true, l0, l1, st);
        memcpy(out + r * rowb, ia, rowb);
        lens[2 * r] = (uint8_t)l0;
        lens[2 * r + 1] = (uint8_t)l1;
        status[r] = (uint8_t)st;
    }
}

// register-resident conjugations of the frontier expansion (conj_row_words of csrc/acx_envrow.cuh, used by
// csrc/acx_expand.cu): the plain form when the search does not reduce cyclically, the rotation form on certified rows
// when it does; everything else through the checked row function
template <bool CERT>
static int conj_words36(uint8_t* ia, uint32_t d, int n) {
    const int m = 36;
    const bool i = d & 1u;
    const uint32_t mag = 1u + ((d >> 3) & 1u), neg = 0x100u - mag;
    const uint32_t g = (d & 4u) ? neg : mag, gi = (d & 4u) ? mag : neg;
    const uint8_t* rel = ia + (i ? m : 0);
    const bool sc = rel[0] == gi, ec = rel[n - 1] == g;
    uint32_t w[18], o[18], fv = 0;
    int fp = -1;
    memcpy(w, ia, 72);
    const int nn = conj_row_words<36, CERT>(w, o, i, n, sc, ec, d, fp, fv);
    memcpy(ia, o, 72);
    if (fp >= 0) ia[fp] = (uint8_t)fv;
    return nn;
}

static void run_conj_words36(const int8_t* in, const uint8_t* moves, int8_t* out, uint8_t* lens, uint8_t* status, int64_t n,
                             int cyc) {
    const int m = 36, rowb = 72;
    std::vector<uint64_t> a(rowb / 8 + 4);
    for (int64_t r = 0; r < n; ++r) {
        uint8_t* ia = reinterpret_cast<uint8_t*>(a.data()) + 8;
        memcpy(ia, in + r * rowb, rowb);
        int l0 = -1, l1 = -1;
        uint32_t st = 99;
        const uint32_t d = decode_move(moves[r]);
        bool done = false;
        if (d & 2u) {
            if (cyc) {
                if (row_is_certified(ia, m, l0, l1)) {
                    conj_words36<true>(ia, d, (d & 1u) ? l1 : l0);
                    st = ACX_ROW_OK;
                    done = true;
                }
            } else {
                Raw64 raw[2];
                uint32_t chk;
                pack_row_static64<36>(ia, raw, chk);
                R64 rr[2];
                int cnt[2];
                if (classify_row64(raw, chk, rr, cnt) == ROW_FAST) {
                    const int nn = conj_words36<false>(ia, d, cnt[d & 1u]);
                    l0 = (d & 1u) ? cnt[0] : nn;
                    l1 = (d & 1u) ? nn : cnt[1];
                    st = ACX_ROW_OK;
                    done = true;
                }
            }
        }
        if (!done) acmove_row64<36>(ia, m, moves[r], cyc != 0, l0, l1, st);
        memcpy(out + r * rowb, ia, rowb);
        lens[2 * r] = (uint8_t)l0;
        lens[2 * r + 1] = (uint8_t)l1;
        status[r] = (uint8_t)st;
    }
}

// variant: 0 = dynamic width (generic planes), 1 = static generic planes (m must be 36 or 128),
//          2 = 64-bit planes dynamic (m <= 64), 3 = 64-bit planes static (m == 36),
//          4 = certified-row env path (m == 36, cyclical only), 5/6 = class-split certified step (acx_envrow.cuh),
//          7 = register-resident conjugations of the frontier expansion (m == 36)
extern "C" int acxh_move_batch(const int8_t* in, const uint8_t* moves, int8_t* out, uint8_t* lens, uint8_t* status,
                               int64_t n, int m, int cyc, int variant, int inplace) {
    if (variant == 4) {
        if (m != 36 || !cyc) return -1;
        run_cert36(in, moves, out, lens, status, n);
        return 0;
    }
    if (variant == 5 || variant == 6) {  // class-split certified step; 6 = padding filled with 0xFF instead of 0
        if (m != 36 || !cyc) return -1;
        run_cert36_split(in, moves, out, lens, status, n, variant == 6 ? 0xFF : 0x00);
        return 0;
    }
    if (variant == 7) {
        if (m != 36) return -1;
        run_conj_words36(in, moves, out, lens, status, n, cyc);
        return 0;
    }
    if (variant == 2) {
        if (m > 64) return -1;
        run64<0>(in, moves, out, lens, status, n, m, cyc);
        return 0;
    }
    if (variant == 3) {
        if (m != 36) return -1;
        run64<36>(in, moves, out, lens, status, n, m, cyc);
        return 0;
    }
    if (variant == 1) {
        if (m == 36) run<2, 36>(in, moves, out, lens, status, n, m, cyc, inplace);
        else if (m == 128) run<4, 128>(in, moves, out, lens, status, n, m, cyc, inplace);
        else return -1;
        return 0;
    }
    if (m <= 32) run<1, 0>(in, moves, out, lens, status, n, m, cyc, inplace);
    else if (m <= 64) run<2, 0>(in, moves, out, lens, status, n, m, cyc, inplace);
    else if (m <= 128) run<4, 0>(in, moves, out, lens, status, n, m, cyc, inplace);
    else return -1;
    return 0;
}

// ---------------------------------------------------------------------------------------------------
// Graph engine (csrc/acx_graph.cuh) on the CPU: the same functors and the same graph_step<> template as
// libacx.so, driven by a sequential executor.  Order-free functors run in a permuted order
// (acxh_graph_set_order: 0 forward, 1 reverse, 2 strided permutation) so that the claim-replacement path
// of the visited-set probe, which a forward sweep never takes, is exercised too.
#include "../../ac-solver-caltech_b200/csrc/acx_graph.cuh"

static int g_order = 0;

struct HostExec {
    template <class F>
    void run(int64_t n, const F& f) {
        for (int64_t i = 0; i < n; ++i) f(i);
    }
    template <class F>
    void run_shuffled(int64_t n, const F& f) {
        if (g_order == 0) {
            for (int64_t i = 0; i < n; ++i) f(i);
        } else if (g_order == 1) {
            for (int64_t i = n - 1; i >= 0; --i) f(i);
        } else {
            if (n <= 0) return;
            int64_t step = 7919;  // a prime that does not divide n: i -> (i * step) % n is a permutation
            if (n % step == 0) step = 7927;
            if (n % step == 0) step = 1;
            for (int64_t i = 0; i < n; ++i) f((int64_t)(((__int128)i * step) % n));
        }
    }
    // one block at a time: phase1 for every thread, then phase2 (what the barrier guarantees); the thread order inside a
    // phase follows g_order
    template <class F>
    void run_blocks(int64_t blocks, int threads, size_t smem_bytes, const F& f) {
        std::vector<uint64_t> smem(smem_bytes / 8 + 2);
        uint8_t* sm = reinterpret_cast<uint8_t*>(smem.data());
        for (int64_t b = 0; b < blocks; ++b) {
            const int64_t bb = g_order == 1 ? blocks - 1 - b : b;
            run_shuffled(threads, [&](int64_t t) { f.phase1(bb, (int)t, sm); });
            run_shuffled(threads, [&](int64_t t) { f.phase2(bb, (int)t, sm); });
        }
    }
    int scan(const uint8_t* flag, int32_t* excl, int64_t n, void*, int64_t) {
        int32_t run = 0;
        for (int64_t i = 0; i < n; ++i) {
            excl[i] = run;
            run += flag[i];
        }
        return 0;
    }
};

#define ACXH_DISPATCH(w, CALL)                          \
    switch (w) {                                        \
        case 2: { constexpr int W = 2; CALL; } break;   \
        case 4: { constexpr int W = 4; CALL; } break;   \
        case 8: { constexpr int W = 8; CALL; } break;   \
        default: return -1;                             \
    }

extern "C" void acxh_graph_set_order(int order) { g_order = order; }
extern "C" int64_t acxh_graph_scan_bytes(int64_t) { return 1; }
extern "C" int acxh_graph_step(const acx_graph_step_t* a, void*) {
    HostExec ex;
    int rc = 0;
    ACXH_DISPATCH(a->width_words, rc = (graph_step<W, HostExec>(ex, *a)));
    return rc;
}
extern "C" int acxh_graph_pack(const int8_t* rows, int64_t n, int m, const int32_t* inst, int w, uint32_t* nodes, uint8_t* ok, void*) {
    HostExec ex;
    ACXH_DISPATCH(w, ex.run(n, GPackRows<W>{rows, m, inst, nodes, ok}));
    return 0;
}
extern "C" int acxh_graph_unpack(const uint32_t* nodes, int64_t n, int w, int m, int8_t* rows, int32_t* inst, void*) {
    HostExec ex;
    ACXH_DISPATCH(w, ex.run(n, GUnpackRows<W>{nodes, m, rows, inst}));
    return 0;
}
extern "C" int acxh_graph_rehash(const uint32_t* nodes, int64_t n, int w, uint64_t* table, int64_t tcap, void*) {
    HostExec ex;
    ACXH_DISPATCH(w, ex.run_shuffled(n, GRehash<W>{nodes, table, tcap}));
    return 0;
}
extern "C" int acxh_graph_count_instances(const uint32_t* nodes, int64_t n, int w, int64_t* counts, int64_t n_inst, void*) {
    HostExec ex;
    ACXH_DISPATCH(w, ex.run(n, GCountInst<W>{nodes, counts, n_inst}));
    return 0;
}

// Miller-Schupp enumeration functor (csrc/acx_ms.cuh) on the CPU
#include "../../ac-solver-caltech_b200/csrc/acx_ms.cuh"
extern "C" int acxh_ms_enumerate(int w_len, int64_t first, int64_t count, uint64_t* key, uint64_t* rel, void*) {
    HostExec ex;
    ex.run(count, MsEnumerate{w_len, first, key, rel});
    return 0;
}

// expand_path_with_cyclic_reductions (csrc/acx_cyclic_expand.cuh) on the CPU, one path per call of the HD function
#include "../../ac-solver-caltech_b200/csrc/acx_cyclic_expand.cuh"
extern "C" int acxh_path_expand_cyclic(const int8_t* roots, const uint8_t* actions, const int32_t* path_len, int64_t P, int T, int m,
                                       uint8_t* out_actions, uint16_t* out_totals, int cap, int32_t* out_len, int32_t* fail_step,
                                       uint8_t* status, void*) {
    std::vector<uint64_t> scratch(2 * m / 8 + 2);
    for (int64_t p = 0; p < P; ++p) {
        uint8_t* row = reinterpret_cast<uint8_t*>(scratch.data());
        memcpy(row, roots + p * 2 * m, 2 * m);
        const int n = std::min(path_len ? path_len[p] : T, T);
        uint32_t st;
        int fail;
        int o;
        if (m <= 32) o = expand_cyclic_path<1, 0>(row, m, actions + p * T, n, out_actions + p * cap, out_totals + p * cap, cap, st, fail);
        else if (m <= 64) o = expand_cyclic_path<2, 0>(row, m, actions + p * T, n, out_actions + p * cap, out_totals + p * cap, cap, st, fail);
        else o = expand_cyclic_path<4, 0>(row, m, actions + p * T, n, out_actions + p * cap, out_totals + p * cap, cap, st, fail);
        out_len[p] = o;
        fail_step[p] = fail;
        status[p] = (uint8_t)st;
    }
    return 0;
}

// one move of the graph engine (g_move + sort) on byte rows, for a direct comparison with the oracle's Presentation::move
template <int W>
static void graph_move_rows(const int8_t* rows, int64_t n, int m, const uint8_t* moves, int classic, int allow_empty, int8_t* out,
                            int32_t* status) {
    std::vector<uint32_t> node(4 * W + 8);
    std::vector<uint8_t> ok(1);
    for (int64_t j = 0; j < n; ++j) {
        // 32-byte aligned scratch is not needed on the host path of g_load / g_store
        GPackRows<W>{rows + j * 2 * m, m, nullptr, node.data(), ok.data()}(0);
        if (!ok[0]) {
            status[j] = -1;
            continue;
        }
        const GPres<W> p = g_unpack<W>(g_load<W>(node.data(), 0));
        GPres<W> c;
        status[j] = g_move<W>(p, moves[j], classic != 0, 32 * W - 8, allow_empty != 0, c);
        if (status[j] != G_MOVE_OK) continue;
        g_store<W>(node.data(), 0, g_pack<W>(c));
        GUnpackRows<W>{node.data(), m, out + j * 2 * m, nullptr}(0);
    }
}

extern "C" int acxh_graph_move(const int8_t* rows, int64_t n, int m, const uint8_t* moves, int classic, int allow_empty, int w,
                               int8_t* out, int32_t* status) {
    ACXH_DISPATCH(w, (graph_move_rows<W>(rows, n, m, moves, classic, allow_empty, out, status)));
    return 0;
}
