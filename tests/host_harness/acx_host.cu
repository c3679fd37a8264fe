// tests/host_harness/acx_host.cu -- TEST INFRASTRUCTURE.
// Runs the product's row arithmetic (csrc/acx_rows.cuh, __host__ __device__) on the CPU so that
// `pytest -m "not gpu"` can compare it with the oracle before any GPU time is spent.  Nothing in
// the product loads this library.
#include <stdint.h>
#include <string.h>

#include <vector>

#include "../../ac-solver-caltech_b200/csrc/acx_rows.cuh"

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

// variant: 0 = dynamic width (generic planes), 1 = static generic planes (m must be 36 or 128),
//          2 = 64-bit planes dynamic (m <= 64), 3 = 64-bit planes static (m == 36),
//          4 = certified-row env path (m == 36, cyclical only)
extern "C" int acxh_move_batch(const int8_t* in, const uint8_t* moves, int8_t* out, uint8_t* lens, uint8_t* status,
                               int64_t n, int m, int cyc, int variant, int inplace) {
    if (variant == 4) {
        if (m != 36 || !cyc) return -1;
        run_cert36(in, moves, out, lens, status, n);
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
