// acx_replay.cu -- batched path replay: ACMove applied along a time axis.
//
// Replaces the Python loops that re-apply a list of actions to an initial presentation, one ACMove per
// step: agents/training.py:74-135 (replay_final_state, actions_to_path), value_search/replay_path.py:107-160,
// value_search/data_extraction.py:53-190.  One THREAD owns one path: the row stays in shared memory for the
// whole path, every step is the same row function as the lockstep kernels (acmove_row64 / acmove_row), and
// every intermediate state and length pair is emitted.  A step whose move empties a relator (the reference
// raises AssertionError, utils.py:264-266), an invalid row or an action outside 0..11 stops that path:
// status[p] names the reason, fail_step[p] the step.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/acx.h"
#include "acx_common.h"
#include "acx_rows.cuh"
#include "acx_cyclic_expand.cuh"

namespace acx {

// MS > 0: static width (rows of 2*MS bytes, 8-byte multiples); MS == 0: any width.
// A warp owns 32 paths.  After every step the 32 rows are written out cooperatively (consecutive lanes write
// consecutive 8-byte words of one row: whole 32-byte sectors instead of 32 scattered words per store).
template <int W, int MS, int TR>
__global__ void __launch_bounds__(TR) k_path_replay(const int8_t* roots, const uint8_t* actions, const int32_t* path_len, int64_t P,
                                                    int T, int m_rt, int cyclical, int8_t* states_out, uint8_t* lens_out,
                                                    int8_t* final_out, uint8_t* final_lens, int32_t* fail_step, uint8_t* status) {
    extern __shared__ __align__(16) uint8_t smem[];
    const int m = MS ? MS : m_rt;
    const int rowb = 2 * m, pitch = (rowb + 7) / 8 * 8;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int64_t p0 = ((int64_t)blockIdx.x * (TR / 32) + wid) * 32;  // first path of this warp
    if (p0 >= P) return;
    const int64_t p = p0 + lane;
    const bool valid = p < P;
    uint8_t* wrows = smem + (size_t)wid * 32 * pitch;
    uint8_t* row = wrows + (size_t)lane * pitch;
    const bool words = (rowb % 8 == 0) && ((((uintptr_t)roots | (uintptr_t)states_out | (uintptr_t)final_out) & 7u) == 0);
    int l0 = 0, l1 = 0;
    if (valid) {
        const uint8_t* src = reinterpret_cast<const uint8_t*>(roots) + p * rowb;
        for (int k = 0; k < rowb; ++k) {
            const uint8_t v = src[k];
            row[k] = v;
            if (v) (k < m ? l0 : l1) += 1;
        }
    }
    const int n = valid ? min(path_len ? path_len[p] : T, T) : 0;
    int nmax = n;
    for (int o = 16; o; o >>= 1) nmax = max(nmax, __shfl_xor_sync(0xFFFFFFFFu, nmax, o));
    uint32_t st = ACX_ROW_OK;
    int fail = -1;
    for (int t = 0; t < nmax; ++t) {
        bool emit = false;
        if (t < n && fail < 0) {
            const int a = actions[p * T + t];
            uint32_t s1 = ACX_ROW_OK;
            if (a >= 12) {
                s1 = ACX_ROW_INVALID_INPUT;
            } else {
                if constexpr (W <= 2) acmove_row64<MS>(row, m, a, cyclical != 0, l0, l1, s1);
                else acmove_row<W, MS>(row, row, m, a, cyclical != 0, l0, l1, s1);
            }
            if (s1 != ACX_ROW_OK) {
                st = s1;
                fail = t;
            } else {
                emit = true;
                if (lens_out) *reinterpret_cast<uchar2*>(lens_out + 2 * (p * T + t)) = make_uchar2((uint8_t)l0, (uint8_t)l1);
            }
        }
        if (states_out) {
            const unsigned em = __ballot_sync(0xFFFFFFFFu, emit);
            __syncwarp();
            uint8_t* base = reinterpret_cast<uint8_t*>(states_out);
            if (words) {
                const int cw = rowb / 8;
                for (int c = lane; c < 32 * cw; c += 32) {
                    const int r = c / cw, k = c - r * cw;
                    if ((em >> r) & 1u)
                        *reinterpret_cast<uint2*>(base + ((p0 + r) * T + t) * rowb + 8 * k) =
                            *reinterpret_cast<const uint2*>(wrows + (size_t)r * pitch + 8 * k);
                }
            } else if (emit) {
                uint8_t* dst = base + (p * T + t) * rowb;
                for (int k = 0; k < rowb; ++k) dst[k] = row[k];
            }
            __syncwarp();
        }
    }
    if (!valid) return;
    if (final_out) {
        uint8_t* dst = reinterpret_cast<uint8_t*>(final_out) + p * rowb;
        if (words) {
            for (int k = 0; k < rowb / 8; ++k) reinterpret_cast<uint2*>(dst)[k] = reinterpret_cast<const uint2*>(row)[k];
        } else {
            for (int k = 0; k < rowb; ++k) dst[k] = row[k];
        }
    }
    if (final_lens) {
        final_lens[2 * p] = (uint8_t)l0;
        final_lens[2 * p + 1] = (uint8_t)l1;
    }
    if (fail_step) fail_step[p] = fail;
    if (status) status[p] = (uint8_t)st;
}

template <int W, int MS, int TR>
static int launch_replay(const int8_t* roots, const uint8_t* actions, const int32_t* path_len, int64_t P, int T, int m, int cyclical,
                         int8_t* states_out, uint8_t* lens_out, int8_t* final_out, uint8_t* final_lens, int32_t* fail_step,
                         uint8_t* status, cudaStream_t st) {
    const size_t smem = (size_t)TR * ((2 * m + 7) / 8 * 8);
    auto kern = k_path_replay<W, MS, TR>;
    kern<<<(unsigned)((P + TR - 1) / TR), TR, smem, st>>>(roots, actions, path_len, P, T, m, cyclical, states_out, lens_out, final_out,
                                                        final_lens, fail_step, status);
    ACX_CUDA(cudaGetLastError());
    return ACX_OK;
}

// expand_path_with_cyclic_reductions for a batch of paths: one thread per path, row in shared memory.
template <int W, int MS, int TR>
__global__ void __launch_bounds__(TR) k_path_expand_cyclic(const int8_t* roots, const uint8_t* actions, const int32_t* path_len, int64_t P,
                                                           int T, int m, uint8_t* out_actions, uint16_t* out_totals, int cap,
                                                           int32_t* out_len, int32_t* fail_step, uint8_t* status) {
    extern __shared__ __align__(16) uint8_t smem[];
    const int rowb = 2 * m, pitch = (rowb + 7) / 8 * 8;
    const int64_t p = (int64_t)blockIdx.x * TR + threadIdx.x;
    if (p >= P) return;
    uint8_t* row = smem + (size_t)threadIdx.x * pitch;
    const uint8_t* src = reinterpret_cast<const uint8_t*>(roots) + p * rowb;
    for (int k = 0; k < rowb; ++k) row[k] = src[k];
    const int n = min(path_len ? path_len[p] : T, T);
    uint32_t st;
    int fail;
    const int o = expand_cyclic_path<W, MS>(row, m, actions + p * T, n, out_actions + p * cap, out_totals + p * cap, cap, st, fail);
    out_len[p] = o;
    fail_step[p] = fail;
    status[p] = (uint8_t)st;
}

template <int W, int MS>
static int launch_expand_cyclic(const int8_t* roots, const uint8_t* actions, const int32_t* path_len, int64_t P, int T, int m,
                                uint8_t* out_actions, uint16_t* out_totals, int cap, int32_t* out_len, int32_t* fail_step,
                                uint8_t* status, cudaStream_t st) {
    constexpr int TR = 64;
    const size_t smem = (size_t)TR * ((2 * m + 7) / 8 * 8);
    k_path_expand_cyclic<W, MS, TR><<<(unsigned)((P + TR - 1) / TR), TR, smem, st>>>(roots, actions, path_len, P, T, m, out_actions,
                                                                                     out_totals, cap, out_len, fail_step, status);
    ACX_CUDA(cudaGetLastError());
    return ACX_OK;
}

// The 14 handcrafted features of value_search/feature_extraction.py:11-91, one thread per row, float32 out.
// Counts run over the first len(r) letters of each relator (:50-61); the two ratios are computed in double and
// rounded once to float32, as numpy does when it builds the float32 array from Python floats (:68-91).
__global__ void k_rows_features(const int8_t* rows, int64_t n, int m, float* out) {
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const int8_t* r = rows + j * 2 * m;
    int len[2] = {0, 0};
    for (int k = 0; k < m; ++k) {
        len[0] += r[k] != 0;
        len[1] += r[m + k] != 0;
    }
    int cnt[2][4] = {{0, 0, 0, 0}, {0, 0, 0, 0}};  // x, x^-1, y, y^-1
    for (int i = 0; i < 2; ++i)
        for (int k = 0; k < len[i]; ++k) {
            const int v = r[i * m + k];
            cnt[i][0] += v == 1;
            cnt[i][1] += v == -1;
            cnt[i][2] += v == 2;
            cnt[i][3] += v == -2;
        }
    const int total = len[0] + len[1];
    const int mn = min(len[0], len[1]), mx = max(len[0], len[1]);
    float* o = out + j * 14;
    o[0] = (float)total;
    o[1] = (float)len[0];
    o[2] = (float)len[1];
    for (int i = 0; i < 2; ++i)
        for (int c = 0; c < 4; ++c) o[3 + 4 * i + c] = (float)cnt[i][c];
    o[11] = (float)((cnt[0][0] - cnt[0][1]) + (cnt[1][0] - cnt[1][1]));
    o[12] = total > 0 ? (float)((double)len[0] / (double)total) : 0.5f;
    o[13] = mn > 0 ? (float)((double)mx / (double)mn) : (float)mx;
}

}  // namespace acx

using namespace acx;

extern "C" int acx_rows_features(const int8_t* rows, int64_t n, int m, float* features_out, acx_stream_t stream) {
    if (n < 0 || m < 1 || m > ACX_MAX_RELATOR_LENGTH) return acx_fail(ACX_ERR_ARG, "acx_rows_features: need n >= 0, 1 <= m <= 128");
    if (n == 0) return ACX_OK;
    if (!rows || !features_out) return acx_fail(ACX_ERR_ARG, "acx_rows_features: null pointer");
    k_rows_features<<<(unsigned)((n + 127) / 128), 128, 0, (cudaStream_t)stream>>>(rows, n, m, features_out);
    ACX_CUDA(cudaGetLastError());
    return ACX_OK;
}


extern "C" int acx_path_replay(const int8_t* roots, const uint8_t* actions, const int32_t* path_len, int64_t n_paths, int max_steps,
                               int m, int cyclical, int8_t* states_out, uint8_t* lens_out, int8_t* final_out, uint8_t* final_lens,
                               int32_t* fail_step, uint8_t* status, acx_stream_t stream) {
    if (n_paths < 0 || max_steps < 0 || m < 1 || m > ACX_MAX_RELATOR_LENGTH) return acx_fail(ACX_ERR_ARG, "acx_path_replay: need n_paths >= 0, max_steps >= 0, 1 <= m <= 128");
    if (n_paths == 0) return ACX_OK;
    if (!roots || (max_steps > 0 && !actions)) return acx_fail(ACX_ERR_ARG, "acx_path_replay: null pointer");
    cudaStream_t st = (cudaStream_t)stream;
    if (m == 36) return launch_replay<2, 36, 64>(roots, actions, path_len, n_paths, max_steps, m, cyclical, states_out, lens_out, final_out, final_lens, fail_step, status, st);
    if (m <= 32) return launch_replay<1, 0, 64>(roots, actions, path_len, n_paths, max_steps, m, cyclical, states_out, lens_out, final_out, final_lens, fail_step, status, st);
    if (m <= 64) return launch_replay<2, 0, 64>(roots, actions, path_len, n_paths, max_steps, m, cyclical, states_out, lens_out, final_out, final_lens, fail_step, status, st);
    return launch_replay<4, 0, 64>(roots, actions, path_len, n_paths, max_steps, m, cyclical, states_out, lens_out, final_out, final_lens, fail_step, status, st);
}

extern "C" int acx_path_expand_cyclic(const int8_t* roots, const uint8_t* actions, const int32_t* path_len, int64_t n_paths,
                                      int max_steps, int m, uint8_t* out_actions, uint16_t* out_totals, int out_cap, int32_t* out_len,
                                      int32_t* fail_step, uint8_t* status, acx_stream_t stream) {
    if (n_paths < 0 || max_steps < 0 || out_cap < 0 || m < 1 || m > ACX_MAX_RELATOR_LENGTH)
        return acx_fail(ACX_ERR_ARG, "acx_path_expand_cyclic: need n_paths >= 0, max_steps >= 0, out_cap >= 0, 1 <= m <= 128");
    if (n_paths == 0) return ACX_OK;
    if (!roots || (max_steps > 0 && !actions) || (out_cap > 0 && (!out_actions || !out_totals)) || !out_len || !fail_step || !status)
        return acx_fail(ACX_ERR_ARG, "acx_path_expand_cyclic: null pointer");
    cudaStream_t st = (cudaStream_t)stream;
    if (m <= 32) return launch_expand_cyclic<1, 0>(roots, actions, path_len, n_paths, max_steps, m, out_actions, out_totals, out_cap, out_len, fail_step, status, st);
    if (m <= 64) return launch_expand_cyclic<2, 0>(roots, actions, path_len, n_paths, max_steps, m, out_actions, out_totals, out_cap, out_len, fail_step, status, st);
    return launch_expand_cyclic<4, 0>(roots, actions, path_len, n_paths, max_steps, m, out_actions, out_totals, out_cap, out_len, fail_step, status, st);
}
