// acx_moves.cu -- lockstep ACMove / ACEnv.step kernels for sm_100a and their C ABI.
//
// Data layout in HBM: the reference's own -- rows of 2*m int8, densely packed ([N, 2m]).
// One THREAD owns one presentation.  A CTA of TR threads owns a tile of TR rows:
//   k_rows_tma    persistent CTAs; tiles move global<->shared with 1-D bulk async copies
//                 (cp.async.bulk, mbarrier completion) through an S-stage ring, so per-row
//                 shared-memory reads/writes are 64-bit and conflict-free while every HBM
//                 transaction is a full-line bulk transfer.  Rows are rewritten IN PLACE in
//                 the stage buffer and stored back from it.
//   k_rows_simple same row code, plain cooperative copies; used for the tail rows, unaligned
//                 pointers and widths without a static instantiation.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/acx.h"
#include "acx_common.h"
#include "acx_envtile.cuh"

namespace acx {

struct StepParams {
    const int8_t* in;  // [n, 2m]
    int8_t* out;       // [n, 2m] (may alias in)
    const uint8_t* actions;
    uint8_t* lens;    // [n,2] or null
    uint8_t* status;  // [n] or null
    int64_t n;
    int64_t row0;  // first row handled by this launch
    int m;
    int cyclical;
    EnvParams env;  // ACEnv.step extras (ENV kernels)
};

// m <= 64 takes the 64-bit plane path, wider rows the generic multi-word planes
template <int W, int MS>
__device__ __forceinline__ void row_move(uint8_t* srow, int m, int move, bool cyc, int& l0, int& l1, uint32_t& status) {
    if constexpr (W <= 2) acmove_row64<MS>(srow, m, move, cyc, l0, l1, status);
    else acmove_row<W, MS>(srow, srow, m, move, cyc, l0, l1, status);
}

// per-row outputs of the checked kernels (the certified tile kernel stages them in shared memory instead)
template <bool ENV>
__device__ __forceinline__ void row_epilogue(const StepParams& p, int64_t row, uint8_t* srow, int m, int l0, int l1,
                                             uint32_t status, uint32_t meta_in, int move) {
    if (ENV) {
        const EnvParams& e = p.env;
        const EnvOut o = env_finish(e, row, srow, m, l0, l1, status, meta_in & 0xFFFFu);
        e.meta[row] = o.meta;
        reinterpret_cast<uint2*>(e.step_out)[row] = make_uint2(o.rew, o.hi);
        if (e.reward_clipped) e.reward_clipped[row] = o.clip;
        if (e.ring_row) e.ring_row[row] = (uint8_t)move;
    } else {
        if (p.lens) *reinterpret_cast<uchar2*>(p.lens + 2 * row) = make_uchar2((uint8_t)l0, (uint8_t)l1);
    }
    if (p.status) p.status[row] = (uint8_t)status;
}

// ---------------------------------------------------------------------------------------------
template <int W, int MS, int TR, int S, bool ENV>
__global__ void __launch_bounds__(TR) k_rows_tma(const StepParams p, int64_t ntiles) {
    constexpr int ROWB = 2 * MS;
    constexpr uint32_t TILEB = (uint32_t)TR * ROWB;
    static_assert(TILEB % 16 == 0, "bulk copies need 16-byte multiples");
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)S * TILEB);
    const int tid = threadIdx.x;
    const uint32_t smem0 = smem_u32(smem);
    const uint32_t bar0 = smem_u32(bars);

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < S; ++s) mbar_init(bar0 + 8 * s, 1);
        mbar_fence_init();
    }
    __syncthreads();
    // Programmatic dependent launch: the CTAs of the next step may become resident and run the prologue above while
    // this step's last CTAs drain; nothing of the previous step is read or written before the wait below.
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

    const int64_t first = blockIdx.x, stride = gridDim.x;
    const int8_t* gin = p.in + p.row0 * ROWB;
    int8_t* gout = p.out + p.row0 * ROWB;
    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < S; ++s) {
            const int64_t t = first + (int64_t)s * stride;
            if (t < ntiles) {
                mbar_expect_tx(bar0 + 8 * s, TILEB);
                bulk_load(smem0 + s * TILEB, gin + t * TILEB, TILEB, bar0 + 8 * s);
            }
        }
    }

    uint32_t s = 0, parity = 0, sprev = S - 1;  // stage of this iteration / its mbarrier phase / previous stage
    bool first_it = true;
    uint8_t* const my_rows = smem + (uint32_t)tid * ROWB;
    for (int64_t t = first; t < ntiles; t += stride) {
        const int64_t row = p.row0 + t * TR + tid;
        const int move = p.actions[row];
        uint32_t meta_in = 0;
        if (ENV) meta_in = p.env.meta[row];
        int l0 = 0, l1 = 0;
        mbar_wait(bar0 + 8 * s, parity);

        uint8_t* srow = my_rows + s * TILEB;
        uint32_t status;
        row_move<W, MS>(srow, MS, move, ENV ? true : (p.cyclical != 0), l0, l1, status);
        row_epilogue<ENV>(p, row, srow, MS, l0, l1, status, meta_in, move);

        fence_proxy_async();
        __syncthreads();
        if (tid == 0) {
            bulk_store(gout + t * TILEB, smem0 + s * TILEB, TILEB);
            bulk_commit();
            if (!first_it) {  // refill the stage whose store was issued one iteration ago
                bulk_wait_read<1>();
                const int64_t tn = t + (int64_t)(S - 1) * stride;
                if (tn < ntiles) {
                    mbar_expect_tx(bar0 + 8 * sprev, TILEB);
                    bulk_load(smem0 + sprev * TILEB, gin + tn * TILEB, TILEB, bar0 + 8 * sprev);
                }
            }
        }
        first_it = false;
        sprev = s;
        if (++s == S) {
            s = 0;
            parity ^= 1u;
        }
    }
    if (tid == 0) bulk_wait_all<0>();
}

// ---------------------------------------------------------------------------------------------
// cooperative copy of `bytes` bytes, any alignment
__device__ __forceinline__ void coop_copy(uint8_t* dst, const uint8_t* src, int bytes, int tid, int nthreads) {
    if ((((uintptr_t)dst | (uintptr_t)src | (uintptr_t)bytes) & 3u) == 0) {
        const uint32_t* s4 = reinterpret_cast<const uint32_t*>(src);
        uint32_t* d4 = reinterpret_cast<uint32_t*>(dst);
        for (int k = tid; k < bytes / 4; k += nthreads) d4[k] = s4[k];
    } else {
        for (int k = tid; k < bytes; k += nthreads) dst[k] = src[k];
    }
}

template <int W, int MS, int TR, bool ENV>
__global__ void __launch_bounds__(TR) k_rows_simple(const StepParams p, int64_t nrows) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int m = MS ? MS : p.m;
    const int rowb = 2 * m;
    const int tid = threadIdx.x;
    const int64_t r0 = (int64_t)blockIdx.x * TR;
    const int rows = (int)min((int64_t)TR, nrows - r0);
    const uint8_t* gin = reinterpret_cast<const uint8_t*>(p.in) + (p.row0 + r0) * rowb;
    uint8_t* gout = reinterpret_cast<uint8_t*>(p.out) + (p.row0 + r0) * rowb;
    coop_copy(smem, gin, rows * rowb, tid, TR);
    __syncthreads();
    if (tid < rows) {
        const int64_t row = p.row0 + r0 + tid;
        const int move = p.actions[row];
        uint32_t meta_in = 0;
        if (ENV) meta_in = p.env.meta[row];
        uint8_t* srow = smem + (size_t)tid * rowb;
        int l0, l1;
        uint32_t status;
        row_move<W, MS>(srow, m, move, ENV ? true : (p.cyclical != 0), l0, l1, status);
        row_epilogue<ENV>(p, row, srow, m, l0, l1, status, meta_in, move);
    }
    __syncthreads();
    coop_copy(gout, smem, rows * rowb, tid, TR);
}

// scalar-API helper: one thread per row straight on global memory, byte path only
__global__ void k_rows_primitive(const int8_t* in, const uint8_t* moves, int8_t* out, uint8_t* lens, uint8_t* status,
                                 int64_t n, int m, int mode) {
    const int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= n) return;
    int l0, l1;
    uint32_t st;
    slow_acmove(in + row * 2 * m, out + row * 2 * m, m, moves ? moves[row] : 0, mode, l0, l1, st);
    if (lens) {
        lens[2 * row] = (uint8_t)l0;
        lens[2 * row + 1] = (uint8_t)l1;
    }
    if (status) status[row] = (uint8_t)st;
}

// certification of rows (see acx_rows.cuh): one thread per row on global memory
__global__ void k_rows_certify(const int8_t* rows, int64_t n, int m, uint8_t* lens, uint8_t* cert) {
    const int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= n) return;
    int l0 = 0, l1 = 0;
    bool ok = false;
    const uint8_t* r = reinterpret_cast<const uint8_t*>(rows) + row * 2 * m;
    if (m <= 64) {
        ok = row_is_certified(r, m, l0, l1);
    } else {
        for (int k = 0; k < m; ++k) {
            l0 += r[k] != 0;
            l1 += r[m + k] != 0;
        }
    }
    if (lens) {
        lens[2 * row] = (uint8_t)l0;
        lens[2 * row + 1] = (uint8_t)l1;
    }
    if (cert) cert[row] = ok ? 1 : 0;
}

// ---------------------------------------------------------------------------------------------
static int current_device() {
    int dev = 0;
    cudaGetDevice(&dev);
    return (dev >= 0 && dev < ACX_MAX_DEVICES) ? dev : 0;
}

static int sm_count() {
    static int cached[ACX_MAX_DEVICES] = {0};
    const int dev = current_device();
    if (cached[dev] == 0) {
        int n = 0;
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        cached[dev] = n > 0 ? n : 148;
    }
    return cached[dev];
}

static int pdl_enabled() {
    static const int pdl = [] {
        const char* e = getenv("ACX_PDL");
        return e ? atoi(e) : 1;
    }();
    return pdl;
}

template <typename Kern, typename Params>
static int launch_pdl(Kern kern, int64_t grid, int block, size_t smem, cudaStream_t st, const Params& p, int64_t ntiles) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(block);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    ACX_CUDA(cudaLaunchKernelEx(&cfg, kern, p, ntiles));
    return ACX_OK;
}

template <int W, int MS, int TR, int S, bool ENV>
static int launch_tma(const StepParams& p, int64_t ntiles, cudaStream_t st) {
    constexpr size_t smem = (size_t)S * TR * 2 * MS + 8 * S;
    auto kern = k_rows_tma<W, MS, TR, S, ENV>;
    static int ctas_per_sm[ACX_MAX_DEVICES] = {0};  // function attributes are per device
    const int dev = current_device();
    if (ctas_per_sm[dev] == 0) {
        ACX_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        int occ = 0;
        ACX_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, TR, smem));
        ctas_per_sm[dev] = occ > 0 ? occ : 1;
    }
    const int64_t grid = std::min<int64_t>(ntiles, (int64_t)sm_count() * ctas_per_sm[dev]);
    return launch_pdl(kern, grid, TR, smem, st, p, ntiles);
}

// certified env rows (m = 36): the class-sorted tile kernel of acx_envtile.cuh
template <int MS, int TR, int S>
static int launch_env_tiles(EnvParams e, int64_t ntiles, bool whole_batch, cudaStream_t st) {
    constexpr size_t smem = EnvTile<MS, TR>::smem_bytes(S);
    auto kern = k_env_tiles<MS, TR, S>;
    static int ctas_per_sm[ACX_MAX_DEVICES] = {0};
    const int dev = current_device();
    if (ctas_per_sm[dev] == 0) {
        ACX_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        int occ = 0;
        ACX_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, TR, smem));
        ctas_per_sm[dev] = occ > 0 ? occ : 1;
    }
    // every CTA gets the same number of tiles (+-1): with 8192 tiles on 1332 resident CTAs, 15 % of the CTAs would
    // otherwise run a 7th tile alone at the end of the launch
    const int64_t resident = (int64_t)sm_count() * ctas_per_sm[dev];
    const int64_t per_cta = (ntiles + resident - 1) / resident;
    const int64_t grid = (ntiles + per_cta - 1) / per_cta;
    // the per-CTA dependency replaces the grid-wide one only when this launch is the whole step
    if (grid > ACX_ENV_CHAIN_SLOTS) e.chain = nullptr;
    if (!e.chain || !whole_batch || !pdl_enabled()) e.chain_wait = 0;
    return launch_pdl(kern, grid, TR, smem, st, e, ntiles);
}

template <int W, int MS, int TR, bool ENV>
static int launch_simple(const StepParams& p, int64_t nrows, cudaStream_t st) {
    const int m = MS ? MS : p.m;
    const size_t smem = (size_t)TR * 2 * m + 16;
    auto kern = k_rows_simple<W, MS, TR, ENV>;
    static bool configured[ACX_MAX_DEVICES] = {false};
    const int dev = current_device();
    if (!configured[dev] && smem > 48 * 1024) {
        ACX_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(TR * 2 * 128 + 16)));
        configured[dev] = true;
    }
    const int64_t grid = (nrows + TR - 1) / TR;
    kern<<<(unsigned)grid, TR, smem, st>>>(p, nrows);
    ACX_CUDA(cudaGetLastError());
    return ACX_OK;
}

static int env_tile_rows() {  // ACX_ENV_TILE=128|256 (tuning knob)
    static int v = 0;
    if (v == 0) {
        const char* e = getenv("ACX_ENV_TILE");
        v = e ? atoi(e) : 128;
        if (v != 256) v = 128;
    }
    return v;
}

static int env_stages() {  // ACX_ENV_STAGES=2|3|4 (tuning knob)
    static int v = 0;
    if (v == 0) {
        const char* e = getenv("ACX_ENV_STAGES");
        v = e ? atoi(e) : 3;
        if (v < 2 || v > 4) v = 3;
    }
    return v;
}

static bool env_tiles_usable(const EnvParams& e) {
    uintptr_t a = (uintptr_t)e.state | (uintptr_t)e.state_out | (uintptr_t)e.meta | (uintptr_t)e.actions | (uintptr_t)e.step_out |
                  (uintptr_t)e.reward_clipped | (uintptr_t)e.status | (uintptr_t)e.ring_row;
    return (a & 15u) == 0;
}

// rows [0, n): full tiles through a bulk-copy kernel when a static instantiation exists and the
// buffers are 16-byte aligned; everything else through the simple kernel.
template <int W, int MS, int TR, int S, bool ENV>
static int run_static(StepParams p, bool certified, cudaStream_t st) {
    const bool aligned = (((uintptr_t)p.in | (uintptr_t)p.out) & 15u) == 0;
    int64_t done_rows = 0;
    if constexpr (ENV && MS == 36) {
        if (certified && env_tiles_usable(p.env)) {
            const int tr = env_tile_rows(), stg = env_stages();
            if (p.n >= tr) {
                const int64_t ntiles = p.n / tr;
                const bool whole = ntiles * tr == p.n;
                EnvParams e = p.env;
                e.row0 = 0;
                int rc;
                if (tr == 256) rc = stg == 2 ? launch_env_tiles<36, 256, 2>(e, ntiles, whole, st) : launch_env_tiles<36, 256, 3>(e, ntiles, whole, st);
                else rc = stg == 2 ? launch_env_tiles<36, 128, 2>(e, ntiles, whole, st)
                        : stg == 4 ? launch_env_tiles<36, 128, 4>(e, ntiles, whole, st) : launch_env_tiles<36, 128, 3>(e, ntiles, whole, st);
                if (rc) return rc;
                done_rows = ntiles * tr;
            }
        }
    }
    if (done_rows == 0 && aligned && p.n >= TR) {
        const int64_t ntiles = p.n / TR;
        p.row0 = 0;
        if (int rc = launch_tma<W, MS, TR, S, ENV>(p, ntiles, st)) return rc;
        done_rows = ntiles * TR;
    }
    if (done_rows < p.n) {
        p.row0 = done_rows;
        return launch_simple<W, MS, TR, ENV>(p, p.n - done_rows, st);
    }
    return ACX_OK;
}

static int tma_stages() {  // ACX_STAGES=3|4 (tuning knob; default 3)
    static int v = 0;
    if (v == 0) {
        const char* e = getenv("ACX_STAGES");
        v = (e && e[0] == '4') ? 4 : 3;
    }
    return v;
}

static int tile_rows() {  // ACX_TILE=64|128|256 (tuning knob; default 128)
    static int v = 0;
    if (v == 0) {
        const char* e = getenv("ACX_TILE");
        v = e ? atoi(e) : 128;
        if (v != 64 && v != 256) v = 128;
    }
    return v;
}

template <bool ENV>
static int run_rows(StepParams p, bool certified, cudaStream_t st) {
    if (p.n == 0) return ACX_OK;
    p.row0 = 0;
    switch (p.m) {
        case 36:
            if (tile_rows() == 256) return run_static<2, 36, 256, 3, ENV>(p, certified, st);
            if (tile_rows() == 64) return run_static<2, 36, 64, 4, ENV>(p, certified, st);
            return tma_stages() == 3 ? run_static<2, 36, 128, 3, ENV>(p, certified, st) : run_static<2, 36, 128, 4, ENV>(p, certified, st);
        case 128: return run_static<4, 128, 64, 4, ENV>(p, certified, st);
        default: break;
    }
    if (p.m <= 32) return launch_simple<1, 0, 128, ENV>(p, p.n, st);
    if (p.m <= 64) return launch_simple<2, 0, 128, ENV>(p, p.n, st);
    return launch_simple<4, 0, 64, ENV>(p, p.n, st);
}

__global__ void k_env_meta_init(const int8_t* state, const int32_t* steps, uint32_t* meta, int64_t n, int m) {
    const int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= n) return;
    const int8_t* r = state + row * 2 * m;
    uint32_t l0 = 0, l1 = 0;
    for (int k = 0; k < m; ++k) {
        l0 += r[k] != 0;
        l1 += r[m + k] != 0;
    }
    meta[row] = ((steps ? (uint32_t)steps[row] : 0u) & 0xFFFFu) | (l0 << 16) | (l1 << 24);
}

}  // namespace acx

using namespace acx;

extern "C" int acx_move_batch(const int8_t* states_in, const uint8_t* moves, int8_t* states_out, uint8_t* lens_out,
                              uint8_t* row_status, int64_t n, int m, int cyclical, acx_stream_t stream) {
    if (n < 0 || m < 1 || m > ACX_MAX_RELATOR_LENGTH) return acx_fail(ACX_ERR_ARG, "acx_move_batch: need n >= 0 and 1 <= m <= 128");
    if (n > 0 && (!states_in || !moves || !states_out)) return acx_fail(ACX_ERR_ARG, "acx_move_batch: null pointer");
    StepParams p;
    memset(&p, 0, sizeof p);
    p.in = states_in;
    p.out = states_out;
    p.actions = moves;
    p.lens = lens_out;
    p.status = row_status;
    p.n = n;
    p.m = m;
    p.cyclical = cyclical;
    return run_rows<false>(p, false, (cudaStream_t)stream);
}

extern "C" int acx_env_step(const acx_env_io* io, acx_stream_t stream) {
    if (!io) return acx_fail(ACX_ERR_ARG, "acx_env_step: null io");
    const int64_t n = io->n;
    const int m = io->m;
    if (n < 0 || m < 1 || m > ACX_MAX_RELATOR_LENGTH || io->horizon < 1 || io->horizon > 65535)
        return acx_fail(ACX_ERR_ARG, "acx_env_step: need n >= 0, 1 <= m <= 128, 1 <= horizon <= 65535");
    if (n > 0 && (!io->state || !io->meta || !io->actions || !io->step_out)) return acx_fail(ACX_ERR_ARG, "acx_env_step: null pointer");
    if (((uintptr_t)io->step_out & 7u) != 0 || ((uintptr_t)io->meta & 3u) != 0)
        return acx_fail(ACX_ERR_ARG, "acx_env_step: step_out must be 8-byte aligned, meta 4-byte aligned");
    if ((io->pool == nullptr) != (io->next_idx == nullptr)) return acx_fail(ACX_ERR_ARG, "acx_env_step: pool and next_idx go together");
    if (io->action_ring && io->ring_len < 1) return acx_fail(ACX_ERR_ARG, "acx_env_step: action_ring needs ring_len >= 1");
    if (io->fin_count && (!io->fin_env || !io->fin_steps || !io->fin_state || io->fin_cap < 0))
        return acx_fail(ACX_ERR_ARG, "acx_env_step: fin_count needs fin_env, fin_steps, fin_state and fin_cap >= 0");
    if (n > 0x7FFFFFFFll && io->fin_count) return acx_fail(ACX_ERR_ARG, "acx_env_step: fin_env holds 32-bit env indices");
    StepParams p;
    memset(&p, 0, sizeof p);
    p.in = io->state;
    p.out = io->state_out ? io->state_out : io->state;
    p.actions = io->actions;
    p.status = io->row_status;
    p.n = n;
    p.m = m;
    p.cyclical = 1;
    EnvParams& e = p.env;
    e.state = io->state;
    e.state_out = io->state_out ? io->state_out : io->state;
    e.meta = io->meta;
    e.actions = io->actions;
    e.step_out = io->step_out;
    e.reward_clipped = io->reward_clipped;
    e.status = io->row_status;
    e.pool = io->pool;
    e.next_idx = io->next_idx;
    if (io->action_ring) {
        int64_t slot = io->t % io->ring_len;
        if (slot < 0) slot += io->ring_len;
        e.ring_row = io->action_ring + slot * n;
    }
    e.fin_count = io->fin_count;
    e.fin_env = io->fin_env;
    e.fin_steps = io->fin_steps;
    e.fin_state = io->fin_state;
    e.fin_cap = io->fin_cap;
    e.err_count = io->err_count;
    e.chain = io->chain;
    e.chain_wait = (io->chain && (io->flags & ACX_ENV_CHAINED)) ? 1 : 0;
    e.n = n;
    e.m = m;
    e.horizon = io->horizon;
    e.clip_lo = io->clip_lo;
    e.clip_hi = io->clip_hi;
    return run_rows<true>(p, (io->flags & ACX_ENV_CERTIFIED) != 0, (cudaStream_t)stream);
}

extern "C" int acx_env_meta_init(const int8_t* state, const int32_t* steps, uint32_t* meta, int64_t n, int m, acx_stream_t stream) {
    if (n < 0 || m < 1 || m > ACX_MAX_RELATOR_LENGTH) return acx_fail(ACX_ERR_ARG, "acx_env_meta_init: need n >= 0, 1 <= m <= 128");
    if (n == 0) return ACX_OK;
    if (!state || !meta) return acx_fail(ACX_ERR_ARG, "acx_env_meta_init: null pointer");
    k_env_meta_init<<<(unsigned)((n + 127) / 128), 128, 0, (cudaStream_t)stream>>>(state, steps, meta, n, m);
    ACX_CUDA(cudaGetLastError());
    return ACX_OK;
}

extern "C" int acx_rows_primitive(const int8_t* states_in, const uint8_t* moves, int8_t* states_out, uint8_t* lens_out,
                                  uint8_t* row_status, int64_t n, int m, int mode, acx_stream_t stream) {
    if (n < 0 || m < 1 || m > 4096) return acx_fail(ACX_ERR_ARG, "acx_rows_primitive: need n >= 0 and 1 <= m <= 4096");
    if (n > 0 && (!states_in || !states_out)) return acx_fail(ACX_ERR_ARG, "acx_rows_primitive: null pointer");
    if (!(mode & ACX_MODE_NO_MOVE) && n > 0 && !moves) return acx_fail(ACX_ERR_ARG, "acx_rows_primitive: moves missing");
    if (n == 0) return ACX_OK;
    k_rows_primitive<<<(unsigned)((n + 63) / 64), 64, 0, (cudaStream_t)stream>>>(states_in, moves, states_out, lens_out,
                                                                              row_status, n, m, mode);
    ACX_CUDA(cudaGetLastError());
    return ACX_OK;
}

extern "C" int acx_rows_certify(const int8_t* rows, int64_t n, int m, uint8_t* lens_out, uint8_t* certified_out,
                                acx_stream_t stream) {
    if (n < 0 || m < 1 || m > ACX_MAX_RELATOR_LENGTH) return acx_fail(ACX_ERR_ARG, "acx_rows_certify: need n >= 0, 1 <= m <= 128");
    if (n == 0) return ACX_OK;
    if (!rows) return acx_fail(ACX_ERR_ARG, "acx_rows_certify: null pointer");
    k_rows_certify<<<(unsigned)((n + 127) / 128), 128, 0, (cudaStream_t)stream>>>(rows, n, m, lens_out, certified_out);
    ACX_CUDA(cudaGetLastError());
    return ACX_OK;
}
