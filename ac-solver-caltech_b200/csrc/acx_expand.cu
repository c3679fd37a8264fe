// acx_expand.cu -- all 12 successors of every parent in one launch (search frontier expansion).
//
// children[f][a] = ACMove(a, parents[f])  for a = 0..11  (greedy.py:76-83, breadth_first.py:69-76)
//
//   k_expand_simple  any width: one thread per (parent, move), parents and children staged in
//                    shared memory, cooperative copies in and out.
//   k_expand_warp    static widths (36, 128): one thread per PARENT; the parent is packed into bit
//                    planes once and the 12 moves are applied to the planes held in registers;
//                    children are staged per warp and written back with coalesced vector stores.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/acx.h"
#include "acx_common.h"
#include "acx_rows.cuh"
#include "acx_expand_warp.cuh"

namespace acx {

struct ExpandParams {
    const int8_t* parents;  // [f, 2m]
    int8_t* children;       // [f, 12, 2m]
    uint8_t* lens;          // [f, 12, 2] or null
    uint8_t* status;        // [f, 12] or null
    int64_t f;
    int m;
    int cyclical;
};

__device__ __forceinline__ void coop_copy_bytes(uint8_t* dst, const uint8_t* src, int64_t bytes, int tid, int nthreads) {
    if ((((uintptr_t)dst | (uintptr_t)src | (uintptr_t)bytes) & 15u) == 0) {
        const uint4* s = reinterpret_cast<const uint4*>(src);
        uint4* d = reinterpret_cast<uint4*>(dst);
        for (int64_t k = tid; k < bytes / 16; k += nthreads) d[k] = s[k];
    } else if ((((uintptr_t)dst | (uintptr_t)src | (uintptr_t)bytes) & 3u) == 0) {
        const uint32_t* s = reinterpret_cast<const uint32_t*>(src);
        uint32_t* d = reinterpret_cast<uint32_t*>(dst);
        for (int64_t k = tid; k < bytes / 4; k += nthreads) d[k] = s[k];
    } else {
        for (int64_t k = tid; k < bytes; k += nthreads) dst[k] = src[k];
    }
}

// PB parents per block, 12*PB threads
template <int W, int PB>
__global__ void __launch_bounds__(12 * PB) k_expand_simple(const ExpandParams p) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int m = p.m, rowb = 2 * m;
    const int tid = threadIdx.x;
    const int64_t f0 = (int64_t)blockIdx.x * PB;
    const int np = (int)min((int64_t)PB, p.f - f0);
    // children first (16-byte aligned at offset 0), parents after them
    uint8_t* sc = smem;
    uint8_t* sp = smem + (((size_t)12 * PB * rowb + 15) & ~(size_t)15);
    coop_copy_bytes(sp, reinterpret_cast<const uint8_t*>(p.parents) + f0 * rowb, (int64_t)np * rowb, tid, 12 * PB);
    __syncthreads();
    const int pi = tid / 12, a = tid - 12 * pi;
    if (pi < np) {
        int l0, l1;
        uint32_t st;
        uint8_t* child = sc + (size_t)tid * rowb;
        if constexpr (W <= 2) {  // 64-bit planes work in place: seed the child slot with the parent
            const uint8_t* par = sp + (size_t)pi * rowb;
            for (int k = 0; k < rowb; ++k) child[k] = par[k];
            acmove_row64<0>(child, m, a, p.cyclical != 0, l0, l1, st);
        } else {
            acmove_row<W, 0>(sp + (size_t)pi * rowb, child, m, a, p.cyclical != 0, l0, l1, st);
        }
        const int64_t c = (f0 + pi) * 12 + a;
        if (p.lens) {
            p.lens[2 * c] = (uint8_t)l0;
            p.lens[2 * c + 1] = (uint8_t)l1;
        }
        if (p.status) p.status[c] = (uint8_t)st;
    }
    __syncthreads();
    coop_copy_bytes(reinterpret_cast<uint8_t*>(p.children) + f0 * 12 * rowb, sc, (int64_t)np * 12 * rowb, tid, 12 * PB);
}

template <int W, int PB>
static int launch_expand_simple(const ExpandParams& p, cudaStream_t st) {
    const size_t smem = (((size_t)12 * PB * 2 * p.m + 15) & ~(size_t)15) + (size_t)PB * 2 * p.m + 16;
    auto kern = k_expand_simple<W, PB>;
    static bool configured[ACX_MAX_DEVICES] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= ACX_MAX_DEVICES) dev = 0;
    if (!configured[dev]) {
        ACX_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)(13 * PB * 2 * ACX_MAX_RELATOR_LENGTH + 64)));
        configured[dev] = true;
    }
    const int64_t grid = (p.f + PB - 1) / PB;
    kern<<<(unsigned)grid, 12 * PB, smem, st>>>(p);
    ACX_CUDA(cudaGetLastError());
    return ACX_OK;
}

template <int VEC>
struct VecT;
template <>
struct VecT<8> {
    using T = uint2;
};
template <>
struct VecT<16> {
    using T = uint4;
};

// one thread per parent; see acx_expand_warp.cuh
template <int W, int MS, int GM, int WARPS>
__global__ void __launch_bounds__(32 * WARPS) k_expand_warp(const ExpandParams p, int64_t ntiles) {
    using Ops = PlaneOps<W, MS>;
    using R = typename Ops::R;
    constexpr int ROWB = 2 * MS;
    constexpr int VEC = (ROWB % 16 == 0) ? 16 : 8;
    using V = typename VecT<VEC>::T;
    constexpr int PSTRIDE = (ROWB % 128 == 0) ? ROWB + 16 : ROWB;  // padded against bank conflicts
    constexpr int CSTRIDE = GM * ROWB + VEC;
    constexpr int WARP_BYTES = 32 * (2 * PSTRIDE + CSTRIDE);
    constexpr int PCH = ROWB / VEC, CCH = GM * ROWB / VEC;
    extern __shared__ __align__(128) uint8_t smem[];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint8_t* par = smem + (size_t)wid * WARP_BYTES;
    uint8_t* base = par + 32 * PSTRIDE;
    uint8_t* cbuf = base + 32 * PSTRIDE;
    const bool cyc = p.cyclical != 0;
    const int64_t warp0 = (int64_t)blockIdx.x * WARPS + wid, nwarps = (int64_t)gridDim.x * WARPS;
    const uint8_t* gpar = reinterpret_cast<const uint8_t*>(p.parents);
    uint8_t* gch = reinterpret_cast<uint8_t*>(p.children);

    for (int64_t tile = warp0; tile < ntiles; tile += nwarps) {
        const int64_t f0 = tile * 32;
        const int nrows = (int)min((int64_t)32, p.f - f0);
        for (int c = lane; c < nrows * PCH; c += 32) {
            const int pr = c / PCH, o = c - pr * PCH;
            *reinterpret_cast<V*>(par + pr * PSTRIDE + o * VEC) = *reinterpret_cast<const V*>(gpar + f0 * ROWB + (size_t)c * VEC);
        }
        __syncwarp();
        const bool active = lane < nrows;
        uint8_t* prow = par + lane * PSTRIDE;
        uint8_t* brow = base + lane * PSTRIDE;
        R r[2], rc[2];
        int cnt[2] = {0, 0};
        int cls = ROW_INVALID;
        if (active) cls = Ops::classify(prow, r, cnt);
        const bool fast = active && cls == ROW_FAST;
        if (fast) {  // base row = the parent with both relators simplified (utils.py:270-281)
            for (int k = 0; k < PCH; ++k) reinterpret_cast<V*>(brow)[k] = reinterpret_cast<const V*>(prow)[k];
            rc[0] = r[0];
            rc[1] = r[1];
            if (cyc) {
                Ops::cyclic(rc[0]);
                Ops::cyclic(rc[1]);
                if (rc[0].n != r[0].n) Ops::unpack(brow, rc[0]);
                if (rc[1].n != r[1].n) Ops::unpack(brow + MS, rc[1]);
            }
        }
#pragma unroll 1
        for (int g = 0; g < 12 / GM; ++g) {
#pragma unroll 1
            for (int j = 0; j < GM; ++j) {
                const int a = g * GM + j;
                uint8_t* child = cbuf + lane * CSTRIDE + j * ROWB;
                if (!active) continue;
                int l0 = cnt[0], l1 = cnt[1];
                uint32_t st = ACX_ROW_OK;
                if (fast) {
                    const uint32_t d = decode_move(a);
                    const bool i = d & 1u, is_conj = d & 2u, inv = d & 4u;
                    R x = Ops::select(i, r[0], r[1]);
                    bool changed;
                    if (!is_conj) {
                        R b = Ops::select(!i, r[0], r[1]);
                        if (inv) b = Ops::inverse(b);
                        changed = Ops::concat(x, b);
                    } else {
                        changed = Ops::conjugate(x, inv ? 1u : 0u, (d >> 3) & 1u);
                    }
                    if (x.n == 0) {  // utils.py:264-266 raises: the child slot holds the parent unchanged
                        st = ACX_ROW_EMPTY_RELATOR;
                        for (int k = 0; k < PCH; ++k) reinterpret_cast<V*>(child)[k] = reinterpret_cast<const V*>(prow)[k];
                    } else {
                        for (int k = 0; k < PCH; ++k) reinterpret_cast<V*>(child)[k] = reinterpret_cast<const V*>(brow)[k];
                        if (changed) {
                            if (cyc) Ops::cyclic(x);
                            Ops::unpack(child + (d & 1u) * MS, x);
                        }
                        const int na = changed ? x.n : (i ? rc[1].n : rc[0].n);
                        l0 = i ? rc[0].n : na;
                        l1 = i ? na : rc[1].n;
                    }
                } else {  // rows outside the plane domain: the checked row function on a copy of the parent
                    for (int k = 0; k < PCH; ++k) reinterpret_cast<V*>(child)[k] = reinterpret_cast<const V*>(prow)[k];
                    if constexpr (W <= 2) acmove_row64<MS>(child, MS, a, cyc, l0, l1, st);
                    else acmove_row<W, MS>(child, child, MS, a, cyc, l0, l1, st);
                }
                const int64_t c = (f0 + lane) * 12 + a;
                if (p.lens) *reinterpret_cast<uchar2*>(p.lens + 2 * c) = make_uchar2((uint8_t)l0, (uint8_t)l1);
                if (p.status) p.status[c] = (uint8_t)st;
            }
            __syncwarp();
            uint8_t* gdst = gch + (f0 * 12 + (int64_t)g * GM) * ROWB;
            for (int c = lane; c < nrows * CCH; c += 32) {
                const int pr = c / CCH, o = c - pr * CCH;
                *reinterpret_cast<V*>(gdst + (size_t)pr * 12 * ROWB + o * VEC) =
                    *reinterpret_cast<const V*>(cbuf + pr * CSTRIDE + o * VEC);
            }
            __syncwarp();
        }
    }
}

template <int W, int MS, int GM, int WARPS>
static int launch_expand_warp(const ExpandParams& p, cudaStream_t st) {
    constexpr int ROWB = 2 * MS;
    constexpr int VEC = (ROWB % 16 == 0) ? 16 : 8;
    constexpr int PSTRIDE = (ROWB % 128 == 0) ? ROWB + 16 : ROWB;
    constexpr int CSTRIDE = GM * ROWB + VEC;
    constexpr size_t smem = (size_t)WARPS * 32 * (2 * PSTRIDE + CSTRIDE);
    auto kern = k_expand_warp<W, MS, GM, WARPS>;
    static int ctas_per_sm[ACX_MAX_DEVICES] = {0};
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= ACX_MAX_DEVICES) dev = 0;
    if (ctas_per_sm[dev] == 0) {
        ACX_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        int occ = 0;
        ACX_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, 32 * WARPS, smem));
        ctas_per_sm[dev] = occ > 0 ? occ : 1;
    }
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int64_t ntiles = (p.f + 31) / 32;
    const int64_t grid = std::min<int64_t>((ntiles + WARPS - 1) / WARPS, (int64_t)sms * ctas_per_sm[dev]);
    kern<<<(unsigned)grid, 32 * WARPS, smem, st>>>(p, ntiles);
    ACX_CUDA(cudaGetLastError());
    return ACX_OK;
}

}  // namespace acx

using namespace acx;

extern "C" int acx_expand12(const int8_t* parents, int8_t* children, uint8_t* child_lens, uint8_t* child_status,
                            int64_t f, int m, int cyclical, acx_stream_t stream) {
    if (f < 0 || m < 1 || m > ACX_MAX_RELATOR_LENGTH) return acx_fail(ACX_ERR_ARG, "acx_expand12: need f >= 0 and 1 <= m <= 128");
    if (f > 0 && (!parents || !children)) return acx_fail(ACX_ERR_ARG, "acx_expand12: null pointer");
    if (f == 0) return ACX_OK;
    ExpandParams p;
    p.parents = parents;
    p.children = children;
    p.lens = child_lens;
    p.status = child_status;
    p.f = f;
    p.m = m;
    p.cyclical = cyclical;
    cudaStream_t st = (cudaStream_t)stream;
    const bool aligned = (((uintptr_t)parents | (uintptr_t)children) & 15u) == 0;
    if (aligned && m == 36) {
        static const int cfg = [] {  // tuning knob: ACX_EXPAND36 = GM*10 + WARPS
            const char* e = getenv("ACX_EXPAND36");
            return e ? atoi(e) : 48;
        }();
        switch (cfg) {
            case 24: return launch_expand_warp<2, 36, 2, 4>(p, st);
            case 28: return launch_expand_warp<2, 36, 2, 8>(p, st);
            case 42: return launch_expand_warp<2, 36, 4, 2>(p, st);
            case 64: return launch_expand_warp<2, 36, 6, 4>(p, st);
            case 34: return launch_expand_warp<2, 36, 3, 4>(p, st);
            case 44: return launch_expand_warp<2, 36, 4, 4>(p, st);
            default: return launch_expand_warp<2, 36, 4, 8>(p, st);
        }
    }
    if (aligned && m == 128) return launch_expand_warp<4, 128, 1, 4>(p, st);
    if (m <= 32) return launch_expand_simple<1, 8>(p, st);
    if (m <= 64) return launch_expand_simple<2, 8>(p, st);
    return launch_expand_simple<4, 8>(p, st);
}
