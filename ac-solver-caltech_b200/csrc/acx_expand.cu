// acx_expand.cu -- all 12 successors of every parent in one launch (search frontier expansion).
//
// children[f][a] = ACMove(a, parents[f])  for a = 0..11  (greedy.py:76-83, breadth_first.py:69-76)
//
//   k_expand_simple  any width: one thread per (parent, move), parents and children staged in
//                    shared memory, cooperative copies in and out.
//   k_expand_warp    static widths (36, 128): one thread per PARENT; the parent is packed into bit
//                    planes once and the 12 moves are applied to the planes held in registers;
//                    children are staged per warp and written back with coalesced vector stores.
//                    Width 36 also keeps the simplified parent ROW in registers (18 words): a conjugation
//                    child is that row with one relator shifted by a byte (conj_row_words, acx_envrow.cuh),
//                    a concatenation child swaps in the unpacked relator, and either goes to the stage as
//                    nine 8-byte stores -- no shared-memory loads per child, one lens / status store per
//                    group of four moves, 16-byte stores of the children to global memory.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/acx.h"
#include "acx_common.h"
#include "acx_rows.cuh"
#include "acx_envrow.cuh"
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

// Shared-memory layout of one warp of k_expand_warp.  REG (64-bit planes, width 36): the simplified parent row lives
// in registers, no base-row copy in shared memory.
template <int W, int MS, int GM>
struct ExpandTile {
    static constexpr int ROWB = 2 * MS;
    static constexpr int VEC = (ROWB % 16 == 0) ? 16 : 8;
    static constexpr bool REG = (W <= 2) && (VEC == 8);
    static constexpr int PSTRIDE = (ROWB % 128 == 0) ? ROWB + 16 : ROWB;  // padded against bank conflicts
    static constexpr int CSTRIDE = GM * ROWB + VEC;
    static constexpr int WARP_BYTES = 32 * ((REG ? 1 : 2) * PSTRIDE + CSTRIDE);
    static constexpr int PCH = ROWB / VEC, CCH = GM * ROWB / VEC;
    static constexpr bool WIDE_OUT = (VEC == 8) && ((GM * ROWB) % 16 == 0);  // 16-byte global stores of the children
};

// one thread per parent; see acx_expand_warp.cuh
template <int W, int MS, int GM, int WARPS>
__global__ void __launch_bounds__(32 * WARPS) k_expand_warp(const ExpandParams p, int64_t ntiles) {
    using Ops = PlaneOps<W, MS>;
    using R = typename Ops::R;
    using L = ExpandTile<W, MS, GM>;
    constexpr int ROWB = L::ROWB, VEC = L::VEC, PSTRIDE = L::PSTRIDE, CSTRIDE = L::CSTRIDE, PCH = L::PCH, CCH = L::CCH;
    constexpr int NW = MS / 4;
    using V = typename VecT<VEC>::T;
    extern __shared__ __align__(128) uint8_t smem[];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint8_t* par = smem + (size_t)wid * L::WARP_BYTES;
    uint8_t* base = par + 32 * PSTRIDE;                      // !REG only
    uint8_t* cbuf = par + (L::REG ? 1 : 2) * 32 * PSTRIDE;
    const bool cyc = p.cyclical != 0;
    const int64_t warp0 = (int64_t)blockIdx.x * WARPS + wid, nwarps = (int64_t)gridDim.x * WARPS;
    const uint8_t* gpar = reinterpret_cast<const uint8_t*>(p.parents);
    uint8_t* gch = reinterpret_cast<uint8_t*>(p.children);
    // one lens / status store per move group when the group is 4 or 2 moves and the arrays are aligned for it
    constexpr bool META_VEC = (GM == 4 || GM == 2);
    const bool meta_vec = META_VEC && (((uintptr_t)p.lens & (uintptr_t)(2 * GM - 1)) == 0) && (((uintptr_t)p.status & (uintptr_t)(GM - 1)) == 0);

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
        uint32_t w[L::REG ? 2 * NW : 1];  // REG: the base row = the parent with both relators simplified (utils.py:270-281)
        if (fast) {
            rc[0] = r[0];
            rc[1] = r[1];
            if (cyc) {
                Ops::cyclic(rc[0]);
                Ops::cyclic(rc[1]);
            }
            if constexpr (L::REG) {
#pragma unroll
                for (int k = 0; k < NW; ++k) {
                    const uint2 v = reinterpret_cast<const uint2*>(prow)[k];
                    w[2 * k] = v.x;
                    w[2 * k + 1] = v.y;
                }
                if (cyc) {
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        if (rc[h].n != r[h].n) {
                            uint32_t u[NW];
                            unpack_words64<MS>(u, rc[h]);
#pragma unroll
                            for (int k = 0; k < NW; ++k) w[h * NW + k] = u[k];
                        }
                    }
                }
            } else {
                for (int k = 0; k < PCH; ++k) reinterpret_cast<V*>(brow)[k] = reinterpret_cast<const V*>(prow)[k];
                if (cyc) {
                    if (rc[0].n != r[0].n) Ops::unpack(brow, rc[0]);
                    if (rc[1].n != r[1].n) Ops::unpack(brow + MS, rc[1]);
                }
            }
        }
#pragma unroll 1
        for (int g = 0; g < 12 / GM; ++g) {
            uint64_t lens_acc = 0ull;
            uint32_t stat_acc = 0u;
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
                    if constexpr (L::REG) {
                        uint32_t o[2 * NW];
                        int fix_pos = -1;
                        uint32_t fix_val = 0u;
                        bool parent_row = false;
                        // a conjugation never needs the planes when the search does not reduce cyclically, or when the
                        // relator is cyclically reduced already (then the reduced conjugate is a rotation)
                        const R64 xr = Ops::select(i, r[0], r[1]);
                        const int nb = i ? rc[1].n : rc[0].n;
                        if (is_conj && (!cyc || nb == xr.n)) {
                            const uint32_t gs = inv ? 1u : 0u, gt = (d >> 3) & 1u;
                            const uint32_t s0 = (uint32_t)xr.S & 1u, t0 = (uint32_t)xr.T & 1u;
                            const uint32_t sl = (uint32_t)shr64(xr.S, xr.n - 1) & 1u, tl = (uint32_t)shr64(xr.T, xr.n - 1) & 1u;
                            const bool sc = (t0 == gt) && (s0 != gs), ec = (tl == gt) && (sl == gs);  // ac_moves.py:119-120
                            int na;
                            if (cyc) na = conj_row_words<MS, true>(w, o, i, xr.n, sc, ec, d, fix_pos, fix_val);
                            else na = conj_row_words<MS, false>(w, o, i, xr.n, sc, ec, d, fix_pos, fix_val);
                            l0 = i ? rc[0].n : na;
                            l1 = i ? na : rc[1].n;
                        } else {
                            R64 x = xr;
                            bool changed;
                            if (!is_conj) {
                                R64 b = Ops::select(!i, r[0], r[1]);
                                if (inv) b = Ops::inverse(b);
                                changed = Ops::concat(x, b);
                            } else {
                                changed = Ops::conjugate(x, inv ? 1u : 0u, (d >> 3) & 1u);
                            }
                            if (x.n == 0) {  // utils.py:264-266 raises: the child slot holds the parent unchanged
                                st = ACX_ROW_EMPTY_RELATOR;
                                parent_row = true;
                            } else {
                                uint32_t u[NW];
                                if (changed) {
                                    if (cyc) Ops::cyclic(x);
                                    unpack_words64<MS>(u, x);
                                }
#pragma unroll
                                for (int k = 0; k < NW; ++k) {
                                    o[k] = (changed && !i) ? u[k] : w[k];
                                    o[NW + k] = (changed && i) ? u[k] : w[NW + k];
                                }
                                const int na = changed ? x.n : nb;
                                l0 = i ? rc[0].n : na;
                                l1 = i ? na : rc[1].n;
                            }
                        }
                        if (parent_row) {
                            for (int k = 0; k < PCH; ++k) reinterpret_cast<V*>(child)[k] = reinterpret_cast<const V*>(prow)[k];
                        } else {
#pragma unroll
                            for (int k = 0; k < NW; ++k) reinterpret_cast<uint2*>(child)[k] = make_uint2(o[2 * k], o[2 * k + 1]);
                            if (fix_pos >= 0) child[fix_pos] = (uint8_t)fix_val;
                        }
                    } else {
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
                    }
                } else {  // rows outside the plane domain: the checked row function on a copy of the parent
                    for (int k = 0; k < PCH; ++k) reinterpret_cast<V*>(child)[k] = reinterpret_cast<const V*>(prow)[k];
                    if constexpr (W <= 2) acmove_row64<MS>(child, MS, a, cyc, l0, l1, st);
                    else acmove_row<W, MS>(child, child, MS, a, cyc, l0, l1, st);
                }
                if (meta_vec) {
                    lens_acc |= (uint64_t)((uint32_t)(l0 & 0xFF) | ((uint32_t)(l1 & 0xFF) << 8)) << (16 * j);
                    stat_acc |= (st & 0xFFu) << (8 * j);
                } else {
                    const int64_t c = (f0 + lane) * 12 + a;
                    if (p.lens) *reinterpret_cast<uchar2*>(p.lens + 2 * c) = make_uchar2((uint8_t)l0, (uint8_t)l1);
                    if (p.status) p.status[c] = (uint8_t)st;
                }
            }
            if (META_VEC && meta_vec && active) {
                const int64_t c = (f0 + lane) * 12 + (int64_t)g * GM;
                if constexpr (GM == 4) {
                    if (p.lens) *reinterpret_cast<uint2*>(p.lens + 2 * c) = make_uint2((uint32_t)lens_acc, (uint32_t)(lens_acc >> 32));
                    if (p.status) *reinterpret_cast<uint32_t*>(p.status + c) = stat_acc;
                } else {
                    if (p.lens) *reinterpret_cast<uint32_t*>(p.lens + 2 * c) = (uint32_t)lens_acc;
                    if (p.status) *reinterpret_cast<uint16_t*>(p.status + c) = (uint16_t)stat_acc;
                }
            }
            __syncwarp();
            uint8_t* gdst = gch + (f0 * 12 + (int64_t)g * GM) * ROWB;
            if constexpr (L::WIDE_OUT) {  // two 8-byte shared loads feed one 16-byte global store
                constexpr int C16 = GM * ROWB / 16;
                for (int c = lane; c < nrows * C16; c += 32) {
                    const int pr = c / C16, o = c - pr * C16;
                    const uint2* src = reinterpret_cast<const uint2*>(cbuf + pr * CSTRIDE + o * 16);
                    const uint2 lo = src[0], hi = src[1];
                    *reinterpret_cast<uint4*>(gdst + (size_t)pr * 12 * ROWB + o * 16) = make_uint4(lo.x, lo.y, hi.x, hi.y);
                }
            } else {
                for (int c = lane; c < nrows * CCH; c += 32) {
                    const int pr = c / CCH, o = c - pr * CCH;
                    *reinterpret_cast<V*>(gdst + (size_t)pr * 12 * ROWB + o * VEC) =
                        *reinterpret_cast<const V*>(cbuf + pr * CSTRIDE + o * VEC);
                }
            }
            __syncwarp();
        }
    }
}

template <int W, int MS, int GM, int WARPS>
static int launch_expand_warp(const ExpandParams& p, cudaStream_t st) {
    constexpr size_t smem = (size_t)WARPS * ExpandTile<W, MS, GM>::WARP_BYTES;
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
            return e ? atoi(e) : 44;
        }();
        switch (cfg) {
            case 24: return launch_expand_warp<2, 36, 2, 4>(p, st);
            case 28: return launch_expand_warp<2, 36, 2, 8>(p, st);
            case 42: return launch_expand_warp<2, 36, 4, 2>(p, st);
            case 64: return launch_expand_warp<2, 36, 6, 4>(p, st);
            case 34: return launch_expand_warp<2, 36, 3, 4>(p, st);
            case 48: return launch_expand_warp<2, 36, 4, 8>(p, st);
            default: return launch_expand_warp<2, 36, 4, 4>(p, st);
        }
    }
    if (aligned && m == 128) return launch_expand_warp<4, 128, 1, 4>(p, st);
    if (m <= 32) return launch_expand_simple<1, 8>(p, st);
    if (m <= 64) return launch_expand_simple<2, 8>(p, st);
    return launch_expand_simple<4, 8>(p, st);
}
