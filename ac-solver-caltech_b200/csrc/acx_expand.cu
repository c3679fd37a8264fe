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
#include <string.h>

#include "../../include/acx.h"
#include "acx_common.h"
#include "acx_rows.cuh"

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
    if (m <= 32) return launch_expand_simple<1, 8>(p, st);
    if (m <= 64) return launch_expand_simple<2, 8>(p, st);
    return launch_expand_simple<4, 8>(p, st);
}
