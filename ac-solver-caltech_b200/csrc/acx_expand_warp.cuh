// acx_expand_warp.cuh -- tuned 12-move expansion: one THREAD per parent.
// The parent is packed into bit planes once; each of the 12 moves is applied to the planes held
// in registers; only the relator a move touches is unpacked, on top of a "base" child row (the
// parent with both relators simplified) that is copied shared->shared.  A warp stages GM moves
// of its 32 parents in shared memory (padded rows: conflict-free) and writes them back with
// coalesced vector stores: children[f][g*GM .. g*GM+GM) is one contiguous run per parent.
#pragma once
#include "acx_rows.cuh"

namespace acx {

template <int W, int MS>
ACX_HD void unpack_rel_static(uint8_t* rel, const Rel<W>& r) {
    constexpr int NW = MS / 4;
    const BV<W> N = bv_mask<W>(r.n);
    const BV<W> X = bv_xor<W>(r.S, r.T);
    uint32_t* p = reinterpret_cast<uint32_t*>(rel);
#pragma unroll
    for (int k = 0; k < NW; ++k) {
        const int g = k / 8, sh = 4 * (k % 8);
        const uint32_t nn = (N.w[g] >> sh) & 0xFu, ss = (r.S.w[g] >> sh) & 0xFu, xx = (X.w[g] >> sh) & 0xFu;
        p[k] = spread4(nn) + spread4(xx) + 253u * spread4(ss);
    }
}

// plane operations behind one interface: 64-bit planes (m <= 64) or W-word planes
template <int W, int MS>
struct PlaneOps {
    using R = Rel<W>;
    static __device__ __forceinline__ int classify(const uint8_t* row, R (&r)[2], int (&cnt)[2]) {
        RawRel<W> raw[2];
        uint32_t chk;
        pack_row_static<W, MS>(row, raw, chk);
        return classify_row<W>(raw, chk, r, cnt);
    }
    static __device__ __forceinline__ R inverse(const R& r) { return rel_inverse<W>(r); }
    static __device__ __forceinline__ bool concat(R& a, const R& b) { return rel_concat<W>(a, b, MS); }
    static __device__ __forceinline__ bool conjugate(R& a, uint32_t gs, uint32_t gt) { return rel_conjugate<W>(a, gs, gt, MS); }
    static __device__ __forceinline__ void cyclic(R& r) { rel_cyclic_reduce<W>(r); }
    static __device__ __forceinline__ void unpack(uint8_t* rel, const R& r) { unpack_rel_static<W, MS>(rel, r); }
    static __device__ __forceinline__ R select(bool second, const R& r0, const R& r1) { return rel_select<W>(second, r0, r1); }
};

template <int MS>
struct PlaneOps<2, MS> {
    using R = R64;
    static __device__ __forceinline__ int classify(const uint8_t* row, R (&r)[2], int (&cnt)[2]) {
        Raw64 raw[2];
        uint32_t chk;
        pack_row_static64<MS>(row, raw, chk);
        return classify_row64(raw, chk, r, cnt);
    }
    static __device__ __forceinline__ R inverse(const R& r) { return r64_inverse(r); }
    static __device__ __forceinline__ bool concat(R& a, const R& b) { return r64_concat(a, b, MS); }
    static __device__ __forceinline__ bool conjugate(R& a, uint32_t gs, uint32_t gt) { return r64_conjugate(a, gs, gt, MS); }
    static __device__ __forceinline__ void cyclic(R& r) { r64_cyclic_reduce(r); }
    static __device__ __forceinline__ void unpack(uint8_t* rel, const R& r) { unpack_rel_static64<MS>(rel, r); }
    static __device__ __forceinline__ R select(bool second, const R& r0, const R& r1) {
        R o;
        o.S = second ? r1.S : r0.S;
        o.T = second ? r1.T : r0.T;
        o.n = second ? r1.n : r0.n;
        return o;
    }
};

}  // namespace acx
