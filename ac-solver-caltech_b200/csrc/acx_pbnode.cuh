// acx_pbnode.cuh -- the 32-byte bit-plane node / candidate record of the plane bfs (acx_pbfs.cu, acx_pbfs2.cu):
//
//     w0 = S(r0) | len(r0) << 40      w1 = T(r0) | len(r1) << 40      w2 = S(r1)      w3 = T(r1)
//
// a candidate record is the same 32 bytes with its discovery order (48 bits) and row status (2 bits) in the unused top
// bits of w2 / w3.  One 256-bit access per node / record (LDG.256 / STG.256): a whole DRAM sector per lane.
#pragma once
#include <stdint.h>

#include "acx_rows.cuh"

namespace acx {

static constexpr int PB_MAX_WORLD = 32;
static constexpr int PB_MAX_M = 36;
static constexpr uint64_t PB_PLANE = (1ull << 36) - 1;  // plane bits
static constexpr uint64_t PB_W01 = (1ull << 46) - 1;    // plane + length field
static constexpr uint64_t PB_PROV = 1ull << 47;
static constexpr uint64_t PB_VAL = PB_PROV - 1;

struct PWords {
    uint64_t w[4];
};

// one 256-bit access per node / record (LDG.256 / STG.256 on sm_100): a whole 32-byte sector per lane
__device__ __forceinline__ PWords pb_load(const uint64_t* p) {
    PWords o;
    asm volatile("ld.global.v4.b64 {%0, %1, %2, %3}, [%4];" : "=l"(o.w[0]), "=l"(o.w[1]), "=l"(o.w[2]), "=l"(o.w[3]) : "l"(p) : "memory");
    return o;
}

// the same load served by L2: for data another kernel may be writing while this one runs (an L1 line could be stale)
__device__ __forceinline__ PWords pb_load_cg(const uint64_t* p) {
    PWords o;
    asm volatile("ld.global.cg.v4.b64 {%0, %1, %2, %3}, [%4];" : "=l"(o.w[0]), "=l"(o.w[1]), "=l"(o.w[2]), "=l"(o.w[3]) : "l"(p) : "memory");
    return o;
}

__device__ __forceinline__ void pb_store(uint64_t* p, const PWords& v) {
    asm volatile("st.global.v4.b64 [%0], {%1, %2, %3, %4};" ::"l"(p), "l"(v.w[0]), "l"(v.w[1]), "l"(v.w[2]), "l"(v.w[3]) : "memory");
}

__device__ __forceinline__ PWords pb_state(const PWords& v) {  // drop order / status
    PWords o;
    o.w[0] = v.w[0] & PB_W01;
    o.w[1] = v.w[1] & PB_W01;
    o.w[2] = v.w[2] & PB_PLANE;
    o.w[3] = v.w[3] & PB_PLANE;
    return o;
}

__device__ __forceinline__ int64_t pb_order(const PWords& v) {
    return (int64_t)(((v.w[2] >> 36) & 0xFFFFFFFull) | (((v.w[3] >> 36) & 0xFFFFFull) << 28));
}
__device__ __forceinline__ uint32_t pb_status(const PWords& v) { return (uint32_t)(v.w[3] >> 56) & 3u; }
__device__ __forceinline__ int pb_total(const PWords& v) { return (int)((v.w[0] >> 40) & 63u) + (int)((v.w[1] >> 40) & 63u); }

__device__ __forceinline__ PWords pb_pack(const R64& r0, const R64& r1) {
    PWords o;
    o.w[0] = r0.S | ((uint64_t)r0.n << 40);
    o.w[1] = r0.T | ((uint64_t)r1.n << 40);
    o.w[2] = r1.S;
    o.w[3] = r1.T;
    return o;
}

__device__ __forceinline__ uint64_t pb_hash(const PWords& s) {
    uint64_t h = 0x9E3779B97F4A7C15ull;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        h ^= s.w[k];
        h *= 0xFF51AFD7ED558CCDull;
        h ^= h >> 32;
    }
    h *= 0xC4CEB9FE1A85EC53ull;
    h ^= h >> 29;
    return h;
}

__device__ __forceinline__ bool pb_equal(const PWords& a, const PWords& b) {
    return ((a.w[0] ^ b.w[0]) | (a.w[1] ^ b.w[1]) | (a.w[2] ^ b.w[2]) | (a.w[3] ^ b.w[3])) == 0ull;
}

// 31 bits of the hash modulo the world size, as a 32-bit remainder (a 64-bit one is a subroutine call per candidate)
__device__ __forceinline__ int pb_owner(uint64_t h, int world) {
    return world == 1 ? 0 : (int)((uint32_t)((h >> 20) & 0x7FFFFFFFull) % (uint32_t)world);
}

}  // namespace acx
