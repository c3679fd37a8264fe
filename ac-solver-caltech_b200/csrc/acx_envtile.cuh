// acx_envtile.cuh -- k_env_tiles: lockstep ACEnv.step over CERTIFIED rows (the bench kernel, K2).
//
// One CTA owns tiles of TR rows; a persistent grid walks the tiles through an S-stage ring in shared memory.
// EVERYTHING a tile needs moves with 1-D bulk async copies (TMA, UBLKCP): in = state rows (TR * 2m bytes), one
// 4-byte meta word per env (count_steps | len0 | len1) and the action bytes; out = state rows (rewritten in place
// in the stage), meta, the 8-byte step record, and optionally clipped reward / row status / the action-ring row.
// No thread issues a global load or store on the steady path, so HBM sees only full-line bulk transfers and no
// per-row latency is exposed (the round-1 kernel read steps / lengths / action with LDGs: +7 % time without any
// arithmetic, tools/copy_ceiling.cu).
//
// Rows of a tile are SORTED BY MOVE CLASS before they are processed (ballots + one permutation byte per row in
// shared memory): concatenations first, conjugations after.  A warp then runs one class only -- conjugation on
// the bytes (acx_envrow.cuh, ~50 instructions) or concatenation on bit planes (~350) -- instead of every warp
// executing both under divergence (round 1: 16 of 32 lanes active per issued instruction).
//
// Launch chaining: every launch marks a parity counter per CTA (odd while it runs).  With p.chain_wait the
// kernel does not wait for the whole previous grid (griddepcontrol.wait); CTA b waits until CTA b of the
// previous launch on the same buffers has finished (acquire / release at gpu scope).  Tile t always belongs to CTA t % grid, so this is
// the exact row-level dependency between two consecutive steps, and the tail of step k overlaps the head of
// step k+1.  Valid only when every other input of the step (actions, pool, next_idx) was complete before the
// previous env step was launched -- the caller asserts that by passing the counters (acx.h: acx_env_io.chain).
#pragma once
#include "acx_envrow.cuh"
#include "acx_tma.cuh"

namespace acx {

struct EnvParams {
    int8_t* state;      // rows are read from here ...
    int8_t* state_out;  // ... and written here (== state: in place)
    uint32_t* meta;
    const uint8_t* actions;
    uint8_t* step_out;
    float* reward_clipped;
    uint8_t* status;
    const int8_t* pool;
    const int32_t* next_idx;
    uint8_t* ring_row;  // this step's row of the action ring, or null
    int32_t* fin_count;
    int32_t* fin_env;
    int32_t* fin_steps;
    int8_t* fin_state;
    int32_t fin_cap;
    int32_t* err_count;
    uint32_t* chain;  // per-CTA parity counters (odd = a launch is running on that CTA's tiles), or null
    int chain_wait;   // 1: wait on the counters instead of the whole previous grid (see the header comment)
    int64_t n;
    int64_t row0;
    int m;
    int horizon;
    float clip_lo, clip_hi;
};

struct EnvOut {
    uint32_t meta, rew, hi;
    float clip;
};

// ac_env.py:97-111 on a row that already sits (post-move) in shared memory, fused with the reset of finished envs
// (ac_env.py:113-129 via agents/training.py:351-352) and the finished-episode side buffer.
__device__ __forceinline__ EnvOut env_finish(const EnvParams& p, int64_t row, uint8_t* srow, int m, int l0, int l1,
                                             uint32_t status, uint32_t steps_in) {
    const int total = l0 + l1;
    bool done = false;
    if (total == 2) {  // is_presentation_trivial (utils.py:57-87): |letters| == {1,2}
        int a = (int8_t)srow[0], b = (int8_t)srow[m];
        a = a < 0 ? -a : a;
        b = b < 0 ? -b : b;
        done = (a == 1 && b == 2) || (a == 2 && b == 1);
    }
    const int32_t max_reward = p.horizon * m * 2;    // ac_env.py:76
    const int32_t rew = done ? max_reward : -total;  // ac_env.py:100
    uint32_t steps = steps_in + 1u;                  // ac_env.py:102
    const bool trunc = (int)steps >= p.horizon;      // ac_env.py:103
    if (done || trunc) {
        const bool words = ((2 * m) & 3) == 0;
        if (p.fin_count) {  // final observation + episode length of the finished env (gymnasium: final_observation)
            const int32_t k = atomicAdd(p.fin_count, 1);
            if (k < p.fin_cap) {
                p.fin_env[k] = (int32_t)row;
                p.fin_steps[k] = (int32_t)steps;
                int8_t* dst = p.fin_state + (int64_t)k * 2 * m;
                if (words && (((uintptr_t)dst | (uintptr_t)srow) & 3u) == 0) {
                    for (int q = 0; q < m / 2; ++q) reinterpret_cast<uint32_t*>(dst)[q] = reinterpret_cast<const uint32_t*>(srow)[q];
                } else {
                    for (int q = 0; q < 2 * m; ++q) dst[q] = (int8_t)srow[q];
                }
            }
        }
        if (p.pool) {
            const int8_t* src = p.pool + (int64_t)p.next_idx[row] * 2 * m;
            int c0 = 0, c1 = 0;
            if (words && (((uintptr_t)src | (uintptr_t)srow) & 3u) == 0) {
                const int hw = m / 2;  // words per row = 2m / 4
                for (int q = 0; q < hw; ++q) {
                    const uint32_t v = __ldg(reinterpret_cast<const uint32_t*>(src) + q);
                    reinterpret_cast<uint32_t*>(srow)[q] = v;
                }
                for (int q = 0; q < m; ++q) {
                    c0 += srow[q] != 0;
                    c1 += srow[m + q] != 0;
                }
            } else {
                for (int q = 0; q < m; ++q) {
                    const int8_t v0 = src[q], v1 = src[m + q];
                    srow[q] = (uint8_t)v0;
                    srow[m + q] = (uint8_t)v1;
                    c0 += v0 != 0;
                    c1 += v1 != 0;
                }
            }
            l0 = c0;
            l1 = c1;
            steps = 0u;
        }
    }
    if (status != 0u && p.err_count) atomicAdd(p.err_count, 1);
    EnvOut o;
    o.meta = steps | ((uint32_t)l0 << 16) | ((uint32_t)l1 << 24);
    o.rew = (uint32_t)rew;
    o.hi = (done ? 1u : 0u) | (trunc ? 0x100u : 0u) | ((uint32_t)l0 << 16) | ((uint32_t)l1 << 24);
    o.clip = fminf(fmaxf((float)rew, p.clip_lo), p.clip_hi);
    return o;
}

__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_release_gpu_add(uint32_t* p, uint32_t v) {
    asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }

template <int MS, int TR>
struct EnvTile {
    static constexpr uint32_t ROWB = 2 * MS;
    static constexpr uint32_t STATE = TR * ROWB, META = TR * 4, ACT = TR, REC = TR * 8, CLIP = TR * 4, STAT = TR;
    static constexpr uint32_t OFF_META = STATE, OFF_ACT = OFF_META + META, OFF_REC = OFF_ACT + ACT, OFF_CLIP = OFF_REC + REC,
                              OFF_STAT = OFF_CLIP + CLIP, STAGE = OFF_STAT + STAT;
    static constexpr uint32_t PAD = 16;  // the byte path reads the word in front of a relator
    static_assert(STATE % 16 == 0 && TR % 16 == 0, "bulk copies need 16-byte multiples");
    static constexpr size_t smem_bytes(int S) { return PAD + (size_t)S * STAGE + TR + 8 * S; }
};

template <int MS, int TR, int S>
__global__ void __launch_bounds__(TR) k_env_tiles(const EnvParams p, int64_t ntiles) {
    using L = EnvTile<MS, TR>;
    constexpr int NWARP = TR / 32;
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t* const stages = smem + L::PAD;
    uint8_t* const perm = stages + (size_t)S * L::STAGE;
    uint64_t* const bars = reinterpret_cast<uint64_t*>(perm + TR);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t stage0 = smem_u32(stages), bar0 = smem_u32(bars);

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < S; ++s) mbar_init(bar0 + 8 * s, 1);
        mbar_fence_init();
    }
    __syncthreads();
    if (!p.chain_wait) asm volatile("griddepcontrol.wait;" ::: "memory");
    if (p.chain) {
        if (tid == 0) {
            // row-level dependency on the previous launch: its CTA of the same index owned the same tiles
            if (p.chain_wait)
                while (ld_acquire_gpu(p.chain + blockIdx.x) & 1u) __nanosleep(32);
            atomicAdd(p.chain + blockIdx.x, 1u);  // odd: this CTA is running
            __threadfence();
            fence_proxy_async_all();
        }
        __syncthreads();  // the counter is odd before ANY thread of this CTA lets the next launch start
    }
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

    const int64_t first = blockIdx.x, stride = gridDim.x;
    const int64_t tile0 = p.row0 / TR;  // p.row0 is a multiple of TR
    auto issue = [&](uint32_t st, int64_t t) {
        const uint32_t base = stage0 + st * L::STAGE, bar = bar0 + 8 * st;
        const int64_t g = tile0 + t;
        mbar_expect_tx(bar, L::STATE + L::META + L::ACT);
        bulk_load(base, p.state + g * L::STATE, L::STATE, bar);
        bulk_load(base + L::OFF_META, reinterpret_cast<const uint8_t*>(p.meta) + g * L::META, L::META, bar);
        bulk_load(base + L::OFF_ACT, p.actions + g * L::ACT, L::ACT, bar);
    };
    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < S; ++s) {
            const int64_t t = first + (int64_t)s * stride;
            if (t < ntiles) issue(s, t);
        }
    }

    uint32_t s = 0, parity = 0, sprev = S - 1;
    bool first_it = true;
    const unsigned ltmask = (1u << lane) - 1u;
    for (int64_t t = first; t < ntiles; t += stride) {
        mbar_wait(bar0 + 8 * s, parity);
        uint8_t* const st = stages + (size_t)s * L::STAGE;
        const uint8_t* const acts = st + L::OFF_ACT;

        // ---- sort the rows of the tile by move class: concatenations (ids 0..3) first ----
        const bool cat = acts[tid] < 4;
        const unsigned mine = __ballot_sync(0xFFFFFFFFu, cat);
        int before = 0, ncat = 0;
#pragma unroll
        for (int g = 0; g < NWARP; ++g) {
            const unsigned b = (g == warp) ? mine : __ballot_sync(0xFFFFFFFFu, acts[32 * g + lane] < 4);
            const int c = __popc(b);
            before += (g < warp) ? c : 0;
            ncat += c;
        }
        const int rk = __popc(mine & ltmask);
        const int slot = cat ? (before + rk) : (ncat + (tid - before - rk));
        perm[slot] = (uint8_t)tid;
        __syncthreads();
        const int r = perm[tid];

        // ---- one row ----
        uint8_t* const srow = st + (uint32_t)r * L::ROWB;
        uint32_t* const metas = reinterpret_cast<uint32_t*>(st + L::OFF_META);
        const uint32_t meta = metas[r];
        int l0 = (int)((meta >> 16) & 0xFFu), l1 = (int)(meta >> 24);
        const uint32_t status = cert_row_step<MS>(srow, acts[r], l0, l1);
        const int64_t row = p.row0 + t * TR + r;
        const EnvOut o = env_finish(p, row, srow, MS, l0, l1, status, meta & 0xFFFFu);
        metas[r] = o.meta;
        reinterpret_cast<uint2*>(st + L::OFF_REC)[r] = make_uint2(o.rew, o.hi);
        if (p.reward_clipped) reinterpret_cast<float*>(st + L::OFF_CLIP)[r] = o.clip;
        if (p.status) st[L::OFF_STAT + r] = (uint8_t)status;

        fence_proxy_async();
        __syncthreads();
        if (tid == 0) {
            const uint32_t base = stage0 + s * L::STAGE;
            const int64_t g = tile0 + t;
            bulk_store(p.state_out + g * L::STATE, base, L::STATE);
            bulk_store(reinterpret_cast<uint8_t*>(p.meta) + g * L::META, base + L::OFF_META, L::META);
            bulk_store(p.step_out + g * L::REC, base + L::OFF_REC, L::REC);
            if (p.reward_clipped) bulk_store(reinterpret_cast<uint8_t*>(p.reward_clipped) + g * L::CLIP, base + L::OFF_CLIP, L::CLIP);
            if (p.status) bulk_store(p.status + g * L::STAT, base + L::OFF_STAT, L::STAT);
            if (p.ring_row) bulk_store(p.ring_row + g * L::ACT, base + L::OFF_ACT, L::ACT);
            bulk_commit();
            if (!first_it) {  // refill the stage whose stores were issued one iteration ago
                bulk_wait_read<1>();
                const int64_t tn = t + (int64_t)(S - 1) * stride;
                if (tn < ntiles) issue(sprev, tn);
            }
        }
        first_it = false;
        sprev = s;
        if (++s == S) {
            s = 0;
            parity ^= 1u;
        }
    }
    if (tid == 0) {
        bulk_wait_all<0>();
        if (p.chain) {
            fence_proxy_async_all();
            red_release_gpu_add(p.chain + blockIdx.x, 1u);  // even: done, the rows of this CTA are in memory
        }
    }
}

}  // namespace acx
