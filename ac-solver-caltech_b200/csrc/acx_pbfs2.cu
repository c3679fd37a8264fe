// acx_pbfs2.cu -- the device-driven round of the hash-partitioned plane bfs (driver: acx/pbfs.py).
//
// Same search semantics as acx_pbfs.cu (breadth_first.py:52-95: FIFO discovery order, success test before the visited
// test, budget test after each popped node), but a round never comes back to the host:
//
//   k_pbfs_expand2   one thread per parent, a CTA of 256 parents.  The children of four moves at a time are SORTED BY
//                    OWNER RANK in shared memory and leave the CTA as contiguous runs, one per owner: 16-byte stores of
//                    consecutive threads, i.e. whole 128-byte lines over NVLink instead of one scattered 32-byte record
//                    per lane (round 1: 380 GB/s of the 900 GB/s egress).  A run is padded to four records with null
//                    records so that every run starts on a line.  The child that undoes the move which discovered the
//                    parent is its grandparent (cyclical = 0: the 12 moves are pairwise inverse) and is not sent.
//   k_pbfs_probe2    visited-set probe with the winner resolution FUSED in: a candidate sets its bit in the per-parent
//                    mask BEFORE it claims a slot (and clears it if the claim fails); a candidate of lower order that
//                    replaces a claim clears the loser's bit.  No separate resolve pass over the records and the table.
//   k_pbfs_decide    after the masks are all-reduced: exclusive scan of the new-node counts, then ONE thread replays the
//                    reference's decision for the round (first success / error by order, the budget cut of
//                    breadth_first.py:91-95) into a control block in device memory.
//   k_pbfs_commit2   winners become nodes unless the control block says the search stopped in this round.
// Every kernel returns at once when the control block holds a stop code, so the host enqueues whole levels without
// reading anything back and looks at the control block once per level.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/acx.h"
#include "acx_common.h"
#include "acx_pbnode.cuh"

#include <cub/device/device_scan.cuh>
#include <cub/iterator/transform_input_iterator.cuh>

namespace acx {

// control block (int64 words), identical on every rank
enum { CTL_STOP = 0, CTL_NODES = 1, CTL_EXPANDED = 2, CTL_GID_BASE = 3, CTL_POS = 4, CTL_PARENT_BASE = 5, CTL_OVERFLOW = 6, CTL_ROUNDS = 7,
       CTL_STAGED = 8, CTL_STAGED_LOCAL = 9 };
enum { STOP_NONE = 0, STOP_SOLVED = 1, STOP_ERROR = 2, STOP_BUDGET = 3 };
static constexpr unsigned long long PB_SENT = 1ull << 62;
static constexpr uint64_t PB_IDX = (1ull << 46) - 1;  // candidate index inside a provisional entry

struct PbDst2 {
    uint8_t* region[PB_MAX_WORLD];
};

static constexpr int EX_THREADS = 256, EX_GM = 4;
static constexpr int EX_SLOTS = EX_THREADS * EX_GM + 4 * PB_MAX_WORLD;  // staged records incl. run padding

// inverse of move a (ac_moves.py:188-206): 0<->2, 1<->3 (r_i r_j^{+-1}); 4<->8, 5<->9, 6<->10, 7<->11 (g <-> g^-1)
__device__ __forceinline__ int pb_inverse_move(int a) { return a < 4 ? (a ^ 2) : (a < 8 ? a + 4 : a - 4); }

__global__ void __launch_bounds__(EX_THREADS) k_pbfs_expand2(const uint64_t* __restrict__ nodes, const uint8_t* __restrict__ action,
                                                            const int64_t* __restrict__ loc, const int64_t* __restrict__ gids,
                                                            int64_t ord_base, int64_t f, int m, int cyclical, int world,
                                                            int64_t region_cap, unsigned long long* cursor, const PbDst2 dst,
                                                            const int64_t* ctl) {
    if (ctl[CTL_STOP] != 0) return;
    extern __shared__ __align__(128) uint8_t ex_smem[];
    uint4* stage = reinterpret_cast<uint4*>(ex_smem);                                   // EX_SLOTS records of 2 x uint4
    unsigned* s_wcnt = reinterpret_cast<unsigned*>(ex_smem + (size_t)EX_SLOTS * 32);    // [8 warps][PB_MAX_WORLD]
    unsigned* s_off = s_wcnt + 8 * PB_MAX_WORLD;                                        // [PB_MAX_WORLD] start of the owner's run in `stage`
    unsigned* s_pad = s_off + PB_MAX_WORLD;                                             // [PB_MAX_WORLD] padded run length
    unsigned long long* s_base = reinterpret_cast<unsigned long long*>(s_pad + PB_MAX_WORLD);  // [PB_MAX_WORLD] slot in the owner's region
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned ltmask = (1u << lane) - 1u;
    const bool cyc = cyclical != 0;
    for (int64_t k0 = (int64_t)blockIdx.x * EX_THREADS; k0 < f; k0 += (int64_t)gridDim.x * EX_THREADS) {
        const int64_t k = k0 + tid;
        const bool active = k < f;
        R64 r[2];
        r[0].S = r[0].T = r[1].S = r[1].T = 0ull;
        r[0].n = r[1].n = 1;
        int64_t ord0 = 0;
        int skip = -1;
        if (active) {
            const int64_t id = loc[k];
            const PWords v = pb_load(nodes + 4 * id);
            r[0].S = v.w[0] & PB_PLANE;
            r[0].T = v.w[1] & PB_PLANE;
            r[0].n = (int)((v.w[0] >> 40) & 63u);
            r[1].S = v.w[2] & PB_PLANE;
            r[1].T = v.w[3] & PB_PLANE;
            r[1].n = (int)((v.w[1] >> 40) & 63u);
            ord0 = (gids[k] - ord_base) * 12;
            const int pa = action[id];
            if (!cyc && pa < 12) skip = pb_inverse_move(pa);  // that child is the grandparent: visited
        }
#pragma unroll 1
        for (int g = 0; g < 12 / EX_GM; ++g) {
            PWords cw[EX_GM];
            int own[EX_GM];
            unsigned rank[EX_GM];
#pragma unroll
            for (int j = 0; j < EX_GM; ++j) {
                const int a = g * EX_GM + j;
                own[j] = -1;
                rank[j] = 0;
                if (!active || a == skip) continue;
                const uint32_t d = decode_move(a);
                const bool i = d & 1u, is_conj = d & 2u, inv = d & 4u;
                R64 x = i ? r[1] : r[0];
                bool changed;
                if (!is_conj) {
                    R64 b = i ? r[0] : r[1];
                    if (inv) b = r64_inverse(b);
                    changed = r64_concat(x, b, m);
                } else {
                    changed = r64_conjugate(x, inv ? 1u : 0u, (d >> 3) & 1u, m);
                }
                uint64_t st = 0;
                if (x.n == 0) {  // utils.py:264-266 raises: report, the record carries the parent
                    st = ACX_ROW_EMPTY_RELATOR;
                    x = i ? r[1] : r[0];
                } else {
                    if (changed && cyc) r64_cyclic_reduce(x);
                    const R64& old = i ? r[1] : r[0];
                    if (!changed || (x.n == old.n && x.S == old.S && x.T == old.T)) continue;  // the parent itself: visited
                }
                cw[j] = i ? pb_pack(r[0], x) : pb_pack(x, r[1]);
                own[j] = pb_owner(pb_hash(cw[j]), world);
                const uint64_t ord = (uint64_t)(ord0 + a);
                cw[j].w[2] |= (ord & 0xFFFFFFFull) << 36;
                cw[j].w[3] |= (((ord >> 28) & 0xFFFFFull) << 36) | (st << 56);
            }
            // rank of every record inside its (warp, owner) group; lane o keeps the warp's count for owner o
            unsigned mycnt = 0;
            for (int o = 0; o < world; ++o) {
                unsigned run = 0;
#pragma unroll
                for (int j = 0; j < EX_GM; ++j) {
                    const unsigned b = __ballot_sync(0xFFFFFFFFu, own[j] == o);
                    if (own[j] == o) rank[j] = run + __popc(b & ltmask);
                    run += __popc(b);
                }
                if (lane == o) mycnt = run;
            }
            if (lane < world) s_wcnt[warp * PB_MAX_WORLD + lane] = mycnt;
            __syncthreads();
            if (tid < world) {  // thread o: prefix over the warps, padded run, one reservation in owner o's region
                unsigned tot = 0;
                for (int w = 0; w < EX_THREADS / 32; ++w) {
                    const unsigned c = s_wcnt[w * PB_MAX_WORLD + tid];
                    s_wcnt[w * PB_MAX_WORLD + tid] = tot;
                    tot += c;
                }
                const unsigned pad = (tot + 3u) & ~3u;
                s_pad[tid] = pad | (tot << 16);
                s_base[tid] = pad ? atomicAdd(cursor + tid, (unsigned long long)pad) : 0ull;
            }
            __syncthreads();
            if (tid == 0) {
                unsigned off = 0;
                for (int o = 0; o < world; ++o) {
                    s_off[o] = off;
                    off += s_pad[o] & 0xFFFFu;
                }
            }
            __syncthreads();
#pragma unroll
            for (int j = 0; j < EX_GM; ++j) {
                if (own[j] < 0) continue;
                const unsigned pos = s_off[own[j]] + s_wcnt[warp * PB_MAX_WORLD + own[j]] + rank[j];
                stage[2 * pos] = make_uint4((uint32_t)cw[j].w[0], (uint32_t)(cw[j].w[0] >> 32), (uint32_t)cw[j].w[1], (uint32_t)(cw[j].w[1] >> 32));
                stage[2 * pos + 1] = make_uint4((uint32_t)cw[j].w[2], (uint32_t)(cw[j].w[2] >> 32), (uint32_t)cw[j].w[3], (uint32_t)(cw[j].w[3] >> 32));
            }
            if (tid < world) {  // null records pad the run to a whole line
                const unsigned pad = s_pad[tid] & 0xFFFFu, tot = s_pad[tid] >> 16;
                for (unsigned q = tot; q < pad; ++q) {
                    stage[2 * (s_off[tid] + q)] = make_uint4(0u, 0u, 0u, 0u);
                    stage[2 * (s_off[tid] + q) + 1] = make_uint4(0u, 0u, 0u, 0u);
                }
            }
            __syncthreads();
            for (int o = 0; o < world; ++o) {  // the run of owner o: consecutive threads, consecutive 16-byte pieces
                const unsigned pad = s_pad[o] & 0xFFFFu;
                if (pad == 0) continue;
                const unsigned long long base = s_base[o];
                if (base + pad > (unsigned long long)region_cap) {
                    if (tid == 0) cursor[world] = 1ull;  // a region overflowed: the host raises
                    continue;
                }
                uint4* out = reinterpret_cast<uint4*>(dst.region[o] + base * 32ull);
                const uint4* in = stage + 2 * s_off[o];
                for (unsigned q = tid; q < 2 * pad; q += EX_THREADS) out[q] = in[q];
            }
            __syncthreads();
        }
    }
}

struct Pb2Params {
    const uint64_t* recs_prev;    // the records of the previous round (its claims may still be provisional in the table)
    uint64_t parity;              // of this round: provisional entries carry it in bit 46 of the value
    const uint64_t* recs;         // nreg regions of rcap records
    const int64_t* region_counts; // device: records (incl. null padding) in region s = region_counts[s * count_stride]
    int64_t count_stride;
    int nreg;
    int64_t rcap;
    uint64_t* table;
    int64_t tcap;
    const uint64_t* nodes;
    int64_t* slot;
    unsigned long long* mins;
    unsigned* mask32;  // 16 bits per parent, two parents per word
    const int32_t* excl;
    int64_t parent_base;
    int64_t* ctl;
    unsigned long long* n_local;
    uint64_t* nodes_w;
    int64_t* gid;
    int64_t* parent;
    uint8_t* action;
};

__device__ __forceinline__ void mask_set(unsigned* mask32, int64_t ord) {
    const int64_t p = ord / 12;
    atomicOr(mask32 + (p >> 1), (1u << (int)(ord % 12)) << (16 * (int)(p & 1)));
}
__device__ __forceinline__ void mask_clear(unsigned* mask32, int64_t ord) {
    const int64_t p = ord / 12;
    atomicAnd(mask32 + (p >> 1), ~((1u << (int)(ord % 12)) << (16 * (int)(p & 1))));
}

// Table slot = tag(16) | PROV(1) | value(47): value = node id, or (PROV) the index of the candidate that currently
// claims the state in this round.  slot[j] >= 0 and the mask bit of j are set exactly for the claim holders when the
// kernel ends: the first discovery (smallest order) of every new state.
template <int MINB>
__global__ void __launch_bounds__(256, MINB) k_pbfs_probe2(const Pb2Params p) {
    if (p.ctl[CTL_STOP] != 0) return;
    const int64_t s = blockIdx.y, i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;  // region, record in the region
    if (i >= p.region_counts[s * p.count_stride]) return;
    const int64_t j = s * p.rcap + i;
    const PWords rec = pb_load(p.recs + 4 * j);
    if ((rec.w[0] | rec.w[1]) == 0ull) {  // null record (run padding)
        p.slot[j] = -3;
        return;
    }
    const int64_t ord = pb_order(rec);
    if (pb_status(rec)) {
        atomicMin(p.mins + 1, (unsigned long long)ord);
        p.slot[j] = -2;
        return;
    }
    if (pb_total(rec) == 2) {  // breadth_first.py:84-85, tested before the visited set
        atomicMin(p.mins, (unsigned long long)ord);
        p.slot[j] = -2;
        return;
    }
    const PWords st = pb_state(rec);
    const uint64_t h = pb_hash(st);
    const uint64_t tag = ((h >> 48) | 1ull) << 48;
    const uint64_t mine = tag | PB_PROV | (p.parity << 46) | (uint64_t)j;
    const uint64_t mask = (uint64_t)p.tcap - 1;
    uint64_t slot = (h ^ (h >> 23)) & mask;
    for (;;) {
        uint64_t e = __ldcg(p.table + slot);
        if (e == 0ull) {
            // tentative winner first, claim second: whoever replaces the claim later finds both in place
            mask_set(p.mask32, ord);
            p.slot[j] = (int64_t)slot;
            __threadfence();
            const uint64_t old = atomicCAS(reinterpret_cast<unsigned long long*>(p.table + slot), 0ull, (unsigned long long)mine);
            if (old == 0ull) return;
            mask_clear(p.mask32, ord);  // nobody saw the claim: undo
            e = old;
        }
        if ((e >> 48) == (tag >> 48)) {
            if ((e & PB_PROV) && ((e >> 46) & 1ull) != p.parity) {
                // a claim of the PREVIOUS round that its commit (running next to this probe) has not turned into a node
                // yet: every candidate of that round precedes this one in the reference's order, so an equal state is visited
                const PWords orec = pb_load(p.recs_prev + 4 * (int64_t)(e & PB_IDX));
                if (pb_equal(st, pb_state(orec))) {
                    p.slot[j] = -1;
                    return;
                }
            } else if (e & PB_PROV) {
                for (;;) {
                    const int64_t other = (int64_t)(e & PB_IDX);
                    const PWords orec = pb_load(p.recs + 4 * other);
                    if (!pb_equal(st, pb_state(orec))) break;
                    const int64_t oord = pb_order(orec);
                    if (oord < ord) {  // a duplicate inside the round, discovered earlier in the reference's order
                        p.slot[j] = -1;
                        return;
                    }
                    mask_set(p.mask32, ord);
                    p.slot[j] = (int64_t)slot;
                    __threadfence();
                    const uint64_t old = atomicCAS(reinterpret_cast<unsigned long long*>(p.table + slot), (unsigned long long)e,
                                                   (unsigned long long)mine);
                    if (old == e) {  // the claim is mine now: the loser's bit and slot go
                        mask_clear(p.mask32, oord);
                        p.slot[other] = -1;
                        return;
                    }
                    mask_clear(p.mask32, ord);
                    e = old;  // somebody else replaced it first: look again
                }
            } else {
                const int64_t id = (int64_t)(e & PB_VAL);
                if (pb_equal(st, pb_load_cg(p.nodes + 4 * id))) {  // the previous round's commit may be appending nodes right now
                    p.slot[j] = -1;
                    return;
                }
            }
        }
        slot = (slot + 1) & mask;
    }
}

struct PopcountMask16 {
    __host__ __device__ __forceinline__ int32_t operator()(const uint16_t& m) const {
#ifdef __CUDA_ARCH__
        return (int32_t)__popc((unsigned)m);
#else
        return (int32_t)__builtin_popcount((unsigned)m);
#endif
    }
};

// one thread: the reference's decision for the round (acx/bfs_sharded.py ShardedBFS.run restated on the device)
__global__ void k_pbfs_decide(const uint16_t* mask16, const int32_t* excl, int64_t n_parents, const unsigned long long* mins,
                              int64_t parent_base, int64_t budget, int64_t* ctl, const int64_t* counts, int world, int rank) {
    if (ctl[CTL_STOP] != 0) return;
    // counts[s * (world + 1) + o] = records rank s staged for owner o this round, [.. + world] = its overflow flag
    for (int g = 0; g < world; ++g) {
        if (counts[g * (world + 1) + world]) ctl[CTL_OVERFLOW] = 1;
        ctl[CTL_STAGED] += counts[rank * (world + 1) + g];
    }
    ctl[CTL_STAGED_LOCAL] += counts[rank * (world + 1) + rank];
    const int64_t n_nodes = ctl[CTL_NODES];
    const int64_t n_new = (int64_t)excl[n_parents - 1] + __popc((unsigned)mask16[n_parents - 1]);
    const unsigned long long c_s = mins[0], c_e = mins[1];
    long long best = -1;
    int what = STOP_NONE;
    if (c_s < PB_SENT) {
        best = (long long)c_s;
        what = STOP_SOLVED;
    }
    if (c_e < PB_SENT && (what == STOP_NONE || (long long)c_e < best)) {
        best = (long long)c_e;
        what = STOP_ERROR;
    }
    int64_t p_star = -1;
    if (n_nodes + n_new >= budget) {
        // first parent after which len(tree_nodes) >= max_nodes_to_explore (breadth_first.py:91-95): the first index whose
        // exclusive count reaches the target is one past that parent
        const int64_t target = budget - n_nodes;
        int64_t lo = 0, hi = n_parents;  // first idx with excl[idx] >= target
        while (lo < hi) {
            const int64_t mid = (lo + hi) >> 1;
            if ((int64_t)excl[mid] >= target) hi = mid;
            else lo = mid + 1;
        }
        p_star = lo - 1 > 0 ? lo - 1 : 0;
        const long long pos = 12 * p_star + 11;
        if (what == STOP_NONE || pos < best) {  // a success / error at the same position comes first (action order)
            best = pos;
            what = STOP_BUDGET;
        }
    }
    ctl[CTL_ROUNDS] += 1;
    if (what == STOP_NONE) {
        ctl[CTL_GID_BASE] = n_nodes;
        ctl[CTL_NODES] = n_nodes + n_new;
        ctl[CTL_EXPANDED] += n_parents;
        return;
    }
    ctl[CTL_STOP] = what;
    ctl[CTL_POS] = best;
    ctl[CTL_PARENT_BASE] = parent_base;
    const int64_t p = best / 12;
    const int a = (int)(best % 12);
    if (what == STOP_BUDGET) {
        ctl[CTL_EXPANDED] += p + 1;
        ctl[CTL_NODES] = n_nodes + excl[p] + __popc((unsigned)mask16[p]);
    } else {
        ctl[CTL_EXPANDED] += p + 1;
        ctl[CTL_NODES] = n_nodes + excl[p] + __popc((unsigned)mask16[p] & ((1u << a) - 1u));
    }
}

__global__ void __launch_bounds__(256) k_pbfs_commit2(const Pb2Params p) {
    if (p.ctl[CTL_STOP] != 0) return;
    const int64_t s = blockIdx.y, i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;  // region, record in the region
    if (i >= p.region_counts[s * p.count_stride]) return;
    const int64_t j = s * p.rcap + i;
    const int64_t sl = p.slot[j];
    if (sl < 0) return;
    const PWords rec = pb_load(p.recs + 4 * j);
    const int64_t ord = pb_order(rec);
    const int64_t par = ord / 12;
    const int act = (int)(ord % 12);
    const unsigned m16 = (p.mask32[par >> 1] >> (16 * (int)(par & 1))) & 0xFFFFu;
    const int64_t gid = p.ctl[CTL_GID_BASE] + p.excl[par] + __popc(m16 & ((1u << act) - 1u));
    const int64_t id = (int64_t)atomicAdd(p.n_local, 1ull);
    const PWords st = pb_state(rec);
    pb_store(p.nodes_w + 4 * id, st);
    p.gid[id] = gid;
    p.parent[id] = p.parent_base + par;
    p.action[id] = (uint8_t)act;
    // the probe of the next round may run next to this kernel: the node must be visible before the slot names it
    __threadfence();
    // the slot keeps its tag (recomputed from the state: no read of the random table line) and now names the node
    __stcg(p.table + sl, (((pb_hash(st) >> 48) | 1ull) << 48) | (uint64_t)id);
}

}  // namespace acx

using namespace acx;

extern "C" int acx_pbfs_expand2(const uint64_t* nodes, const uint8_t* action, const int64_t* loc, const int64_t* gids,
                                int64_t ord_base, int64_t f, int m, int cyclical, int world, int64_t region_cap,
                                uint64_t* cursor, uint8_t* const* dst_regions, const int64_t* ctl, acx_stream_t stream) {
    if (f < 0 || m < 1 || m > 36 || world < 1 || world > PB_MAX_WORLD || region_cap < 4 || (region_cap & 3))
        return acx_fail(ACX_ERR_ARG, "acx_pbfs_expand2: bad argument (1 <= m <= 36, 1 <= world <= 32, region_cap a multiple of 4)");
    if (f == 0) return ACX_OK;
    if (f > (1ll << 44)) return acx_fail(ACX_ERR_ARG, "acx_pbfs_expand2: order numbers are 48 bits");
    if (!nodes || !action || !loc || !gids || !cursor || !dst_regions || !ctl) return acx_fail(ACX_ERR_ARG, "acx_pbfs_expand2: null pointer");
    PbDst2 dst = {};
    for (int o = 0; o < world; ++o) {
        if (!dst_regions[o] || ((uintptr_t)dst_regions[o] & 127u))
            return acx_fail(ACX_ERR_ARG, "acx_pbfs_expand2: destination regions must be non-null and 128-byte aligned");
        dst.region[o] = dst_regions[o];
    }
    if ((uintptr_t)nodes & 31u) return acx_fail(ACX_ERR_ARG, "acx_pbfs_expand2: nodes must be 32-byte aligned");
    const size_t smem = (size_t)EX_SLOTS * 32 + (8 * PB_MAX_WORLD + 2 * PB_MAX_WORLD) * sizeof(unsigned) + PB_MAX_WORLD * sizeof(unsigned long long);
    static int ctas_per_sm[ACX_MAX_DEVICES] = {0};
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= ACX_MAX_DEVICES) dev = 0;
    if (ctas_per_sm[dev] == 0) {
        ACX_CUDA(cudaFuncSetAttribute(k_pbfs_expand2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        int occ = 0;
        ACX_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_pbfs_expand2, EX_THREADS, smem));
        ctas_per_sm[dev] = occ > 0 ? occ : 1;
    }
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    int64_t grid = (f + EX_THREADS - 1) / EX_THREADS;
    if (grid > (int64_t)sms * ctas_per_sm[dev]) grid = (int64_t)sms * ctas_per_sm[dev];
    k_pbfs_expand2<<<(unsigned)grid, EX_THREADS, smem, (cudaStream_t)stream>>>(nodes, action, loc, gids, ord_base, f, m, cyclical, world,
                                                                              region_cap, reinterpret_cast<unsigned long long*>(cursor), dst, ctl);
    ACX_CUDA(cudaGetLastError());
    return ACX_OK;
}

static int pb2_common(Pb2Params& p, const uint64_t* recs, const int64_t* region_counts_dev, int64_t count_stride, int n_regions,
                      int64_t region_cap, const char* who) {
    if (!recs || !region_counts_dev || count_stride < 1 || n_regions < 1 || n_regions > PB_MAX_WORLD || region_cap < 1)
        return acx_fail(ACX_ERR_ARG, "%s: need records, device region counts, 1 <= n_regions <= 32", who);
    p.recs = recs;
    p.region_counts = region_counts_dev;
    p.count_stride = count_stride;
    p.nreg = n_regions;
    p.rcap = region_cap;
    return ACX_OK;
}

extern "C" int acx_pbfs_probe2(const uint64_t* recs, const uint64_t* recs_prev, int parity, const int64_t* region_counts_dev,
                               int64_t count_stride, int n_regions, int64_t region_cap, uint64_t* table, int64_t tcap, const uint64_t* nodes, int64_t* slot, uint64_t* mins,
                               uint32_t* mask32, const int64_t* ctl, acx_stream_t stream) {
    if (tcap < 2 || (tcap & (tcap - 1))) return acx_fail(ACX_ERR_ARG, "acx_pbfs_probe2: table capacity must be a power of two");
    Pb2Params p = {};
    if (int rc = pb2_common(p, recs, region_counts_dev, count_stride, n_regions, region_cap, "acx_pbfs_probe2")) return rc;
    if (!table || !nodes || !slot || !mins || !mask32 || !ctl || !recs_prev) return acx_fail(ACX_ERR_ARG, "acx_pbfs_probe2: null pointer");
    if ((int64_t)n_regions * region_cap > (int64_t)PB_IDX) return acx_fail(ACX_ERR_ARG, "acx_pbfs_probe2: too many records per round");
    p.recs_prev = recs_prev;
    p.parity = (uint64_t)(parity & 1);
    p.table = table; p.tcap = tcap; p.nodes = nodes; p.slot = slot;
    p.mins = reinterpret_cast<unsigned long long*>(mins); p.mask32 = mask32; p.ctl = const_cast<int64_t*>(ctl);
    const dim3 grid((unsigned)((region_cap + 255) / 256), (unsigned)n_regions);
    static const int occ = [] {  // ACX_PROBE_OCC=5|6|8 resident CTAs per SM (registers 44 / 40 / 32): tuning knob
        const char* e = getenv("ACX_PROBE_OCC");
        return e ? atoi(e) : 6;
    }();
    if (occ >= 8) k_pbfs_probe2<8><<<grid, 256, 0, (cudaStream_t)stream>>>(p);
    else if (occ == 6) k_pbfs_probe2<6><<<grid, 256, 0, (cudaStream_t)stream>>>(p);
    else k_pbfs_probe2<5><<<grid, 256, 0, (cudaStream_t)stream>>>(p);
    ACX_CUDA(cudaGetLastError());
    return ACX_OK;
}

// temp == NULL: scratch size in bytes.  mask16: uint16 per parent (after the all-reduce); excl_out int32 [n_parents].
extern "C" int64_t acx_pbfs_decide(const uint16_t* mask16, int32_t* excl_out, int64_t n_parents, const uint64_t* mins,
                                   int64_t parent_base, int64_t budget, int64_t* ctl, const int64_t* counts, int world, int rank,
                                   void* temp, int64_t temp_bytes, acx_stream_t stream) {
    if (n_parents < 1 || n_parents > 0x7FFFFFFFll / 12) return acx_fail(ACX_ERR_ARG, "acx_pbfs_decide: 1 <= n_parents < 2^31 / 12");
    cub::TransformInputIterator<int32_t, PopcountMask16, const uint16_t*> in(mask16, PopcountMask16());
    size_t need = 0;
    cudaError_t e = cub::DeviceScan::ExclusiveSum(nullptr, need, in, excl_out, (int)n_parents, (cudaStream_t)stream);
    if (e != cudaSuccess) return acx_fail(ACX_ERR_CUDA, "cub::DeviceScan (size query): %s", cudaGetErrorString(e));
    if (!temp) return (int64_t)need;
    if (!mask16 || !excl_out || !mins || !ctl || !counts) return acx_fail(ACX_ERR_ARG, "acx_pbfs_decide: null pointer");
    if (world < 1 || world > PB_MAX_WORLD || rank < 0 || rank >= world) return acx_fail(ACX_ERR_ARG, "acx_pbfs_decide: bad world / rank");
    if ((size_t)temp_bytes < need) return acx_fail(ACX_ERR_ARG, "acx_pbfs_decide: scratch too small");
    size_t tb = (size_t)temp_bytes;
    e = cub::DeviceScan::ExclusiveSum(temp, tb, in, excl_out, (int)n_parents, (cudaStream_t)stream);
    if (e != cudaSuccess) return acx_fail(ACX_ERR_CUDA, "cub::DeviceScan: %s", cudaGetErrorString(e));
    k_pbfs_decide<<<1, 1, 0, (cudaStream_t)stream>>>(mask16, excl_out, n_parents, reinterpret_cast<const unsigned long long*>(mins),
                                                    parent_base, budget, ctl, counts, world, rank);
    ACX_CUDA(cudaGetLastError());
    return ACX_OK;
}

extern "C" int acx_pbfs_commit2(const uint64_t* recs, const int64_t* region_counts_dev, int64_t count_stride, int n_regions, int64_t region_cap,
                                const int64_t* slot, const int32_t* excl, const uint32_t* mask32, int64_t parent_base,
                                const int64_t* ctl, uint64_t* n_local, uint64_t* table, uint64_t* nodes, int64_t* gid,
                                int64_t* parent, uint8_t* action, acx_stream_t stream) {
    Pb2Params p = {};
    if (int rc = pb2_common(p, recs, region_counts_dev, count_stride, n_regions, region_cap, "acx_pbfs_commit2")) return rc;
    if (!slot || !excl || !mask32 || !ctl || !n_local || !table || !nodes || !gid || !parent || !action)
        return acx_fail(ACX_ERR_ARG, "acx_pbfs_commit2: null pointer");
    p.slot = const_cast<int64_t*>(slot); p.excl = excl; p.mask32 = const_cast<unsigned*>(mask32); p.parent_base = parent_base;
    p.ctl = const_cast<int64_t*>(ctl); p.n_local = reinterpret_cast<unsigned long long*>(n_local); p.table = table;
    p.nodes_w = nodes; p.gid = gid; p.parent = parent; p.action = action;
    const dim3 grid((unsigned)((region_cap + 255) / 256), (unsigned)n_regions);
    k_pbfs_commit2<<<grid, 256, 0, (cudaStream_t)stream>>>(p);
    ACX_CUDA(cudaGetLastError());
    return ACX_OK;
}
