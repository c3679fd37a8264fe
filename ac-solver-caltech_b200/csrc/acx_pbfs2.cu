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
//   k_pbfs_probe3    visited-set probe in which a claim is FINAL: the candidate reserves its place in the node store
//                    first and claims the table slot with that place, so the slot is written once in its life (the machine
//                    serves 36 G random memory operations / s, reads and writes alike: tools/random_access.cu -- the first
//                    version wrote the slot twice, provisional claim and node id).  Which of the equal candidates of a
//                    round discovered the state first is settled beside the node (atomicMin of the order, in gid[]), and
//                    the per-parent mask follows by XOR.
//   k_pbfs_decide    after the masks are all-reduced: exclusive scan of the new-node counts, then ONE thread replays the
//                    reference's decision for the round (first success / error by order, the budget cut of
//                    breadth_first.py:91-95) into a control block in device memory.
//   k_pbfs_commit3   the nodes of the round get their global id, parent and action from the winning order -- a
//                    streaming pass over the round's part of the node store -- unless the control block says the search
//                    stopped in this round.
// Every kernel returns at once when the control block holds a stop code, so the host enqueues whole levels without
// reading anything back and looks at the control block once per level.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/acx.h"
#include "acx_common.h"
#include "acx_pbnode.cuh"
#include "acx_tma.cuh"

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

template <int MINB, bool BULK>
__global__ void __launch_bounds__(EX_THREADS, MINB) k_pbfs_expand2(const uint64_t* __restrict__ nodes, const uint8_t* __restrict__ action,
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
        const bool active = k < f && gids[k] >= 0;  // gid -1: a hole of the node store (acx_pbfs_probe3), not a node
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
                if (BULK) bulk_wait_read<0>();  // the bulk store this thread issued for the previous group has read the stage
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
            if (BULK) {
                // the run of owner o leaves as ONE bulk async copy (shared -> global, local or peer memory over NVLink),
                // issued by thread o: no thread waits for the stores, the next four moves are computed meanwhile
                fence_proxy_async();
                __syncthreads();
                if (tid < world) {
                    const unsigned pad = s_pad[tid] & 0xFFFFu;
                    const unsigned long long base = s_base[tid];
                    if (pad != 0) {
                        if (base + pad > (unsigned long long)region_cap) cursor[world] = 1ull;  // a region overflowed: the host raises
                        else bulk_store(dst.region[tid] + base * 32ull, smem_u32(stage + 2 * s_off[tid]), pad * 32u);
                    }
                    bulk_commit();
                }
            } else {
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
            }
            __syncthreads();
        }
    }
    if (BULK && tid < world) bulk_wait_all<0>();  // the stage must outlive the copies that read it
}

struct Pb3Params {
    const uint64_t* recs;         // nreg regions of rcap records
    const int64_t* region_counts; // device: records (incl. null padding) in region s = region_counts[s * count_stride]
    int64_t count_stride;
    int nreg;
    int64_t rcap;
    uint64_t* table;              // slot = tag(16) | place of the node in the local store (48)
    int64_t tcap;
    uint64_t* nodes;
    int64_t node_cap;
    unsigned long long* gid;      // during the round: PB_IN_ROUND | smallest order seen for the node; -1 = never used
    unsigned long long* n_local;
    const int64_t* round_start;   // device: n_local before this round's probe
    const int64_t* round_end;     // device: n_local behind it (commit)
    unsigned long long* mins;
    unsigned* mask32;             // 16 bits per parent, two parents per word
    const int32_t* excl;
    int64_t parent_base;
    int64_t* ctl;
    int64_t* parent;
    uint8_t* action;
    int64_t* round_log;           // commit: {first place, end place, first global id, new nodes} of the round, or NULL
};

static constexpr unsigned long long PB_IN_ROUND = 1ull << 62;  // keeps the orders of a round apart from real global ids
static constexpr uint64_t PB_PLACE = (1ull << 48) - 1;

// The winner mask is kept with XOR: a candidate toggles its own bit when it becomes the node's smallest order, and
// toggles the bit of the order it displaced.  A bit is toggled at most once by its owner and at most once by its
// displacer, and XOR commutes: when the kernel ends bit (p, a) is set exactly for the first discovery of every new
// state, whatever order the toggles became visible in -- no fence anywhere in the probe.
__device__ __forceinline__ void mask_flip(unsigned* mask32, int64_t ord) {
    if (((uint64_t)ord >> 32) == 0ull) {  // every round of less than 2^32 / 12 parents: a 32-bit division (the 64-bit one was a third of the kernel's instructions)
        const uint32_t o = (uint32_t)ord, p = o / 12u, a = o - 12u * p;
        atomicXor(mask32 + (p >> 1), (1u << a) << (16u * (p & 1u)));
        return;
    }
    const int64_t p = ord / 12;
    atomicXor(mask32 + (p >> 1), (1u << (int)(ord % 12)) << (16 * (int)(p & 1)));
}

__device__ __forceinline__ void settle_order(const Pb3Params& p, int64_t place, int64_t ord) {
    const unsigned long long mine = PB_IN_ROUND | (unsigned long long)ord;
    const unsigned long long old = atomicMin(p.gid + place, mine);
    if (old > mine) {
        mask_flip(p.mask32, ord);
        if (old != ~0ull) mask_flip(p.mask32, (int64_t)(old & ~PB_IN_ROUND));
    }
}

// Walk the probe sequence of one candidate from `slot` (whose content `e` is already loaded).  Returns true when the
// candidate is settled (claimed its node, or met an equal state), false when it stands at an EMPTY slot and holds no
// place to claim it with (`spare` < 0): the caller reserves one and calls again.
__device__ __forceinline__ bool probe_walk(const Pb3Params& p, const PWords& st, int64_t ord, uint64_t tag, uint64_t tmask, int64_t rs,
                                           uint64_t& slot, uint64_t& e, int64_t& spare) {
    for (;;) {
        if (e == 0ull) {
            if (spare < 0) return false;
            const uint64_t old = atomicCAS(reinterpret_cast<unsigned long long*>(p.table + slot), 0ull, (unsigned long long)(tag | (uint64_t)spare));
            if (old == 0ull) {  // the state is new and this is its node, for good
                pb_store(p.nodes + 4 * spare, st);
                settle_order(p, spare, ord);
                spare = -1;
                return true;
            }
            e = old;
        }
        if ((e >> 48) == (tag >> 48)) {
            const int64_t place = (int64_t)(e & PB_PLACE);
            PWords nd = pb_load_cg(p.nodes + 4 * place);
            if (place >= rs) {  // claimed in this round: its owner may be between the claim and the store
                while (nd.w[0] == 0ull) {
                    __nanosleep(32);
                    nd = pb_load_cg(p.nodes + 4 * place);
                }
            }
            if (pb_equal(st, nd)) {
                // visited -- or discovered earlier in this very round: the smaller order is the reference's first discovery
                if (place >= rs) settle_order(p, place, ord);
                return true;
            }
        }
        slot = (slot + 1) & tmask;
        e = __ldcg(p.table + slot);
    }
}

// Persistent grid: a thread walks over the records of its region with a stride.  Places in the node store are handed
// out where the CTA is converged, to exactly the lanes that stand at an empty slot without a place, from a pool the CTA
// refills with ONE atomic on the store's counter for several iterations at a time (every atomic on that one address
// costs about ten cycles in L2 and they serialize: with one per warp and iteration the counter alone set the kernel's
// duration).  A place whose claim then lost its race (an equal candidate claimed the slot first) is kept for the thread's
// next claim; what a thread or the pool still holds at the end stays a hole in the node store (gid -1, never expanded;
// the refills shrink to the exact need in the CTA's last iterations).
template <int NT, int MINB>
__global__ void __launch_bounds__(NT, MINB) k_pbfs_probe3(const Pb3Params p) {
    if (p.ctl[CTL_STOP] != 0) return;
    __shared__ int s_wneed[NT / 32];
    __shared__ long long s_range[4];  // old pool base, places taken from it, new chunk base
    __shared__ long long s_pool[2];   // base, left
    const int64_t s = blockIdx.y;  // region
    const int64_t cnt = p.region_counts[s * p.count_stride];
    const int64_t T = (int64_t)gridDim.x * NT;
    const int64_t rs = *p.round_start;
    const uint64_t* recs = p.recs + 4 * s * p.rcap;
    const uint64_t tmask = (uint64_t)p.tcap - 1;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int64_t spare = -1;
    if (threadIdx.x == 0) s_pool[0] = s_pool[1] = 0;
    // A large round is walked with a stride: iteration `it` of thread t takes record it * T + t (coalesced record loads).
    // A small round -- fewer than 17 records per thread -- is walked in blocks instead: thread t takes the 17 CONSECUTIVE
    // records [17 t, 17 t + 17), one per iteration, and the rest of the grid idles.  Equal candidates sit close to each other
    // in a region (children of sibling parents, mostly a multiple of 32 records apart); handled by neighbouring threads in
    // the same iteration they race for the empty slot and the loser's place is wasted unless the thread claims again
    // later, which in a small round it never does (one record per thread: a fifth of the store became holes).  17 is odd:
    // records 32 apart fall into different iterations.
    constexpr int64_t RB = 17;
    const int64_t per_thread = (cnt + T - 1) / T;
    const bool blocked = per_thread < RB;
    const int64_t gt = (int64_t)blockIdx.x * NT + threadIdx.x;
    if ((int64_t)blockIdx.x * NT * (blocked ? RB : 1) >= cnt) return;  // block-uniform
    const int64_t R = blocked ? RB : (cnt - (int64_t)blockIdx.x * NT + T - 1) / T;  // iterations of THIS CTA
    for (int64_t it = 0; it < R; ++it) {  // block-uniform: the CTA meets at the barriers
        const int64_t i = blocked ? gt * RB + it : it * T + gt;
        bool live = false, settled = true;
        PWords st;
        st.w[0] = st.w[1] = st.w[2] = st.w[3] = 0ull;
        int64_t ord = 0;
        uint64_t tag = 0, slot = 0, e = 0;
        if (i < cnt) {
            const PWords rec = pb_load(recs + 4 * i);
            if ((rec.w[0] | rec.w[1]) != 0ull) {  // not a null record (run padding)
                ord = pb_order(rec);
                if (pb_status(rec)) {
                    atomicMin(p.mins + 1, (unsigned long long)ord);
                } else if (pb_total(rec) == 2) {  // breadth_first.py:84-85, tested before the visited set
                    atomicMin(p.mins, (unsigned long long)ord);
                } else {
                    live = true;
                    st = pb_state(rec);
                    const uint64_t h = pb_hash(st);
                    tag = ((h >> 48) | 1ull) << 48;
                    slot = (h ^ (h >> 23)) & tmask;
                    e = __ldcg(p.table + slot);
                }
            }
        }
        if (live) settled = probe_walk(p, st, ord, tag, tmask, rs, slot, e, spare);
        const unsigned need = __ballot_sync(0xFFFFFFFFu, !settled);
        if (lane == 0) s_wneed[warp] = __popc(need);
        __syncthreads();
        if (threadIdx.x == 0) {
            int total = 0;
            for (int w = 0; w < NT / 32; ++w) {
                const int c = s_wneed[w];
                s_wneed[w] = total;
                total += c;
            }
            long long pb = s_pool[0], left = s_pool[1];
            s_range[0] = pb;
            if (left >= total) {
                s_range[1] = total;
                s_range[2] = 0;
                s_pool[0] = pb + total;
                s_pool[1] = left - total;
            } else {  // the pool runs dry: what it holds goes first, the rest comes from a new chunk
                const int64_t iters_left = R - it;  // this one included
                const long long mult = iters_left - 1 > 8 ? 8 : (iters_left - 1 < 1 ? 1 : iters_left - 1);
                const long long want = (long long)(total - left) + (mult - 1) * total;
                const long long nb = (long long)atomicAdd(p.n_local, (unsigned long long)want);
                s_range[1] = left;
                s_range[2] = nb;
                s_pool[0] = nb + (total - left);
                s_pool[1] = want - (total - left);
            }
        }
        __syncthreads();
        if (!settled) {
            const long long r = s_wneed[warp] + __popc(need & ((1u << lane) - 1u));
            spare = r < s_range[1] ? s_range[0] + r : s_range[2] + (r - s_range[1]);
            if (spare >= p.node_cap) {  // the node store is full: the host raises
                p.ctl[CTL_OVERFLOW] = 1;
                spare = -1;
            } else {
                probe_walk(p, st, ord, tag, tmask, rs, slot, e, spare);
            }
        }
        __syncthreads();  // s_wneed / s_range are rewritten in the next iteration
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

// the nodes the probe appended in this round: order -> (global id, parent, action)
__global__ void __launch_bounds__(256) k_pbfs_commit3(const Pb3Params p) {
    if (p.ctl[CTL_STOP] != 0) return;
    const int64_t re = *p.round_end < p.node_cap ? *p.round_end : p.node_cap;
    const int64_t gid_base = p.ctl[CTL_GID_BASE];
    if (p.round_log && blockIdx.x == 0 && threadIdx.x == 0) {  // where the round's nodes are: the next level is grouped from it
        p.round_log[0] = *p.round_start;
        p.round_log[1] = re;
        p.round_log[2] = gid_base;
        p.round_log[3] = p.ctl[CTL_NODES] - gid_base;
    }
  for (int64_t place = *p.round_start + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; place < re; place += (int64_t)gridDim.x * blockDim.x) {
    const unsigned long long o = p.gid[place];
    if (o == ~0ull) continue;  // a hole: reserved, never claimed
    const int64_t ord = (int64_t)(o & ~PB_IN_ROUND);
    int64_t par;
    int act;
    if (((uint64_t)ord >> 32) == 0ull) {
        const uint32_t o32 = (uint32_t)ord, p32 = o32 / 12u;
        par = p32;
        act = (int)(o32 - 12u * p32);
    } else {
        par = ord / 12;
        act = (int)(ord % 12);
    }
    const unsigned m16 = (p.mask32[par >> 1] >> (16 * (int)(par & 1))) & 0xFFFFu;
    p.gid[place] = (unsigned long long)(gid_base + p.excl[par] + __popc(m16 & ((1u << act) - 1u)));
    p.parent[place] = p.parent_base + par;
    p.action[place] = (uint8_t)act;
  }
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
    static const int want = [] {  // ACX_EXPAND_OCC=3|4 resident CTAs per SM (80 / 64 registers): tuning knob
        const char* e = getenv("ACX_EXPAND_OCC");
        return e ? atoi(e) : 4;
    }();
    static const bool bulk = [] {  // ACX_EXPAND_BULK=0: per-thread 16-byte stores instead of one bulk async copy per run
        const char* e = getenv("ACX_EXPAND_BULK");
        return !(e && atoi(e) == 0);
    }();
    auto kern = bulk ? (want >= 4 ? k_pbfs_expand2<4, true> : k_pbfs_expand2<3, true>) : (want >= 4 ? k_pbfs_expand2<4, false> : k_pbfs_expand2<3, false>);
    static int ctas_per_sm[ACX_MAX_DEVICES] = {0};
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= ACX_MAX_DEVICES) dev = 0;
    if (ctas_per_sm[dev] == 0) {
        ACX_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        int occ = 0;
        ACX_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, EX_THREADS, smem));
        ctas_per_sm[dev] = occ > 0 ? occ : 1;
    }
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    int64_t grid = (f + EX_THREADS - 1) / EX_THREADS;
    if (grid > (int64_t)sms * ctas_per_sm[dev]) grid = (int64_t)sms * ctas_per_sm[dev];
    kern<<<(unsigned)grid, EX_THREADS, smem, (cudaStream_t)stream>>>(nodes, action, loc, gids, ord_base, f, m, cyclical, world, region_cap,
                                                                    reinterpret_cast<unsigned long long*>(cursor), dst, ctl);
    ACX_CUDA(cudaGetLastError());
    return ACX_OK;
}

extern "C" int acx_pbfs_probe3(const uint64_t* recs, const int64_t* region_counts_dev, int64_t count_stride, int n_regions, int64_t region_cap,
                               uint64_t* table, int64_t tcap, uint64_t* nodes, int64_t node_cap, int64_t* gid, uint64_t* n_local,
                               const int64_t* round_start_dev, uint64_t* mins, uint32_t* mask32, int64_t* ctl, acx_stream_t stream) {
    if (tcap < 2 || (tcap & (tcap - 1))) return acx_fail(ACX_ERR_ARG, "acx_pbfs_probe3: table capacity must be a power of two");
    if (!recs || !region_counts_dev || count_stride < 1 || n_regions < 1 || n_regions > PB_MAX_WORLD || region_cap < 1)
        return acx_fail(ACX_ERR_ARG, "acx_pbfs_probe3: need records, device region counts, 1 <= n_regions <= 32");
    if (!table || !nodes || !gid || !n_local || !round_start_dev || !mins || !mask32 || !ctl || node_cap < 1)
        return acx_fail(ACX_ERR_ARG, "acx_pbfs_probe3: null pointer");
    Pb3Params p = {};
    p.recs = recs; p.region_counts = region_counts_dev; p.count_stride = count_stride; p.nreg = n_regions; p.rcap = region_cap;
    p.table = table; p.tcap = tcap; p.nodes = nodes; p.node_cap = node_cap; p.gid = reinterpret_cast<unsigned long long*>(gid);
    p.n_local = reinterpret_cast<unsigned long long*>(n_local); p.round_start = round_start_dev;
    p.mins = reinterpret_cast<unsigned long long*>(mins); p.mask32 = mask32; p.ctl = ctl;
    static const int occ = [] {  // ACX_PROBE_OCC=2..8 resident CTAs of 256 threads per SM: tuning knob
        const char* e = getenv("ACX_PROBE_OCC");
        return e ? atoi(e) : 8;
    }();
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int per_sm = occ >= 8 ? 8 : (occ >= 6 ? 6 : (occ >= 4 ? 4 : 2));
    // persistent: the resident CTAs of the device, split over the regions (their record counts are near equal)
    int64_t gx = ((int64_t)sms * per_sm + n_regions - 1) / n_regions;
    if (gx > (region_cap + 255) / 256) gx = (region_cap + 255) / 256;
    const dim3 grid((unsigned)gx, (unsigned)n_regions);
    if (per_sm == 8) k_pbfs_probe3<256, 8><<<grid, 256, 0, (cudaStream_t)stream>>>(p);
    else if (per_sm == 6) k_pbfs_probe3<256, 6><<<grid, 256, 0, (cudaStream_t)stream>>>(p);
    else if (per_sm == 4) k_pbfs_probe3<256, 4><<<grid, 256, 0, (cudaStream_t)stream>>>(p);
    else k_pbfs_probe3<256, 2><<<grid, 256, 0, (cudaStream_t)stream>>>(p);
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

extern "C" int acx_pbfs_commit3(const int64_t* round_start_dev, const int64_t* round_end_dev, int64_t max_new, int64_t node_cap,
                                const int32_t* excl, const uint32_t* mask32, int64_t parent_base, const int64_t* ctl, int64_t* gid,
                                int64_t* parent, uint8_t* action, int64_t* round_log, acx_stream_t stream) {
    if (!round_start_dev || !round_end_dev || !excl || !mask32 || !ctl || !gid || !parent || !action || max_new < 1 || node_cap < 1)
        return acx_fail(ACX_ERR_ARG, "acx_pbfs_commit3: bad argument");
    Pb3Params p = {};
    p.round_start = round_start_dev; p.round_end = round_end_dev; p.node_cap = node_cap;
    p.excl = excl; p.mask32 = const_cast<unsigned*>(mask32); p.parent_base = parent_base;
    p.ctl = const_cast<int64_t*>(ctl); p.gid = reinterpret_cast<unsigned long long*>(gid); p.parent = parent; p.action = action;
    p.round_log = round_log;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    int64_t grid = (max_new + 255) / 256;  // the round's size is known on the device only: a grid-stride loop over it
    if (grid > (int64_t)sms * 8) grid = (int64_t)sms * 8;
    k_pbfs_commit3<<<(unsigned)grid, 256, 0, (cudaStream_t)stream>>>(p);
    ACX_CUDA(cudaGetLastError());
    return ACX_OK;
}
