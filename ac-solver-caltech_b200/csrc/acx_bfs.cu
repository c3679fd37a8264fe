// acx_bfs.cu -- level-synchronous BFS building blocks: a GPU open-addressing visited table with
// first-discovery-order semantics, used by the single-GPU and the hash-partitioned multi-GPU bfs.
//
// Reference semantics reproduced (ac_solver/search/breadth_first.py:15-97): the FIFO search
// discovers nodes in the order (parent in discovery order, action 0..11).  Within one level the
// candidates are numbered c = parent_rank * 12 + action, which IS that order; a candidate is new
// iff its state is neither in the table nor carried by a candidate with a smaller number.  New
// candidates get node ids by an exclusive scan of the `new` flags in c order (host side, a
// torch.cumsum), so ids == the reference's discovery order and the next frontier is a contiguous
// id range.  The budget rule and the `total length == 2` early exit are evaluated on the scanned
// flags (acx/bfs.py).
//
// Table slot (64 bit): 0 = empty, else  tag(16) | provisional(1) | value(47)
//   committed:   value = node index in the node store
//   provisional: value = candidate index of this round (states are compared against `cand`)
// Key equality is always decided by a full row compare; the tag only filters.
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>

#include "../../include/acx.h"
#include "acx_common.h"
#include "acx_rows.cuh"
#include "acx_expand_warp.cuh"

namespace acx {

static constexpr int ACX_BFS_MAX_WORLD = 32;
static constexpr uint64_t PROV_BIT = 1ull << 47;
static constexpr uint64_t VAL_MASK = PROV_BIT - 1;

__device__ __forceinline__ uint64_t bfs_row_hash(const uint32_t* w, int nwords) {
    uint64_t h = 0x9E3779B97F4A7C15ull;
    for (int k = 0; k < nwords; ++k) {
        h ^= w[k];
        h *= 0xFF51AFD7ED558CCDull;
        h ^= h >> 29;
    }
    return h;
}

__device__ __forceinline__ bool bfs_rows_equal(const uint32_t* a, const uint32_t* b, int nwords) {
    uint32_t d = 0;
    for (int k = 0; k < nwords; ++k) d |= a[k] ^ b[k];
    return d == 0u;
}

// Rows of any width (the single-GPU level engine, beam / value-guided search): 2m % 4 == 0 reads 32-bit words, other
// widths (AK(2) at max_relator_length = 7, the reference tests' example) assemble the same words from bytes, the last
// one zero-filled.  rowb is a kernel parameter, so the branch is uniform.
__device__ __forceinline__ uint32_t bfs_word_b(const int8_t* row, int k, int rowb) {
    uint32_t w = 0;
    for (int b = 0; b < 4; ++b) {
        const int i = 4 * k + b;
        if (i < rowb) w |= (uint32_t)(uint8_t)row[i] << (8 * b);
    }
    return w;
}

__device__ __forceinline__ uint64_t bfs_row_hash_any(const int8_t* row, int rowb) {
    if ((rowb & 3) == 0) return bfs_row_hash(reinterpret_cast<const uint32_t*>(row), rowb / 4);
    uint64_t h = 0x9E3779B97F4A7C15ull;
    for (int k = 0; k < (rowb + 3) / 4; ++k) {
        h ^= bfs_word_b(row, k, rowb);
        h *= 0xFF51AFD7ED558CCDull;
        h ^= h >> 29;
    }
    return h;
}

__device__ __forceinline__ bool bfs_rows_equal_any(const int8_t* a, const int8_t* b, int rowb) {
    if ((rowb & 3) == 0) return bfs_rows_equal(reinterpret_cast<const uint32_t*>(a), reinterpret_cast<const uint32_t*>(b), rowb / 4);
    int d = 0;
    for (int k = 0; k < rowb; ++k) d |= a[k] ^ b[k];
    return d == 0;
}

struct ProbeParams {
    const int8_t* cand;       // [n, rowb] candidate rows of this round
    const uint8_t* lens;      // [n, 2]
    const uint8_t* status;    // [n] row status from the expansion (may be null)
    const int64_t* order;     // [n] global discovery number of each candidate, or null (= its index)
    int64_t n;
    int rowb;
    uint64_t* table;
    int64_t tcap;             // power of two
    const int8_t* nodes;      // [*, rowb] committed node rows
    int64_t* slot_out;        // [n]  slot index; -1 = already visited; -2 = not a candidate (success / error row)
    unsigned long long* min_success;  // smallest order number with total length 2 (init: all ones)
    unsigned long long* min_error;    // smallest order number whose row status is non-zero (the reference raises there)
    uint64_t* hash_out;       // [n] optional: the row hash (owner partition key)
};

// phase 1: every candidate claims or joins a slot
__global__ void k_bfs_probe(const ProbeParams p) {
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= p.n) return;
    const int8_t* row = p.cand + j * p.rowb;
    const int64_t ord = p.order ? p.order[j] : j;
    if (p.status && p.status[j] != 0) {
        atomicMin(p.min_error, (unsigned long long)ord);
        p.slot_out[j] = -2;
        return;
    }
    if ((int)p.lens[2 * j] + (int)p.lens[2 * j + 1] == 2) {  // breadth_first.py:84-85: success before the visited test
        atomicMin(p.min_success, (unsigned long long)ord);
        p.slot_out[j] = -2;
        return;
    }
    const uint64_t h = bfs_row_hash_any(row, p.rowb);
    if (p.hash_out) p.hash_out[j] = h;
    const uint64_t tag = ((h >> 48) | 1ull) << 48;
    const uint64_t mine = tag | PROV_BIT | (uint64_t)j;
    const uint64_t mask = (uint64_t)p.tcap - 1;
    uint64_t slot = (h ^ (h >> 23)) & mask;
    for (;;) {
        uint64_t e = __ldcg(p.table + slot);
        if (e == 0ull) {
            const uint64_t old = atomicCAS(reinterpret_cast<unsigned long long*>(p.table + slot), 0ull, (unsigned long long)mine);
            if (old == 0ull) {
                p.slot_out[j] = (int64_t)slot;
                return;
            }
            e = old;  // somebody else took it: look at what is there now
        }
        if ((e >> 48) == (tag >> 48)) {
            if (e & PROV_BIT) {
                // another candidate of this round: same state?  then the smaller order number owns the slot
                for (;;) {
                    const int64_t other = (int64_t)(e & VAL_MASK);
                    if (other == j) {
                        p.slot_out[j] = (int64_t)slot;
                        return;
                    }
                    if (!bfs_rows_equal_any(row, p.cand + other * p.rowb, p.rowb)) break;
                    const int64_t oord = p.order ? p.order[other] : other;
                    if (oord < ord) {  // it stays; we are a duplicate within the round (resolved in phase 2)
                        p.slot_out[j] = (int64_t)slot;
                        return;
                    }
                    const uint64_t old = atomicCAS(reinterpret_cast<unsigned long long*>(p.table + slot), (unsigned long long)e,
                                                   (unsigned long long)mine);
                    if (old == e) {
                        p.slot_out[j] = (int64_t)slot;
                        return;
                    }
                    e = old;  // changed under us: it still carries an equal state (only equal states replace each other)
                    if (!(e & PROV_BIT)) break;
                }
            } else {
                const int64_t id = (int64_t)(e & VAL_MASK);
                if (bfs_rows_equal_any(row, p.nodes + id * p.rowb, p.rowb)) {
                    p.slot_out[j] = -1;
                    return;
                }
            }
        }
        slot = (slot + 1) & mask;
    }
}

// phase 2: a candidate is new iff it still owns its slot
__global__ void k_bfs_resolve(const uint64_t* table, const int64_t* slot, int64_t n, uint8_t* is_new) {
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const int64_t s = slot[j];
    uint8_t v = 0;
    if (s >= 0) {
        const uint64_t e = table[s];
        v = ((e & PROV_BIT) && (int64_t)(e & VAL_MASK) == j) ? 1 : 0;
    }
    is_new[j] = v;
}

struct CommitParams {
    const int8_t* cand;
    const uint8_t* lens;
    const int64_t* slot;
    const uint8_t* is_new;
    const int64_t* new_rank;  // [n] exclusive scan of is_new
    const int64_t* order;     // [n] discovery number of each candidate relative to parent_base*12, or null (= j)
    int64_t parent_base;
    int64_t n, limit;         // only candidates j < limit are committed
    int rowb;
    uint64_t* table;
    int8_t* nodes;            // row of new node k goes to nodes[(node_base + new_rank) * rowb]
    int64_t* parent;          // [*] parent node id (global numbering)
    uint8_t* action;
    uint16_t* total;
    int64_t node_base;
};

// phase 3: new candidates become nodes; their slots switch from provisional to committed
__global__ void k_bfs_commit(const CommitParams p) {
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= p.n) return;
    const int64_t s = p.slot[j];
    if (!p.is_new[j]) return;
    if (j >= p.limit) return;  // beyond the cut: the search ends here
    const int64_t id = p.node_base + p.new_rank[j];
    if ((p.rowb & 3) == 0) {
        const uint32_t* src = reinterpret_cast<const uint32_t*>(p.cand + j * p.rowb);
        uint32_t* dst = reinterpret_cast<uint32_t*>(p.nodes + id * p.rowb);
        for (int k = 0; k < p.rowb / 4; ++k) dst[k] = src[k];
    } else {
        for (int k = 0; k < p.rowb; ++k) p.nodes[id * p.rowb + k] = p.cand[j * p.rowb + k];
    }
    const int64_t ord = p.order ? p.order[j] : j;
    p.parent[id] = p.parent_base + ord / 12;
    p.action[id] = (uint8_t)(ord % 12);
    p.total[id] = (uint16_t)((int)p.lens[2 * j] + (int)p.lens[2 * j + 1]);
    const uint64_t e = p.table[s];
    p.table[s] = (e & ~(PROV_BIT | VAL_MASK)) | (uint64_t)id;
}

__global__ void k_rows_hash(const int8_t* rows, int64_t n, int rowb, uint64_t* out) {
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    out[j] = bfs_row_hash_any(rows + j * rowb, rowb);
}

// root insertion
__global__ void k_bfs_seed(const int8_t* row, int rowb, uint64_t* table, int64_t tcap, int64_t id) {
    const uint64_t h = bfs_row_hash_any(row, rowb);
    const uint64_t tag = ((h >> 48) | 1ull) << 48;
    const uint64_t mask = (uint64_t)tcap - 1;
    uint64_t slot = (h ^ (h >> 23)) & mask;
    while (table[slot] != 0ull) slot = (slot + 1) & mask;
    table[slot] = tag | (uint64_t)id;
}

// ---------------------------------------------------------------------------------------------
// Sharded bfs: candidate records travel between GPUs as  row | order (8) | lens (2) | status (1) | pad,
// rec = align8(rowb) + 16 bytes.  owner(state) = ((hash >> 20) & 0x7fffffff) % world.
__device__ __forceinline__ int bfs_owner(uint64_t h, int world) { return (int)(((h >> 20) & 0x7FFFFFFFull) % (uint64_t)world); }

// pass 1: owner of every candidate + per-owner counts (block histogram, one global atomic per owner per block)
__global__ void k_bfs_bucket_count(const int8_t* cand, int64_t n, int rowb, int world, uint8_t* owner, unsigned long long* counts) {
    __shared__ unsigned int s_cnt[64];
    if (threadIdx.x < 64) s_cnt[threadIdx.x] = 0u;
    __syncthreads();
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j < n) {
        const int o = bfs_owner(bfs_row_hash(reinterpret_cast<const uint32_t*>(cand + j * rowb), rowb / 4), world);
        owner[j] = (uint8_t)o;
        atomicAdd(&s_cnt[o], 1u);
    }
    __syncthreads();
    if (threadIdx.x < world && s_cnt[threadIdx.x]) atomicAdd(&counts[threadIdx.x], (unsigned long long)s_cnt[threadIdx.x]);
}

// pass 2: scatter the records into the send buffer, grouped by owner (order inside a group is free:
// order numbers travel with the rows and the receiver resolves first discovery by them)
__global__ void k_bfs_bucket_scatter(const int8_t* cand, const uint8_t* lens, const uint8_t* status, const int64_t* parent_ord,
                                     const uint8_t* owner, int64_t n, int rowb, int rec, int world, const int64_t* offsets,
                                     unsigned long long* cursor, uint8_t* send) {
    __shared__ unsigned int s_cnt[64];
    __shared__ unsigned long long s_base[64];
    if (threadIdx.x < 64) s_cnt[threadIdx.x] = 0u;
    __syncthreads();
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int o = 0;
    unsigned int r = 0;
    if (j < n) {
        o = owner[j];
        r = atomicAdd(&s_cnt[o], 1u);
    }
    __syncthreads();
    if (threadIdx.x < world && s_cnt[threadIdx.x])
        s_base[threadIdx.x] = atomicAdd(&cursor[threadIdx.x], (unsigned long long)s_cnt[threadIdx.x]);
    __syncthreads();
    if (j >= n) return;
    uint8_t* dst = send + ((int64_t)offsets[o] + (int64_t)s_base[o] + r) * rec;
    const uint32_t* src = reinterpret_cast<const uint32_t*>(cand + j * rowb);
    uint32_t* d4 = reinterpret_cast<uint32_t*>(dst);
    for (int k = 0; k < rowb / 4; ++k) d4[k] = src[k];
    const int off = (rowb + 7) / 8 * 8;
    if (off != rowb) d4[rowb / 4] = 0u;
    *reinterpret_cast<int64_t*>(dst + off) = parent_ord[j / 12] * 12 + (j % 12);
    *reinterpret_cast<uint32_t*>(dst + off + 8) = (uint32_t)lens[2 * j] | ((uint32_t)lens[2 * j + 1] << 8) | ((uint32_t)status[j] << 16);
    for (int k = off + 12; k < rec; k += 4) *reinterpret_cast<uint32_t*>(dst + k) = 0u;
}

struct PackedParams {
    const uint8_t* recs;  // nreg regions of rcap records each; region s holds pre[s+1] - pre[s] records
    int64_t n;            // records in all regions
    int nreg;
    int64_t rcap;
    int64_t pre[ACX_BFS_MAX_WORLD + 1];
    int rowb, rec;
    uint64_t* table;
    int64_t tcap;
    const int8_t* nodes;
    int64_t* slot;        // [n] scratch
    unsigned long long* mins;  // [2] min success order, min error order
    int* mask;            // [C] per-parent 12-bit "new" mask of this round (atomicOr)
    // commit
    const int64_t* excl_par;  // [C] exclusive scan of the global per-parent new counts
    const int* gmask;         // [C] global masks (after the all-reduce)
    int64_t parent_base, gid_base;
    unsigned long long* n_local;  // device counter of local nodes
    int8_t* nodes_w;
    int64_t* gid;
    int64_t* parent;
    uint8_t* action;
    uint16_t* total;
};

// linear thread index -> record index (regions are filled from their start, with a gap behind each)
__device__ __forceinline__ int64_t packed_index(const PackedParams& p, int64_t t) {
    int s = 0;
    while (s + 1 < p.nreg && t >= p.pre[s + 1]) ++s;
    return (int64_t)s * p.rcap + (t - p.pre[s]);
}

__device__ __forceinline__ int64_t rec_order(const PackedParams& p, int64_t j) {
    return *reinterpret_cast<const int64_t*>(p.recs + j * p.rec + (p.rowb + 7) / 8 * 8);
}

__global__ void k_bfs_probe_packed(const PackedParams p) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= p.n) return;
    const int64_t j = packed_index(p, t);
    const int nwords = p.rowb / 4, off = (p.rowb + 7) / 8 * 8;
    const uint8_t* rec = p.recs + j * p.rec;
    const uint32_t* row = reinterpret_cast<const uint32_t*>(rec);
    const int64_t ord = *reinterpret_cast<const int64_t*>(rec + off);
    const uint32_t meta = *reinterpret_cast<const uint32_t*>(rec + off + 8);
    if ((meta >> 16) & 0xFFu) {
        atomicMin(p.mins + 1, (unsigned long long)ord);
        p.slot[j] = -2;
        return;
    }
    if ((int)(meta & 0xFFu) + (int)((meta >> 8) & 0xFFu) == 2) {
        atomicMin(p.mins, (unsigned long long)ord);
        p.slot[j] = -2;
        return;
    }
    const uint64_t h = bfs_row_hash(row, nwords);
    const uint64_t tag = ((h >> 48) | 1ull) << 48;
    const uint64_t mine = tag | PROV_BIT | (uint64_t)j;
    const uint64_t mask = (uint64_t)p.tcap - 1;
    uint64_t slot = (h ^ (h >> 23)) & mask;
    for (;;) {
        uint64_t e = __ldcg(p.table + slot);
        if (e == 0ull) {
            const uint64_t old = atomicCAS(reinterpret_cast<unsigned long long*>(p.table + slot), 0ull, (unsigned long long)mine);
            if (old == 0ull) {
                p.slot[j] = (int64_t)slot;
                return;
            }
            e = old;
        }
        if ((e >> 48) == (tag >> 48)) {
            if (e & PROV_BIT) {
                for (;;) {
                    const int64_t other = (int64_t)(e & VAL_MASK);
                    if (other == j) {
                        p.slot[j] = (int64_t)slot;
                        return;
                    }
                    if (!bfs_rows_equal(row, reinterpret_cast<const uint32_t*>(p.recs + other * p.rec), nwords)) break;
                    if (rec_order(p, other) < ord) {
                        p.slot[j] = (int64_t)slot;
                        return;
                    }
                    const uint64_t old = atomicCAS(reinterpret_cast<unsigned long long*>(p.table + slot), (unsigned long long)e,
                                                   (unsigned long long)mine);
                    if (old == e) {
                        p.slot[j] = (int64_t)slot;
                        return;
                    }
                    e = old;
                    if (!(e & PROV_BIT)) break;
                }
            } else {
                const int64_t id = (int64_t)(e & VAL_MASK);
                if (bfs_rows_equal(row, reinterpret_cast<const uint32_t*>(p.nodes + id * p.rowb), nwords)) {
                    p.slot[j] = -1;
                    return;
                }
            }
        }
        slot = (slot + 1) & mask;
    }
}

// new candidates set their bit in the per-parent mask of the round
__global__ void k_bfs_resolve_packed(const PackedParams p) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= p.n) return;
    const int64_t j = packed_index(p, t);
    const int64_t s = p.slot[j];
    if (s < 0) return;
    const uint64_t e = p.table[s];
    if ((e & PROV_BIT) && (int64_t)(e & VAL_MASK) == j) {
        const int64_t ord = rec_order(p, j);
        atomicOr(p.mask + ord / 12, 1 << (int)(ord % 12));
    } else {
        p.slot[j] = -1;  // a duplicate inside the round
    }
}

// new candidates become local nodes; their global id follows from the all-reduced masks
__global__ void k_bfs_commit_packed(const PackedParams p) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= p.n) return;
    const int64_t j = packed_index(p, t);
    const int64_t s = p.slot[j];
    if (s < 0) return;
    const int off = (p.rowb + 7) / 8 * 8;
    const uint8_t* rec = p.recs + j * p.rec;
    const int64_t ord = *reinterpret_cast<const int64_t*>(rec + off);
    const uint32_t meta = *reinterpret_cast<const uint32_t*>(rec + off + 8);
    const int64_t par = ord / 12;
    const int act = (int)(ord % 12);
    const int64_t gid = p.gid_base + p.excl_par[par] + __popc((unsigned)p.gmask[par] & ((1u << act) - 1u));
    const int64_t id = (int64_t)atomicAdd(p.n_local, 1ull);
    const uint32_t* src = reinterpret_cast<const uint32_t*>(rec);
    uint32_t* dst = reinterpret_cast<uint32_t*>(p.nodes_w + id * p.rowb);
    for (int k = 0; k < p.rowb / 4; ++k) dst[k] = src[k];
    p.gid[id] = gid;
    p.parent[id] = p.parent_base + par;
    p.action[id] = (uint8_t)act;
    p.total[id] = (uint16_t)((meta & 0xFFu) + ((meta >> 8) & 0xFFu));
    const uint64_t e = p.table[s];
    p.table[s] = (e & ~(PROV_BIT | VAL_MASK)) | (uint64_t)id;
}

// Fused frontier step of the sharded bfs: gather 32 parents per warp by index, apply the 12 moves on bit planes
// (same row code as k_expand_warp), build the packed record of every child in shared memory, hash it, pick its
// owner rank and copy the record straight into that owner's region of the send buffer.  Slots are reserved per
// warp and per group of GM moves: one ballot scan per owner, one atomicAdd per owner per warp-step.
// send = [world][region_cap] records; cursor[o] = records written to region o; cursor[world] = overflow flag.
struct BucketDst {
    uint8_t* region[ACX_BFS_MAX_WORLD];  // this sender's region in the receive buffer of each owner (local or peer memory)
};

template <int MS, int GM, int WARPS>
__global__ void __launch_bounds__(32 * WARPS) k_bfs_expand_bucket(const int8_t* nodes, const int64_t* loc, const int64_t* gids,
                                                                  int64_t ord_base, int64_t f, int cyclical, int world,
                                                                  int64_t region_cap, unsigned long long* cursor,
                                                                  const BucketDst dst, int64_t ntiles) {
    using Ops = PlaneOps<2, MS>;
    using R = R64;
    constexpr int ROWB = 2 * MS, REC = (ROWB + 16 + 15) / 16 * 16, PCH = ROWB / 8, RCH = REC / 16;
    static_assert(ROWB % 8 == 0, "rows must be whole 8-byte words");
    // records are whole 32-byte sectors (96 B at m = 36) and leave in 16-byte stores: full-sector writes over NVLink
    constexpr int CSTRIDE = GM * REC + 8;  // CSTRIDE / 8 odd: conflict-free 64-bit accesses across lanes
    constexpr int WARP_BYTES = 32 * (2 * ROWB + CSTRIDE) + 32 * GM * 4;
    extern __shared__ __align__(128) uint8_t smem[];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint8_t* par = smem + (size_t)wid * WARP_BYTES;
    uint8_t* base = par + 32 * ROWB;
    uint8_t* cbuf = base + 32 * ROWB;
    uint32_t* spos = reinterpret_cast<uint32_t*>(cbuf + 32 * CSTRIDE);
    const bool cyc = cyclical != 0;
    const unsigned ltmask = (1u << lane) - 1u;
    const int64_t warp0 = (int64_t)blockIdx.x * WARPS + wid, nwarps = (int64_t)gridDim.x * WARPS;
    const uint8_t* gn = reinterpret_cast<const uint8_t*>(nodes);

    for (int64_t tile = warp0; tile < ntiles; tile += nwarps) {
        const int64_t f0 = tile * 32;
        const int nrows = (int)min((int64_t)32, f - f0);
        for (int c = lane; c < nrows * PCH; c += 32) {
            const int pr = c / PCH, o = c - pr * PCH;
            *reinterpret_cast<uint2*>(par + pr * ROWB + o * 8) =
                *reinterpret_cast<const uint2*>(gn + loc[f0 + pr] * (int64_t)ROWB + o * 8);
        }
        __syncwarp();
        const bool active = lane < nrows;
        const int64_t ord0 = active ? (gids[f0 + lane] - ord_base) * 12 : 0;
        uint8_t* prow = par + lane * ROWB;
        uint8_t* brow = base + lane * ROWB;
        R r[2], rc[2];
        int cnt[2] = {0, 0};
        int cls = ROW_INVALID;
        if (active) cls = Ops::classify(prow, r, cnt);
        const bool fast = active && cls == ROW_FAST;
        if (fast) {
            for (int k = 0; k < PCH; ++k) reinterpret_cast<uint2*>(brow)[k] = reinterpret_cast<const uint2*>(prow)[k];
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
            int own[GM];
            unsigned rank[GM];
#pragma unroll
            for (int j = 0; j < GM; ++j) {
                const int a = g * GM + j;
                uint8_t* child = cbuf + lane * CSTRIDE + j * REC;
                own[j] = -1;
                rank[j] = 0;
                if (!active) continue;
                int l0 = cnt[0], l1 = cnt[1];
                uint32_t st = ACX_ROW_OK;
                bool moved = true;
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
                    moved = changed;
                    if (x.n == 0) {
                        st = ACX_ROW_EMPTY_RELATOR;
                        for (int k = 0; k < PCH; ++k) reinterpret_cast<uint2*>(child)[k] = reinterpret_cast<const uint2*>(prow)[k];
                    } else {
                        for (int k = 0; k < PCH; ++k) reinterpret_cast<uint2*>(child)[k] = reinterpret_cast<const uint2*>(brow)[k];
                        if (changed) {
                            if (cyc) Ops::cyclic(x);
                            Ops::unpack(child + (d & 1u) * MS, x);
                            const R& old = i ? rc[1] : rc[0];
                            moved = !(x.n == old.n && x.S == old.S && x.T == old.T);
                        }
                        const int na = changed ? x.n : (i ? rc[1].n : rc[0].n);
                        l0 = i ? rc[0].n : na;
                        l1 = i ? na : rc[1].n;
                    }
                } else {
                    for (int k = 0; k < PCH; ++k) reinterpret_cast<uint2*>(child)[k] = reinterpret_cast<const uint2*>(prow)[k];
                    acmove_row64<MS>(child, MS, a, cyc, l0, l1, st);
                }
                // a move that leaves the presentation untouched (e.g. the product would exceed max_relator_length)
                // reproduces the parent, which is already in the visited set: nothing to send
                if (fast && st == ACX_ROW_OK && !moved && rc[0].n == r[0].n && rc[1].n == r[1].n) continue;
                *reinterpret_cast<int64_t*>(child + ROWB) = ord0 + a;
                *reinterpret_cast<uint2*>(child + ROWB + 8) = make_uint2((uint32_t)l0 | ((uint32_t)l1 << 8) | (st << 16), 0u);
                for (int k = ROWB + 16; k < REC; k += 8) *reinterpret_cast<uint2*>(child + k) = make_uint2(0u, 0u);
                own[j] = bfs_owner(bfs_row_hash(reinterpret_cast<const uint32_t*>(child), ROWB / 4), world);
            }
            // slots: per owner, rank of each child among the warp-step's children of that owner
            unsigned mycnt = 0;
            for (int o = 0; o < world; ++o) {
                unsigned run = 0;
#pragma unroll
                for (int j = 0; j < GM; ++j) {
                    const unsigned b = __ballot_sync(0xFFFFFFFFu, own[j] == o);
                    if (own[j] == o) rank[j] = run + __popc(b & ltmask);
                    run += __popc(b);
                }
                if (lane == o) mycnt = run;
            }
            unsigned long long mybase = 0;
            if (lane < world && mycnt) mybase = atomicAdd(cursor + lane, (unsigned long long)mycnt);
#pragma unroll
            for (int j = 0; j < GM; ++j) {
                const unsigned long long b = __shfl_sync(0xFFFFFFFFu, mybase, own[j] < 0 ? 0 : own[j]);
                uint32_t idx = 0xFFFFFFFFu;
                if (own[j] >= 0) {
                    const unsigned long long pos = b + rank[j];
                    if (pos < (unsigned long long)region_cap) idx = ((uint32_t)own[j] << 27) | (uint32_t)pos;
                    else cursor[world] = 1ull;  // a region overflowed: the host raises
                }
                spos[lane * GM + j] = idx;
            }
            __syncwarp();
            for (int it = lane; it < nrows * GM * RCH; it += 32) {
                const int c = it / RCH, k = it - c * RCH;
                const int pr = c / GM, j = c - pr * GM;
                const uint32_t idx = spos[c];
                if (idx != 0xFFFFFFFFu) {  // local or NVLink peer store
                    const uint2* src = reinterpret_cast<const uint2*>(cbuf + pr * CSTRIDE + j * REC + 16 * k);
                    const uint2 lo = src[0], hi = src[1];
                    *reinterpret_cast<uint4*>(dst.region[idx >> 27] + (size_t)(idx & 0x7FFFFFFu) * REC + 16 * k) =
                        make_uint4(lo.x, lo.y, hi.x, hi.y);
                }
            }
            __syncwarp();
        }
    }
}

template <int MS, int GM, int WARPS>
static int launch_expand_bucket(const int8_t* nodes, const int64_t* loc, const int64_t* gids, int64_t ord_base, int64_t f,
                                int cyclical, int world, int64_t region_cap, uint64_t* cursor, const BucketDst& dst,
                                cudaStream_t st) {
    constexpr int ROWB = 2 * MS, REC = (ROWB + 16 + 15) / 16 * 16, CSTRIDE = GM * REC + 8;
    constexpr size_t smem = (size_t)WARPS * (32 * (2 * ROWB + CSTRIDE) + 32 * GM * 4);
    auto kern = k_bfs_expand_bucket<MS, GM, WARPS>;
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
    const int64_t ntiles = (f + 31) / 32;
    int64_t grid = (ntiles + WARPS - 1) / WARPS;
    if (grid > (int64_t)sms * ctas_per_sm[dev]) grid = (int64_t)sms * ctas_per_sm[dev];
    kern<<<(unsigned)grid, 32 * WARPS, smem, st>>>(nodes, loc, gids, ord_base, f, cyclical, world, region_cap,
                                                   reinterpret_cast<unsigned long long*>(cursor), dst, ntiles);
    ACX_CUDA(cudaGetLastError());
    return ACX_OK;
}

}  // namespace acx

using namespace acx;

extern "C" int acx_bfs_seed(const int8_t* row, int m, uint64_t* table, int64_t tcap, int64_t node_id, acx_stream_t stream) {
    if (!row || !table || m < 1 || tcap < 2 || (tcap & (tcap - 1))) return acx_fail(ACX_ERR_ARG, "acx_bfs_seed: bad argument (tcap must be a power of two)");
    k_bfs_seed<<<1, 1, 0, (cudaStream_t)stream>>>(row, 2 * m, table, tcap, node_id);
    ACX_CUDA(cudaGetLastError());
    return ACX_OK;
}

extern "C" int acx_bfs_probe(const int8_t* cand, const uint8_t* lens, const uint8_t* status, const int64_t* order, int64_t n,
                             int m, uint64_t* table, int64_t tcap, const int8_t* nodes, int64_t* slot_out,
                             uint64_t* min_success, uint64_t* min_error, uint64_t* hash_out, uint8_t* is_new_out, acx_stream_t stream) {
    if (n < 0 || m < 1 || tcap < 2 || (tcap & (tcap - 1))) return acx_fail(ACX_ERR_ARG, "acx_bfs_probe: bad argument (tcap must be a power of two)");
    if (n == 0) return ACX_OK;
    if (!cand || !lens || !table || !nodes || !slot_out || !min_success || !min_error || !is_new_out) return acx_fail(ACX_ERR_ARG, "acx_bfs_probe: null pointer");
    ProbeParams p;
    p.cand = cand; p.lens = lens; p.status = status; p.order = order; p.n = n; p.rowb = 2 * m; p.table = table; p.tcap = tcap;
    p.nodes = nodes; p.slot_out = slot_out; p.min_success = reinterpret_cast<unsigned long long*>(min_success);
    p.min_error = reinterpret_cast<unsigned long long*>(min_error);
    p.hash_out = hash_out;
    cudaStream_t st = (cudaStream_t)stream;
    const unsigned grid = (unsigned)((n + 255) / 256);
    k_bfs_probe<<<grid, 256, 0, st>>>(p);
    k_bfs_resolve<<<grid, 256, 0, st>>>(table, slot_out, n, is_new_out);
    ACX_CUDA(cudaGetLastError());
    return ACX_OK;
}

extern "C" int acx_bfs_commit(const int8_t* cand, const uint8_t* lens, const int64_t* slot, const uint8_t* is_new,
                              const int64_t* new_rank, const int64_t* order, int64_t parent_base, int64_t n, int64_t limit,
                              int m, uint64_t* table, int8_t* nodes, int64_t* parent, uint8_t* action, uint16_t* total,
                              int64_t node_base, acx_stream_t stream) {
    if (n < 0 || m < 1) return acx_fail(ACX_ERR_ARG, "acx_bfs_commit: bad argument");
    if (n == 0) return ACX_OK;
    if (!cand || !lens || !slot || !is_new || !new_rank || !table || !nodes || !parent || !action || !total) return acx_fail(ACX_ERR_ARG, "acx_bfs_commit: null pointer");
    CommitParams p;
    p.cand = cand; p.lens = lens; p.slot = slot; p.is_new = is_new; p.new_rank = new_rank; p.order = order;
    p.parent_base = parent_base; p.n = n; p.limit = limit; p.rowb = 2 * m; p.table = table; p.nodes = nodes; p.parent = parent;
    p.action = action; p.total = total; p.node_base = node_base;
    k_bfs_commit<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(p);
    ACX_CUDA(cudaGetLastError());
    return ACX_OK;
}

extern "C" int acx_rows_hash(const int8_t* rows, int64_t n, int m, uint64_t* hash_out, acx_stream_t stream) {
    if (n < 0 || m < 1) return acx_fail(ACX_ERR_ARG, "acx_rows_hash: bad argument");
    if (n == 0) return ACX_OK;
    if (!rows || !hash_out) return acx_fail(ACX_ERR_ARG, "acx_rows_hash: null pointer");
    k_rows_hash<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(rows, n, 2 * m, hash_out);
    ACX_CUDA(cudaGetLastError());
    return ACX_OK;
}

extern "C" int acx_bfs_record_bytes(int m) { return ((2 * m + 7) / 8 * 8 + 16 + 15) / 16 * 16; }

extern "C" int acx_bfs_bucket(const int8_t* cand, const uint8_t* lens, const uint8_t* status, const int64_t* parent_ord,
                              int64_t n, int m, int world, uint8_t* owner_scratch, uint64_t* counts, int64_t* offsets,
                              uint64_t* cursor, uint8_t* send, int phase, acx_stream_t stream) {
    if (n < 0 || m < 1 || (2 * m) % 4 != 0 || world < 1 || world > 64) return acx_fail(ACX_ERR_ARG, "acx_bfs_bucket: bad argument (m even, 1 <= world <= 64)");
    if (n == 0) return ACX_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const unsigned grid = (unsigned)((n + 255) / 256);
    if (phase == 0) {
        if (!cand || !owner_scratch || !counts) return acx_fail(ACX_ERR_ARG, "acx_bfs_bucket: null pointer");
        k_bfs_bucket_count<<<grid, 256, 0, st>>>(cand, n, 2 * m, world, owner_scratch, reinterpret_cast<unsigned long long*>(counts));
    } else {
        if (!cand || !lens || !status || !parent_ord || !owner_scratch || !offsets || !cursor || !send) return acx_fail(ACX_ERR_ARG, "acx_bfs_bucket: null pointer");
        k_bfs_bucket_scatter<<<grid, 256, 0, st>>>(cand, lens, status, parent_ord, owner_scratch, n, 2 * m, acx_bfs_record_bytes(m), world,
                                                   offsets, reinterpret_cast<unsigned long long*>(cursor), send);
    }
    ACX_CUDA(cudaGetLastError());
    return ACX_OK;
}

static int fill_regions(PackedParams& p, const int64_t* region_counts, int n_regions, int64_t region_cap, const char* who) {
    if (!region_counts || n_regions < 1 || n_regions > ACX_BFS_MAX_WORLD || region_cap < 0)
        return acx_fail(ACX_ERR_ARG, "%s: need 1 <= n_regions <= 32 and a host array of region counts", who);
    p.nreg = n_regions;
    p.rcap = region_cap;
    p.pre[0] = 0;
    for (int s = 0; s < n_regions; ++s) {
        if (region_counts[s] < 0 || region_counts[s] > region_cap) return acx_fail(ACX_ERR_ARG, "%s: region count out of range", who);
        p.pre[s + 1] = p.pre[s] + region_counts[s];
    }
    p.n = p.pre[n_regions];
    return ACX_OK;
}

extern "C" int acx_bfs_probe_packed(const uint8_t* recs, const int64_t* region_counts, int n_regions, int64_t region_cap, int m,
                                    uint64_t* table, int64_t tcap, const int8_t* nodes, int64_t* slot, uint64_t* mins,
                                    int32_t* mask, acx_stream_t stream) {
    if (m < 1 || (2 * m) % 4 != 0 || tcap < 2 || (tcap & (tcap - 1))) return acx_fail(ACX_ERR_ARG, "acx_bfs_probe_packed: bad argument");
    PackedParams p = {};
    if (int rc = fill_regions(p, region_counts, n_regions, region_cap, "acx_bfs_probe_packed")) return rc;
    if (p.n == 0) return ACX_OK;
    if (!recs || !table || !nodes || !slot || !mins || !mask) return acx_fail(ACX_ERR_ARG, "acx_bfs_probe_packed: null pointer");
    p.recs = recs; p.rowb = 2 * m; p.rec = acx_bfs_record_bytes(m); p.table = table; p.tcap = tcap; p.nodes = nodes;
    p.slot = slot; p.mins = reinterpret_cast<unsigned long long*>(mins); p.mask = mask;
    const unsigned grid = (unsigned)((p.n + 255) / 256);
    k_bfs_probe_packed<<<grid, 256, 0, (cudaStream_t)stream>>>(p);
    k_bfs_resolve_packed<<<grid, 256, 0, (cudaStream_t)stream>>>(p);
    ACX_CUDA(cudaGetLastError());
    return ACX_OK;
}

extern "C" int acx_bfs_commit_packed(const uint8_t* recs, const int64_t* region_counts, int n_regions, int64_t region_cap, int m,
                                     const int64_t* slot, const int64_t* excl_par, const int32_t* gmask, int64_t parent_base,
                                     int64_t gid_base, uint64_t* n_local, uint64_t* table, int8_t* nodes, int64_t* gid,
                                     int64_t* parent, uint8_t* action, uint16_t* total, acx_stream_t stream) {
    if (m < 1 || (2 * m) % 4 != 0) return acx_fail(ACX_ERR_ARG, "acx_bfs_commit_packed: bad argument");
    PackedParams p = {};
    if (int rc = fill_regions(p, region_counts, n_regions, region_cap, "acx_bfs_commit_packed")) return rc;
    if (p.n == 0) return ACX_OK;
    if (!recs || !slot || !excl_par || !gmask || !n_local || !table || !nodes || !gid || !parent || !action || !total) return acx_fail(ACX_ERR_ARG, "acx_bfs_commit_packed: null pointer");
    p.recs = recs; p.rowb = 2 * m; p.rec = acx_bfs_record_bytes(m); p.table = table; p.slot = const_cast<int64_t*>(slot);
    p.excl_par = excl_par; p.gmask = gmask; p.parent_base = parent_base; p.gid_base = gid_base;
    p.n_local = reinterpret_cast<unsigned long long*>(n_local); p.nodes_w = nodes; p.gid = gid; p.parent = parent; p.action = action;
    p.total = total;
    k_bfs_commit_packed<<<(unsigned)((p.n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(p);
    ACX_CUDA(cudaGetLastError());
    return ACX_OK;
}

extern "C" int acx_bfs_expand_bucket(const int8_t* nodes, const int64_t* loc, const int64_t* gids, int64_t ord_base, int64_t f,
                                     int m, int cyclical, int world, int64_t region_cap, uint64_t* cursor,
                                     uint8_t* const* dst_regions, acx_stream_t stream) {
    if (f < 0 || world < 1 || world > ACX_BFS_MAX_WORLD || region_cap < 1 || region_cap >= (1ll << 27))
        return acx_fail(ACX_ERR_ARG, "acx_bfs_expand_bucket: bad argument (1 <= world <= 32, region_cap < 2^27)");
    if (m != 36) return acx_fail(ACX_ERR_ARG, "acx_bfs_expand_bucket: built for max_relator_length 36 (use acx_expand12 + acx_bfs_bucket otherwise)");
    if (f == 0) return ACX_OK;
    if (!nodes || !loc || !gids || !cursor || !dst_regions) return acx_fail(ACX_ERR_ARG, "acx_bfs_expand_bucket: null pointer");
    BucketDst dst = {};
    for (int o = 0; o < world; ++o) {
        if (!dst_regions[o] || ((uintptr_t)dst_regions[o] & 15u)) return acx_fail(ACX_ERR_ARG, "acx_bfs_expand_bucket: destination regions must be non-null and 16-byte aligned");
        dst.region[o] = dst_regions[o];
    }
    if ((uintptr_t)nodes & 7u) return acx_fail(ACX_ERR_ARG, "acx_bfs_expand_bucket: nodes must be 8-byte aligned");
    return launch_expand_bucket<36, 4, 4>(nodes, loc, gids, ord_base, f, cyclical, world, region_cap, cursor, dst, (cudaStream_t)stream);
}

// Peer memory: buffers another process on the same NVLink domain may write into (CUDA IPC).
extern "C" int acx_peer_alloc(int64_t bytes, void** ptr) {
    if (bytes <= 0 || !ptr) return acx_fail(ACX_ERR_ARG, "acx_peer_alloc: bad argument");
    ACX_CUDA(cudaMalloc(ptr, (size_t)bytes));
    return ACX_OK;
}
extern "C" int acx_peer_free(void* ptr) {
    if (ptr) ACX_CUDA(cudaFree(ptr));
    return ACX_OK;
}
extern "C" int acx_peer_export(const void* ptr, uint8_t* handle64) {
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
    if (!ptr || !handle64) return acx_fail(ACX_ERR_ARG, "acx_peer_export: null pointer");
    cudaIpcMemHandle_t h;
    ACX_CUDA(cudaIpcGetMemHandle(&h, const_cast<void*>(ptr)));
    memcpy(handle64, &h, 64);
    return ACX_OK;
}
extern "C" int acx_peer_open(const uint8_t* handle64, void** ptr) {
    if (!ptr || !handle64) return acx_fail(ACX_ERR_ARG, "acx_peer_open: null pointer");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    ACX_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return ACX_OK;
}
extern "C" int acx_peer_close(void* ptr) {
    if (ptr) ACX_CUDA(cudaIpcCloseMemHandle(ptr));
    return ACX_OK;
}
