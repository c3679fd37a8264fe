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

#include "../../include/acx.h"
#include "acx_common.h"

namespace acx {

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
    const int nwords = p.rowb / 4;
    const uint32_t* row = reinterpret_cast<const uint32_t*>(p.cand + j * p.rowb);
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
    const uint64_t h = bfs_row_hash(row, nwords);
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
                    if (!bfs_rows_equal(row, reinterpret_cast<const uint32_t*>(p.cand + other * p.rowb), nwords)) break;
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
                if (bfs_rows_equal(row, reinterpret_cast<const uint32_t*>(p.nodes + id * p.rowb), nwords)) {
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
    const int nwords = p.rowb / 4;
    const uint32_t* src = reinterpret_cast<const uint32_t*>(p.cand + j * p.rowb);
    uint32_t* dst = reinterpret_cast<uint32_t*>(p.nodes + id * p.rowb);
    for (int k = 0; k < nwords; ++k) dst[k] = src[k];
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
    out[j] = bfs_row_hash(reinterpret_cast<const uint32_t*>(rows + j * rowb), rowb / 4);
}

// root insertion
__global__ void k_bfs_seed(const int8_t* row, int rowb, uint64_t* table, int64_t tcap, int64_t id) {
    const uint32_t* w = reinterpret_cast<const uint32_t*>(row);
    const uint64_t h = bfs_row_hash(w, rowb / 4);
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
    *reinterpret_cast<uint32_t*>(dst + off + 12) = 0u;
}

struct PackedParams {
    const uint8_t* recs;  // [n, rec]
    int64_t n;
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

__device__ __forceinline__ int64_t rec_order(const PackedParams& p, int64_t j) {
    return *reinterpret_cast<const int64_t*>(p.recs + j * p.rec + (p.rowb + 7) / 8 * 8);
}

__global__ void k_bfs_probe_packed(const PackedParams p) {
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= p.n) return;
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
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= p.n) return;
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
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= p.n) return;
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

}  // namespace acx

using namespace acx;

extern "C" int acx_bfs_seed(const int8_t* row, int m, uint64_t* table, int64_t tcap, int64_t node_id, acx_stream_t stream) {
    if (!row || !table || m < 1 || (2 * m) % 4 != 0 || tcap < 2 || (tcap & (tcap - 1))) return acx_fail(ACX_ERR_ARG, "acx_bfs_seed: bad argument (m must be even, tcap a power of two)");
    k_bfs_seed<<<1, 1, 0, (cudaStream_t)stream>>>(row, 2 * m, table, tcap, node_id);
    ACX_CUDA(cudaGetLastError());
    return ACX_OK;
}

extern "C" int acx_bfs_probe(const int8_t* cand, const uint8_t* lens, const uint8_t* status, const int64_t* order, int64_t n,
                             int m, uint64_t* table, int64_t tcap, const int8_t* nodes, int64_t* slot_out,
                             uint64_t* min_success, uint64_t* min_error, uint64_t* hash_out, uint8_t* is_new_out, acx_stream_t stream) {
    if (n < 0 || m < 1 || (2 * m) % 4 != 0 || tcap < 2 || (tcap & (tcap - 1))) return acx_fail(ACX_ERR_ARG, "acx_bfs_probe: bad argument (m must be even, tcap a power of two)");
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
    if (n < 0 || m < 1 || (2 * m) % 4 != 0) return acx_fail(ACX_ERR_ARG, "acx_bfs_commit: bad argument");
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
    if (n < 0 || m < 1 || (2 * m) % 4 != 0) return acx_fail(ACX_ERR_ARG, "acx_rows_hash: bad argument (m must be even)");
    if (n == 0) return ACX_OK;
    if (!rows || !hash_out) return acx_fail(ACX_ERR_ARG, "acx_rows_hash: null pointer");
    k_rows_hash<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(rows, n, 2 * m, hash_out);
    ACX_CUDA(cudaGetLastError());
    return ACX_OK;
}

extern "C" int acx_bfs_record_bytes(int m) { return (2 * m + 7) / 8 * 8 + 16; }

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

extern "C" int acx_bfs_probe_packed(const uint8_t* recs, int64_t n, int m, uint64_t* table, int64_t tcap, const int8_t* nodes,
                                    int64_t* slot, uint64_t* mins, int32_t* mask, acx_stream_t stream) {
    if (n < 0 || m < 1 || (2 * m) % 4 != 0 || tcap < 2 || (tcap & (tcap - 1))) return acx_fail(ACX_ERR_ARG, "acx_bfs_probe_packed: bad argument");
    if (n == 0) return ACX_OK;
    if (!recs || !table || !nodes || !slot || !mins || !mask) return acx_fail(ACX_ERR_ARG, "acx_bfs_probe_packed: null pointer");
    PackedParams p = {};
    p.recs = recs; p.n = n; p.rowb = 2 * m; p.rec = acx_bfs_record_bytes(m); p.table = table; p.tcap = tcap; p.nodes = nodes;
    p.slot = slot; p.mins = reinterpret_cast<unsigned long long*>(mins); p.mask = mask;
    const unsigned grid = (unsigned)((n + 255) / 256);
    k_bfs_probe_packed<<<grid, 256, 0, (cudaStream_t)stream>>>(p);
    k_bfs_resolve_packed<<<grid, 256, 0, (cudaStream_t)stream>>>(p);
    ACX_CUDA(cudaGetLastError());
    return ACX_OK;
}

extern "C" int acx_bfs_commit_packed(const uint8_t* recs, int64_t n, int m, const int64_t* slot, const int64_t* excl_par,
                                     const int32_t* gmask, int64_t parent_base, int64_t gid_base, uint64_t* n_local,
                                     uint64_t* table, int8_t* nodes, int64_t* gid, int64_t* parent, uint8_t* action,
                                     uint16_t* total, acx_stream_t stream) {
    if (n < 0 || m < 1 || (2 * m) % 4 != 0) return acx_fail(ACX_ERR_ARG, "acx_bfs_commit_packed: bad argument");
    if (n == 0) return ACX_OK;
    if (!recs || !slot || !excl_par || !gmask || !n_local || !table || !nodes || !gid || !parent || !action || !total) return acx_fail(ACX_ERR_ARG, "acx_bfs_commit_packed: null pointer");
    PackedParams p = {};
    p.recs = recs; p.n = n; p.rowb = 2 * m; p.rec = acx_bfs_record_bytes(m); p.table = table; p.slot = const_cast<int64_t*>(slot);
    p.excl_par = excl_par; p.gmask = gmask; p.parent_base = parent_base; p.gid_base = gid_base;
    p.n_local = reinterpret_cast<unsigned long long*>(n_local); p.nodes_w = nodes; p.gid = gid; p.parent = parent; p.action = action;
    p.total = total;
    k_bfs_commit_packed<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(p);
    ACX_CUDA(cudaGetLastError());
    return ACX_OK;
}
