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
