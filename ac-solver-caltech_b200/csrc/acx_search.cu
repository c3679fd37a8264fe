// acx_search.cu -- greedy_search / bfs as a batch of independent search instances on one GPU.
//
// Reference: ac_solver/search/greedy.py:15-121 and ac_solver/search/breadth_first.py:15-97.
// Both are inherently sequential per instance (one pop at a time, order-dependent visited set), so
// the parallelism is ACROSS instances (BASELINE config #1: the 1190 Miller-Schupp presentations)
// and across the 12 moves of the popped node.  One WARP owns one instance and runs the whole
// search inside a single kernel launch:
//     pop (heap minimum / FIFO head)  ->  lanes 0..11 apply moves 0..11 with the bit-plane row code
//     -> success test `total length == 2` in move order, before the visited test (greedy.py:91-100)
//     -> visited test + insert (open addressing, 64-bit slots: fingerprint | node id, full-row
//        compare on a fingerprint hit, never hash-only)  ->  push in move order
//     -> budget test on the number of seen nodes once per popped node (greedy.py:115-119).
// Greedy priority = Python tuple order (total_len, depth, state compared element-wise as signed
// ints, greedy.py:104-113); it is packed as a 64-bit key (total | depth | 13 leading letters in an
// order-preserving 3-bit code) with a full (depth, row) compare on a key tie.
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>

#include "../../include/acx.h"
#include "acx_common.h"
#include "acx_rows.cuh"

namespace acx {

struct SearchParams {
    const int8_t* roots;  // [S, 2m]
    int S, m, cyclical, greedy;
    int64_t budget;  // max_nodes_to_explore
    int64_t cap;     // node capacity per instance (>= budget + 12)
    int64_t tcap;    // hash slots per instance (power of two)
    int pitch;       // bytes per stored row (2m rounded up to 4)
    int path_cap;
    // per-instance workspaces, instance s at offset s * (per-instance size)
    uint8_t* nodes;      // [S][cap][pitch]
    int32_t* parent;     // [S][cap]
    int32_t* depth;      // [S][cap]
    uint16_t* total;     // [S][cap]
    uint8_t* action;     // [S][cap]
    uint64_t* key;       // [S][cap]   greedy priority key of node id
    int32_t* heap;       // [S][cap]   greedy: binary heap of node ids
    uint64_t* table;     // [S][tcap]  0 = empty, else fingerprint << 32 | (id + 1)
    // outputs
    int32_t* result;     // [S][8]: solved, n_seen, n_expanded, path_len, error, last_node, last_action, last_len
    int32_t* path;       // [S][path_cap][2]  (action, total length) incl. the (-1, len0) head
};

enum { RES_SOLVED = 0, RES_SEEN, RES_EXPANDED, RES_PATHLEN, RES_ERROR, RES_LASTNODE, RES_LASTACT, RES_LASTLEN };

__device__ __forceinline__ uint64_t row_hash64(const uint32_t* w, int nwords) {
    uint64_t h = 0x9E3779B97F4A7C15ull;
    for (int k = 0; k < nwords; ++k) {
        h ^= w[k];
        h *= 0xFF51AFD7ED558CCDull;
        h ^= h >> 29;
    }
    return h;
}

__device__ __forceinline__ bool rows_equal(const uint32_t* a, const uint32_t* b, int nwords) {
    uint32_t d = 0;
    for (int k = 0; k < nwords; ++k) d |= a[k] ^ b[k];
    return d == 0u;
}

// Python tuple order on the rows: element-wise signed compare; true when x < y
__device__ __forceinline__ bool row_less(const uint32_t* x, const uint32_t* y, int nwords) {
    uint32_t dx = 0, dy = 0;
    for (int k = nwords - 1; k >= 0; --k) {
        const uint32_t a = x[k], b = y[k];
        if (a != b) {
            dx = a;
            dy = b;
        }
    }
    if (dx == dy) return false;
    const int byte = (__ffs(dx ^ dy) - 1) >> 3;
    return (int8_t)(dx >> (8 * byte)) < (int8_t)(dy >> (8 * byte));
}

// total (9 bits) | depth (16 bits, saturating) | 13 leading letters, 3 bits each, order preserving
__device__ __forceinline__ uint64_t priority_key(const uint8_t* row, int rowb, int total, int depth) {
    uint64_t k = ((uint64_t)min(total, 511) << 16) | (uint64_t)min(depth, 65535);
    for (int j = 0; j < 13; ++j) {
        int v = (j < rowb) ? (int)(int8_t)row[j] : 0;
        v = max(-3, min(3, v)) + 3;
        k = (k << 3) | (uint64_t)v;
    }
    return k;
}

struct Inst {
    uint8_t* nodes;
    int32_t* parent;
    int32_t* depth;
    uint16_t* total;
    uint8_t* action;
    uint64_t* key;
    int32_t* heap;
    uint64_t* table;
    int pitch, nwords;
    uint64_t tmask;
};

// Python's tuple order (total, depth, state as signed ints) of greedy.py:104-113, in stages: the (total | depth) field,
// the true depth when the 16-bit field saturated, then the 13 leading letters -- decided by the 3-bit codes only when no
// letter in front of the first differing code was clamped (|letter| >= 3 all map to code 0 / 6) -- else the full rows.
__device__ __forceinline__ bool node_less(const Inst& I, int32_t a, int32_t b) {
    const uint64_t ka = I.key[a], kb = I.key[b];
    const uint64_t ha = ka >> 39, hb = kb >> 39;
    if (ha != hb) return ha < hb;
    if ((ha & 0xFFFFull) == 0xFFFFull) {
        const int da = I.depth[a], db = I.depth[b];
        if (da != db) return da < db;
    }
    const uint64_t low = (1ull << 39) - 1, la = ka & low, lb = kb & low;
    if (la != lb) {
        const int g = (63 - __clzll((long long)(la ^ lb))) / 3;                    // group of the first differing code
        const uint64_t above = low & ~((1ull << (3 * (g + 1))) - 1);               // the codes in front of it (equal in both)
        const uint64_t clamped = ~la & ~((la >> 1) ^ (la >> 2)) & 0x1249249249ull;  // codes 0 and 6
        if ((clamped & above) == 0ull) return la < lb;
    }
    return row_less(reinterpret_cast<const uint32_t*>(I.nodes + (size_t)a * I.pitch),
                    reinterpret_cast<const uint32_t*>(I.nodes + (size_t)b * I.pitch), I.nwords);
}

__device__ void heap_push(const Inst& I, int64_t& hn, int32_t id) {
    int64_t k = hn++;
    while (k > 0) {
        const int64_t p = (k - 1) >> 1;
        const int32_t pid = I.heap[p];
        if (!node_less(I, id, pid)) break;
        I.heap[k] = pid;
        k = p;
    }
    I.heap[k] = id;
}

__device__ int32_t heap_pop(const Inst& I, int64_t& hn) {
    const int32_t top = I.heap[0];
    const int64_t n = --hn;
    if (n == 0) return top;
    const int32_t last = I.heap[n];
    int64_t k = 0;
    for (;;) {
        const int64_t l = 2 * k + 1, r = l + 1;
        if (l >= n) break;
        int64_t c = l;
        int32_t cid = I.heap[l];
        if (r < n) {
            const int32_t rid = I.heap[r];
            if (node_less(I, rid, cid)) {
                c = r;
                cid = rid;
            }
        }
        if (!node_less(I, cid, last)) break;
        I.heap[k] = cid;
        k = c;
    }
    I.heap[k] = last;
    return top;
}

// find `row` in the table; returns true when present, else leaves `slot` at the empty slot
__device__ __forceinline__ bool table_find(const Inst& I, const uint32_t* row, uint64_t h, uint64_t& slot) {
    const uint32_t fp = (uint32_t)(h >> 32) | 1u;
    slot = h & I.tmask;
    for (;;) {
        const uint64_t e = I.table[slot];
        if (e == 0ull) return false;
        if ((uint32_t)(e >> 32) == fp) {
            const int32_t id = (int32_t)(uint32_t)e - 1;
            if (rows_equal(row, reinterpret_cast<const uint32_t*>(I.nodes + (size_t)id * I.pitch), I.nwords)) return true;
        }
        slot = (slot + 1) & I.tmask;
    }
}

// One warp = one search instance.  W64: m <= 64 (64-bit planes) else generic planes (m <= 128).
template <bool W64>
__global__ void __launch_bounds__(32) k_search(const SearchParams p) {
    extern __shared__ __align__(16) uint8_t smem[];
    const int s = blockIdx.x, lane = threadIdx.x;
    const int m = p.m, rowb = 2 * m, pitch = p.pitch, nwords = pitch / 4;
    Inst I;
    I.nodes = p.nodes + (size_t)s * p.cap * pitch;
    I.parent = p.parent + (size_t)s * p.cap;
    I.depth = p.depth + (size_t)s * p.cap;
    I.total = p.total + (size_t)s * p.cap;
    I.action = p.action + (size_t)s * p.cap;
    I.key = p.key + (size_t)s * p.cap;
    I.heap = p.heap + (size_t)s * p.cap;
    I.table = p.table + (size_t)s * p.tcap;
    I.pitch = pitch;
    I.nwords = nwords;
    I.tmask = (uint64_t)p.tcap - 1;
    int32_t* res = p.result + (size_t)s * 8;

    // shared: 12 child rows, then per-child metadata
    uint8_t* crow = smem + (size_t)lane * pitch;  // valid for lane < 12
    uint64_t* c_hash = reinterpret_cast<uint64_t*>(smem + (size_t)12 * pitch);
    int32_t* c_tot = reinterpret_cast<int32_t*>(c_hash + 12);
    int32_t* c_st = c_tot + 12;
    int32_t* c_ex = c_st + 12;

    int64_t n_nodes = 0, hn = 0, head = 0, expanded = 0;
    int32_t cur = 0, last_len = 0, fin_action = 11;
    int solved = 0, err = 0;

    if (lane == 0) {  // root (greedy.py:36-68)
        const int8_t* root = p.roots + (size_t)s * rowb;
        int len0 = 0;
        for (int k = 0; k < pitch; ++k) {
            const int8_t v = (k < rowb) ? root[k] : (int8_t)0;
            I.nodes[k] = (uint8_t)v;
            len0 += v != 0;
        }
        I.parent[0] = -1;
        I.depth[0] = 0;
        I.total[0] = (uint16_t)len0;
        I.action[0] = 0xff;
        const uint32_t* rw = reinterpret_cast<const uint32_t*>(I.nodes);
        const uint64_t h = row_hash64(rw, nwords);
        I.table[h & I.tmask] = ((uint64_t)((uint32_t)(h >> 32) | 1u) << 32) | 1ull;
        if (p.greedy) {
            I.key[0] = priority_key(I.nodes, rowb, len0, 0);
            I.heap[0] = 0;
            hn = 1;
        }
        n_nodes = 1;
        last_len = len0;
    }
    __syncwarp();

    for (;;) {
        int go = 0;
        if (lane == 0) {
            if (p.greedy ? (hn > 0) : (head < n_nodes)) {
                cur = p.greedy ? heap_pop(I, hn) : (int32_t)head++;
                go = 1;
            }
        }
        go = __shfl_sync(0xffffffffu, go, 0);
        if (!go) break;
        cur = __shfl_sync(0xffffffffu, cur, 0);
        __syncwarp();
        if (lane < 12) {  // the 12 successors (greedy.py:76-83)
            const uint32_t* src = reinterpret_cast<const uint32_t*>(I.nodes + (size_t)cur * pitch);
            uint32_t* dst = reinterpret_cast<uint32_t*>(crow);
            for (int k = 0; k < nwords; ++k) dst[k] = src[k];
            int l0, l1;
            uint32_t st;
            if (W64) acmove_row64<0>(crow, m, lane, p.cyclical != 0, l0, l1, st);
            else acmove_row<4, 0>(crow, crow, m, lane, p.cyclical != 0, l0, l1, st);
            const uint64_t h = row_hash64(dst, nwords);
            uint64_t slot;
            c_hash[lane] = h;
            c_tot[lane] = l0 + l1;
            c_st[lane] = (int32_t)st;
            c_ex[lane] = table_find(I, dst, h, slot) ? 1 : 0;
        }
        __syncwarp();
        int stop = 0;
        if (lane == 0) {
            const int depth_cur = I.depth[cur];
            for (int a = 0; a < 12; ++a) {
                if (c_st[a] != 0) {  // the reference raises AssertionError here (utils.py:264-266)
                    err = 1;
                    stop = 1;
                    break;
                }
                last_len = c_tot[a];
                if (last_len == 2) {  // success test before the visited test (greedy.py:91-100)
                    solved = 1;
                    fin_action = a;
                    stop = 1;
                    break;
                }
                if (c_ex[a]) continue;
                const uint32_t* row = reinterpret_cast<const uint32_t*>(smem + (size_t)a * pitch);
                const uint64_t h = c_hash[a];
                uint64_t slot;
                if (table_find(I, row, h, slot)) continue;  // equal to an earlier sibling
                const int32_t id = (int32_t)n_nodes++;
                uint32_t* dst = reinterpret_cast<uint32_t*>(I.nodes + (size_t)id * pitch);
                for (int k = 0; k < nwords; ++k) dst[k] = row[k];
                I.parent[id] = cur;
                I.depth[id] = depth_cur + 1;
                I.total[id] = (uint16_t)last_len;
                I.action[id] = (uint8_t)a;
                I.table[slot] = ((uint64_t)((uint32_t)(h >> 32) | 1u) << 32) | (uint64_t)(uint32_t)(id + 1);
                if (p.greedy) {
                    I.key[id] = priority_key(reinterpret_cast<const uint8_t*>(row), rowb, last_len, depth_cur + 1);
                    heap_push(I, hn, id);
                }
            }
            ++expanded;
            if (!stop && n_nodes >= p.budget) stop = 1;  // greedy.py:115-119
        }
        stop = __shfl_sync(0xffffffffu, stop, 0);
        if (stop) break;
    }

    if (lane == 0) {
        int path_len = 0;
        if (!err && (solved || p.greedy)) {  // bfs failure returns None (breadth_first.py:97)
            const int d = I.depth[cur];
            path_len = d + 2;
            if (path_len <= p.path_cap) {
                int32_t* path = p.path + (size_t)s * p.path_cap * 2;
                path[2 * (d + 1)] = solved ? fin_action : 11;  // greedy.py:121 hands back the last expansion
                path[2 * (d + 1) + 1] = solved ? 2 : last_len;
                int32_t id = cur;
                for (int k = d; k >= 0; --k) {
                    path[2 * k] = (k == 0) ? -1 : (int32_t)I.action[id];
                    path[2 * k + 1] = (int32_t)I.total[id];
                    id = I.parent[id];
                }
            } else {
                path_len = -path_len;
            }
        }
        res[RES_SOLVED] = solved;
        res[RES_SEEN] = (int32_t)n_nodes;
        res[RES_EXPANDED] = (int32_t)expanded;
        res[RES_PATHLEN] = path_len;
        res[RES_ERROR] = err;
        res[RES_LASTNODE] = cur;
        res[RES_LASTACT] = fin_action;
        res[RES_LASTLEN] = last_len;
    }
}

static size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

struct SearchLayout {
    int64_t cap, tcap;
    int pitch;
    size_t off_nodes, off_parent, off_depth, off_total, off_action, off_key, off_heap, off_table, bytes;
};

static SearchLayout search_layout(int S, int m, int64_t budget) {
    SearchLayout L;
    if (budget < 1) budget = 1;
    L.cap = budget + 16;
    L.tcap = 64;
    while (L.tcap < 2 * L.cap) L.tcap <<= 1;
    L.pitch = (2 * m + 3) & ~3;
    size_t o = 0;
    L.off_nodes = o;  o = align_up(o + (size_t)S * L.cap * L.pitch, 256);
    L.off_parent = o; o = align_up(o + (size_t)S * L.cap * 4, 256);
    L.off_depth = o;  o = align_up(o + (size_t)S * L.cap * 4, 256);
    L.off_total = o;  o = align_up(o + (size_t)S * L.cap * 2, 256);
    L.off_action = o; o = align_up(o + (size_t)S * L.cap, 256);
    L.off_key = o;    o = align_up(o + (size_t)S * L.cap * 8, 256);
    L.off_heap = o;   o = align_up(o + (size_t)S * L.cap * 4, 256);
    L.off_table = o;  o = align_up(o + (size_t)S * L.tcap * 8, 256);
    L.bytes = o;
    return L;
}

}  // namespace acx

using namespace acx;

extern "C" int64_t acx_search_workspace_bytes(int n_instances, int m, int64_t budget) {
    if (n_instances < 0 || m < 1 || m > ACX_MAX_RELATOR_LENGTH) return acx_fail(ACX_ERR_ARG, "acx_search_workspace_bytes: bad argument");
    return (int64_t)search_layout(n_instances, m, budget).bytes;
}

extern "C" int acx_search_batch(const int8_t* roots, int n_instances, int m, int64_t budget, int cyclical, int greedy,
                                void* workspace, int64_t workspace_bytes, int32_t* result, int32_t* path, int path_cap,
                                acx_stream_t stream) {
    if (n_instances < 0 || m < 1 || m > ACX_MAX_RELATOR_LENGTH || path_cap < 2)
        return acx_fail(ACX_ERR_ARG, "acx_search_batch: need n_instances >= 0, 1 <= m <= 128, path_cap >= 2");
    if (n_instances == 0) return ACX_OK;
    if (!roots || !workspace || !result || !path) return acx_fail(ACX_ERR_ARG, "acx_search_batch: null pointer");
    if (budget < 1) budget = 1;
    if (budget > 2000000000ll) return acx_fail(ACX_ERR_ARG, "acx_search_batch: budget above 2e9 nodes per instance");
    const SearchLayout L = search_layout(n_instances, m, budget);
    if ((int64_t)L.bytes > workspace_bytes) return acx_fail(ACX_ERR_ARG, "acx_search_batch: workspace too small (%zu needed)", L.bytes);
    cudaStream_t st = (cudaStream_t)stream;
    uint8_t* ws = reinterpret_cast<uint8_t*>(workspace);
    SearchParams p;
    p.roots = roots;
    p.S = n_instances;
    p.m = m;
    p.cyclical = cyclical;
    p.greedy = greedy;
    p.budget = budget;
    p.cap = L.cap;
    p.tcap = L.tcap;
    p.pitch = L.pitch;
    p.path_cap = path_cap;
    p.nodes = ws + L.off_nodes;
    p.parent = reinterpret_cast<int32_t*>(ws + L.off_parent);
    p.depth = reinterpret_cast<int32_t*>(ws + L.off_depth);
    p.total = reinterpret_cast<uint16_t*>(ws + L.off_total);
    p.action = ws + L.off_action;
    p.key = reinterpret_cast<uint64_t*>(ws + L.off_key);
    p.heap = reinterpret_cast<int32_t*>(ws + L.off_heap);
    p.table = reinterpret_cast<uint64_t*>(ws + L.off_table);
    p.result = result;
    p.path = path;
    ACX_CUDA(cudaMemsetAsync(p.table, 0, (size_t)n_instances * L.tcap * 8, st));
    const size_t smem = (size_t)12 * L.pitch + 12 * (8 + 4 + 4 + 4) + 16;
    if (m <= 64) k_search<true><<<n_instances, 32, smem, st>>>(p);
    else k_search<false><<<n_instances, 32, smem, st>>>(p);
    ACX_CUDA(cudaGetLastError());
    return ACX_OK;
}
