// acx_pbfs.cu -- breadth-first search on bit-plane nodes (max_relator_length <= 36).
//
// The byte rows of the reference (int8 [2m], breadth_first.py:52-60) never reach HBM here: a node is
// FOUR 64-bit words -- the sign / type planes of both relators plus their lengths --
//
//     w0 = S(r0) | len(r0) << 40      w1 = T(r0) | len(r1) << 40      w2 = S(r1)      w3 = T(r1)
//
// i.e. 32 bytes = one DRAM sector instead of 72 bytes, and a candidate record on its way to the owner
// rank is the same 32 bytes with its discovery order (48 bits) and row status (2 bits) folded into the
// unused top bits of w2 / w3.  The 12 moves (ac_moves.py:159-213) run on the planes in registers with
// the same row arithmetic as the env-step kernel (acx_rel64.cuh); nothing is packed or unpacked.
// Valid for presentations over {x, y} whose root is freely reduced (and cyclically reduced when
// cyclical=True) -- every Miller-Schupp / AK(n) instance; other roots use the byte-row kernels of
// acx_bfs.cu.  Visited-set semantics are those of acx_bfs.cu: first discovery in the reference's FIFO
// order wins (breadth_first.py:69-95), resolved by order numbers, not by arrival order.
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>

#include "../../include/acx.h"
#include "acx_common.h"
#include "acx_pbnode.cuh"

#include <cub/device/device_scan.cuh>
#include <cub/iterator/transform_input_iterator.cuh>

namespace acx {

// byte rows -> plane nodes (the root; tests).  ok = 0 when the row is outside the plane domain.
__global__ void k_pbfs_pack(const int8_t* rows, int64_t n, int m, int cyclical, uint64_t* out, uint8_t* ok) {
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const uint8_t* row = reinterpret_cast<const uint8_t*>(rows + j * 2 * m);
    Raw64 raw[2];
    uint32_t chk = 0u;
    pack_rel_dynamic64(row, m, raw[0], chk);
    pack_rel_dynamic64(row + m, m, raw[1], chk);
    R64 r[2];
    int cnt[2];
    const int cls = classify_row64(raw, chk, r, cnt);
    bool good = cls == ROW_FAST;
    if (good && cyclical) good = !r64_needs_cyclic(r[0]) && !r64_needs_cyclic(r[1]);
    PWords w = {};
    if (good) w = pb_pack(r[0], r[1]);
    pb_store(out + 4 * j, w);
    ok[j] = good ? 1 : 0;
}

// plane nodes -> byte rows (path replay, tests)
__global__ void k_pbfs_unpack(const uint64_t* nodes, int64_t n, int m, int8_t* rows) {
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const PWords v = pb_load(nodes + 4 * j);
    const int len[2] = {(int)((v.w[0] >> 40) & 63u), (int)((v.w[1] >> 40) & 63u)};
    for (int i = 0; i < 2; ++i) {
        const uint64_t S = v.w[2 * i] & PB_PLANE, T = v.w[2 * i + 1] & PB_PLANE;
        for (int k = 0; k < m; ++k) {
            int letter = 0;
            if (k < len[i]) {
                letter = ((T >> k) & 1ull) ? 2 : 1;
                if ((S >> k) & 1ull) letter = -letter;
            }
            rows[j * 2 * m + i * m + k] = (int8_t)letter;
        }
    }
}

__global__ void k_pbfs_hash(const uint64_t* nodes, int64_t n, uint64_t* out) {
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j < n) out[j] = pb_hash(pb_state(pb_load(nodes + 4 * j)));
}

__global__ void k_pbfs_seed(const uint64_t* node, uint64_t* table, int64_t tcap, int64_t node_id) {
    const PWords s = pb_state(pb_load(node));
    const uint64_t h = pb_hash(s);
    const uint64_t tag = ((h >> 48) | 1ull) << 48;
    const uint64_t mask = (uint64_t)tcap - 1;
    uint64_t slot = (h ^ (h >> 23)) & mask;
    while (table[slot] != 0ull) slot = (slot + 1) & mask;
    table[slot] = tag | (uint64_t)node_id;
}

struct PbDst {
    uint8_t* region[PB_MAX_WORLD];  // this sender's region in the receive buffer of each owner (local or peer memory)
};

// Frontier step: one thread per parent.  Load the 32-byte node, apply the 12 moves to the planes, give every
// child that differs from its parent an order number, an owner rank and a slot in that owner's region, and
// store the 32-byte record there (one full sector per lane; a peer region is written over NVLink).
// Slots are reserved per warp and per group of GM moves: one ballot scan per owner, one atomicAdd per
// owner per warp-step.  cursor[o] = records written to region o; cursor[world] = overflow flag.
template <int GM>
__global__ void __launch_bounds__(256) k_pbfs_expand(const uint64_t* nodes, const int64_t* loc, const int64_t* gids,
                                                     int64_t ord_base, int64_t f, int m, int cyclical, int world,
                                                     int64_t region_cap, unsigned long long* cursor, const PbDst dst,
                                                     int64_t ntiles) {
    const int lane = threadIdx.x & 31;
    const unsigned ltmask = (1u << lane) - 1u;
    const bool cyc = cyclical != 0;
    const int64_t warp0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t tile = warp0; tile < ntiles; tile += nwarps) {
        const int64_t k = tile * 32 + lane;
        const bool active = k < f;
        R64 r[2];
        r[0].S = r[0].T = r[1].S = r[1].T = 0ull;
        r[0].n = r[1].n = 1;
        int64_t ord0 = 0;
        if (active) {
            const PWords v = pb_load(nodes + 4 * loc[k]);
            r[0].S = v.w[0] & PB_PLANE;
            r[0].T = v.w[1] & PB_PLANE;
            r[0].n = (int)((v.w[0] >> 40) & 63u);
            r[1].S = v.w[2] & PB_PLANE;
            r[1].T = v.w[3] & PB_PLANE;
            r[1].n = (int)((v.w[1] >> 40) & 63u);
            ord0 = (gids[k] - ord_base) * 12;
        }
#pragma unroll 1
        for (int g = 0; g < 12 / GM; ++g) {
            PWords cw[GM];
            int own[GM];
            unsigned rank[GM];
#pragma unroll
            for (int j = 0; j < GM; ++j) {
                const int a = g * GM + j;
                own[j] = -1;
                rank[j] = 0;
                if (!active) continue;
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
                    // the move reproduces the parent (length guard, x^k conjugated by x, a conjugation undone by the
                    // cyclic reduction): already visited, nothing to send
                    const R64& old = i ? r[1] : r[0];
                    if (!changed || (x.n == old.n && x.S == old.S && x.T == old.T)) continue;
                }
                cw[j] = i ? pb_pack(r[0], x) : pb_pack(x, r[1]);
                own[j] = pb_owner(pb_hash(cw[j]), world);
                const uint64_t ord = (uint64_t)(ord0 + a);
                cw[j].w[2] |= (ord & 0xFFFFFFFull) << 36;
                cw[j].w[3] |= (((ord >> 28) & 0xFFFFFull) << 36) | (st << 56);
            }
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
                if (own[j] >= 0) {
                    const unsigned long long pos = b + rank[j];
                    if (pos < (unsigned long long)region_cap) pb_store(reinterpret_cast<uint64_t*>(dst.region[own[j]] + pos * 32ull), cw[j]);
                    else cursor[world] = 1ull;  // a region overflowed: the host raises
                }
            }
        }
    }
}

struct PbParams {
    const uint64_t* recs;  // nreg regions of rcap records (4 words each); region s holds pre[s+1] - pre[s]
    int64_t n;
    int nreg;
    int64_t rcap;
    int64_t pre[PB_MAX_WORLD + 1];
    uint64_t* table;
    int64_t tcap;
    const uint64_t* nodes;
    int64_t* slot;
    unsigned long long* mins;
    int* mask;
    const int64_t* excl_par;
    const int* gmask;
    int64_t parent_base, gid_base;
    unsigned long long* n_local;
    uint64_t* nodes_w;
    int64_t* gid;
    int64_t* parent;
    uint8_t* action;
};

__device__ __forceinline__ int64_t pb_index(const PbParams& p, int64_t t) {
    int s = 0;
    while (s + 1 < p.nreg && t >= p.pre[s + 1]) ++s;
    return (int64_t)s * p.rcap + (t - p.pre[s]);
}

// Table slot = tag(16) | PROV(1) | value(47): value = node id, or (PROV) the index of the candidate that
// currently claims the state in this round.  A claim is replaced by an equal candidate of lower order.
__global__ void __launch_bounds__(256) k_pbfs_probe(const PbParams p) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= p.n) return;
    const int64_t j = pb_index(p, t);
    const PWords rec = pb_load(p.recs + 4 * j);
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
    const PWords s = pb_state(rec);
    const uint64_t h = pb_hash(s);
    const uint64_t tag = ((h >> 48) | 1ull) << 48;
    const uint64_t mine = tag | PB_PROV | (uint64_t)j;
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
            if (e & PB_PROV) {
                for (;;) {
                    const int64_t other = (int64_t)(e & PB_VAL);
                    if (other == j) {
                        p.slot[j] = (int64_t)slot;
                        return;
                    }
                    const PWords orec = pb_load(p.recs + 4 * other);
                    if (!pb_equal(s, pb_state(orec))) break;
                    if (pb_order(orec) < ord) {
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
                    if (!(e & PB_PROV)) break;
                }
            }
            if (!(e & PB_PROV) && (e >> 48) == (tag >> 48)) {
                const int64_t id = (int64_t)(e & PB_VAL);
                if (pb_equal(s, pb_load(p.nodes + 4 * id))) {
                    p.slot[j] = -1;
                    return;
                }
            }
        }
        slot = (slot + 1) & mask;
    }
}

__global__ void __launch_bounds__(256) k_pbfs_resolve(const PbParams p) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= p.n) return;
    const int64_t j = pb_index(p, t);
    const int64_t s = p.slot[j];
    if (s < 0) return;
    const uint64_t e = p.table[s];
    if ((e & PB_PROV) && (int64_t)(e & PB_VAL) == j) {
        const int64_t ord = pb_order(pb_load(p.recs + 4 * j));
        atomicOr(p.mask + ord / 12, 1 << (int)(ord % 12));
    } else {
        p.slot[j] = -1;  // a duplicate inside the round
    }
}

__global__ void __launch_bounds__(256) k_pbfs_commit(const PbParams p) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= p.n) return;
    const int64_t j = pb_index(p, t);
    const int64_t s = p.slot[j];
    if (s < 0) return;
    const PWords rec = pb_load(p.recs + 4 * j);
    const int64_t ord = pb_order(rec);
    const int64_t par = ord / 12;
    const int act = (int)(ord % 12);
    const int64_t gid = p.gid_base + p.excl_par[par] + __popc((unsigned)p.gmask[par] & ((1u << act) - 1u));
    const int64_t id = (int64_t)atomicAdd(p.n_local, 1ull);
    const PWords st = pb_state(rec);
    pb_store(p.nodes_w + 4 * id, st);
    p.gid[id] = gid;
    p.parent[id] = p.parent_base + par;
    p.action[id] = (uint8_t)act;
    // the slot keeps its tag (recomputed from the state: no read of the random table line) and now names the node
    p.table[s] = (((pb_hash(st) >> 48) | 1ull) << 48) | (uint64_t)id;
}

static int pb_regions(PbParams& p, const int64_t* region_counts, int n_regions, int64_t region_cap, const char* who) {
    if (!region_counts || n_regions < 1 || n_regions > PB_MAX_WORLD || region_cap < 0)
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

}  // namespace acx

using namespace acx;

extern "C" int acx_pbfs_pack(const int8_t* rows, int64_t n, int m, int cyclical, uint64_t* nodes_out, uint8_t* ok_out,
                             acx_stream_t stream) {
    if (n < 0 || m < 1 || m > PB_MAX_M) return acx_fail(ACX_ERR_ARG, "acx_pbfs_pack: need n >= 0 and 1 <= m <= 36");
    if (n == 0) return ACX_OK;
    if (!rows || !nodes_out || !ok_out) return acx_fail(ACX_ERR_ARG, "acx_pbfs_pack: null pointer");
    k_pbfs_pack<<<(unsigned)((n + 127) / 128), 128, 0, (cudaStream_t)stream>>>(rows, n, m, cyclical, nodes_out, ok_out);
    ACX_CUDA(cudaGetLastError());
    return ACX_OK;
}

extern "C" int acx_pbfs_unpack(const uint64_t* nodes, int64_t n, int m, int8_t* rows_out, acx_stream_t stream) {
    if (n < 0 || m < 1 || m > PB_MAX_M) return acx_fail(ACX_ERR_ARG, "acx_pbfs_unpack: need n >= 0 and 1 <= m <= 36");
    if (n == 0) return ACX_OK;
    if (!nodes || !rows_out) return acx_fail(ACX_ERR_ARG, "acx_pbfs_unpack: null pointer");
    k_pbfs_unpack<<<(unsigned)((n + 127) / 128), 128, 0, (cudaStream_t)stream>>>(nodes, n, m, rows_out);
    ACX_CUDA(cudaGetLastError());
    return ACX_OK;
}

extern "C" int acx_pbfs_hash(const uint64_t* nodes, int64_t n, uint64_t* hash_out, acx_stream_t stream) {
    if (n < 0) return acx_fail(ACX_ERR_ARG, "acx_pbfs_hash: bad argument");
    if (n == 0) return ACX_OK;
    if (!nodes || !hash_out) return acx_fail(ACX_ERR_ARG, "acx_pbfs_hash: null pointer");
    k_pbfs_hash<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(nodes, n, hash_out);
    ACX_CUDA(cudaGetLastError());
    return ACX_OK;
}

extern "C" int acx_pbfs_seed(const uint64_t* node, uint64_t* table, int64_t tcap, int64_t node_id, acx_stream_t stream) {
    if (!node || !table || tcap < 2 || (tcap & (tcap - 1)) || node_id < 0) return acx_fail(ACX_ERR_ARG, "acx_pbfs_seed: bad argument");
    k_pbfs_seed<<<1, 1, 0, (cudaStream_t)stream>>>(node, table, tcap, node_id);
    ACX_CUDA(cudaGetLastError());
    return ACX_OK;
}

extern "C" int acx_pbfs_expand(const uint64_t* nodes, const int64_t* loc, const int64_t* gids, int64_t ord_base, int64_t f, int m,
                               int cyclical, int world, int64_t region_cap, uint64_t* cursor, uint8_t* const* dst_regions,
                               int ctas_per_sm_limit, acx_stream_t stream) {
    if (f < 0 || m < 1 || m > PB_MAX_M || world < 1 || world > PB_MAX_WORLD || region_cap < 1)
        return acx_fail(ACX_ERR_ARG, "acx_pbfs_expand: bad argument (1 <= m <= 36, 1 <= world <= 32)");
    if (f == 0) return ACX_OK;
    if (f > (1ll << 44)) return acx_fail(ACX_ERR_ARG, "acx_pbfs_expand: order numbers are 48 bits");
    if (!nodes || !loc || !gids || !cursor || !dst_regions) return acx_fail(ACX_ERR_ARG, "acx_pbfs_expand: null pointer");
    PbDst dst = {};
    for (int o = 0; o < world; ++o) {
        if (!dst_regions[o] || ((uintptr_t)dst_regions[o] & 31u)) return acx_fail(ACX_ERR_ARG, "acx_pbfs_expand: destination regions must be non-null and 32-byte aligned");
        dst.region[o] = dst_regions[o];
    }
    if ((uintptr_t)nodes & 31u) return acx_fail(ACX_ERR_ARG, "acx_pbfs_expand: nodes must be 32-byte aligned");
    static int ctas_per_sm[ACX_MAX_DEVICES] = {0};
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= ACX_MAX_DEVICES) dev = 0;
    auto kern = k_pbfs_expand<4>;
    if (ctas_per_sm[dev] == 0) {
        int occ = 0;
        ACX_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, 256, 0));
        ctas_per_sm[dev] = occ > 0 ? occ : 1;
    }
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int64_t ntiles = (f + 31) / 32;
    int64_t grid = (ntiles + 7) / 8;
    int resident = ctas_per_sm[dev];
    if (ctas_per_sm_limit > 0 && ctas_per_sm_limit < resident) resident = ctas_per_sm_limit;  // leave room for a concurrent kernel
    if (grid > (int64_t)sms * resident) grid = (int64_t)sms * resident;
    kern<<<(unsigned)grid, 256, 0, (cudaStream_t)stream>>>(nodes, loc, gids, ord_base, f, m, cyclical, world, region_cap,
                                                          reinterpret_cast<unsigned long long*>(cursor), dst, ntiles);
    ACX_CUDA(cudaGetLastError());
    return ACX_OK;
}

extern "C" int acx_pbfs_probe(const uint64_t* recs, const int64_t* region_counts, int n_regions, int64_t region_cap,
                              uint64_t* table, int64_t tcap, const uint64_t* nodes, int64_t* slot, uint64_t* mins, int32_t* mask,
                              acx_stream_t stream) {
    if (tcap < 2 || (tcap & (tcap - 1))) return acx_fail(ACX_ERR_ARG, "acx_pbfs_probe: table capacity must be a power of two");
    PbParams p = {};
    if (int rc = pb_regions(p, region_counts, n_regions, region_cap, "acx_pbfs_probe")) return rc;
    if (p.n == 0) return ACX_OK;
    if (!recs || !table || !nodes || !slot || !mins || !mask) return acx_fail(ACX_ERR_ARG, "acx_pbfs_probe: null pointer");
    p.recs = recs; p.table = table; p.tcap = tcap; p.nodes = nodes; p.slot = slot;
    p.mins = reinterpret_cast<unsigned long long*>(mins); p.mask = mask;
    const unsigned grid = (unsigned)((p.n + 255) / 256);
    k_pbfs_probe<<<grid, 256, 0, (cudaStream_t)stream>>>(p);
    k_pbfs_resolve<<<grid, 256, 0, (cudaStream_t)stream>>>(p);
    ACX_CUDA(cudaGetLastError());
    return ACX_OK;
}

extern "C" int acx_pbfs_commit(const uint64_t* recs, const int64_t* region_counts, int n_regions, int64_t region_cap,
                               const int64_t* slot, const int64_t* excl_par, const int32_t* gmask, int64_t parent_base,
                               int64_t gid_base, uint64_t* n_local, uint64_t* table, uint64_t* nodes, int64_t* gid,
                               int64_t* parent, uint8_t* action, acx_stream_t stream) {
    PbParams p = {};
    if (int rc = pb_regions(p, region_counts, n_regions, region_cap, "acx_pbfs_commit")) return rc;
    if (p.n == 0) return ACX_OK;
    if (!recs || !slot || !excl_par || !gmask || !n_local || !table || !nodes || !gid || !parent || !action) return acx_fail(ACX_ERR_ARG, "acx_pbfs_commit: null pointer");
    p.recs = recs; p.table = table; p.slot = const_cast<int64_t*>(slot); p.excl_par = excl_par; p.gmask = gmask;
    p.parent_base = parent_base; p.gid_base = gid_base; p.n_local = reinterpret_cast<unsigned long long*>(n_local);
    p.nodes_w = nodes; p.gid = gid; p.parent = parent; p.action = action;
    k_pbfs_commit<<<(unsigned)((p.n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(p);
    ACX_CUDA(cudaGetLastError());
    return ACX_OK;
}

namespace acx {
struct PopcountMask {
    __host__ __device__ __forceinline__ int64_t operator()(const int32_t& m) const {
#ifdef __CUDA_ARCH__
        return (int64_t)__popc((unsigned)m);
#else
        return (int64_t)__builtin_popcount((unsigned)m);
#endif
    }
};
__global__ void k_mask_total(const int32_t* mask, const int64_t* excl, int64_t n, int64_t* total) {
    total[0] = excl[n - 1] + __popc((unsigned)mask[n - 1]);
}
}  // namespace acx

// exclusive scan of popcount(mask[p]) over the parents of a round (one pass, cub): excl_out[p] = new nodes
// discovered by parents before p, total_out[0] = new nodes of the round.  temp == null: return the scratch size.
extern "C" int64_t acx_bfs_mask_scan(const int32_t* mask, int64_t n, int64_t* excl_out, int64_t* total_out, void* temp,
                                     int64_t temp_bytes, acx_stream_t stream) {
    if (n < 0 || n > 0x7FFFFFFFll) return acx_fail(ACX_ERR_ARG, "acx_bfs_mask_scan: bad argument");
    cub::TransformInputIterator<int64_t, PopcountMask, const int32_t*> in(mask, PopcountMask());
    size_t need = 0;
    cudaError_t e = cub::DeviceScan::ExclusiveSum(nullptr, need, in, excl_out, (int)n, (cudaStream_t)stream);
    if (e != cudaSuccess) return acx_fail(ACX_ERR_CUDA, "cub::DeviceScan (size query): %s", cudaGetErrorString(e));
    if (!temp) return (int64_t)need;
    if (n == 0) return ACX_OK;
    if (!mask || !excl_out || !total_out) return acx_fail(ACX_ERR_ARG, "acx_bfs_mask_scan: null pointer");
    if ((size_t)temp_bytes < need) return acx_fail(ACX_ERR_ARG, "acx_bfs_mask_scan: scratch too small");
    size_t tb = (size_t)temp_bytes;
    e = cub::DeviceScan::ExclusiveSum(temp, tb, in, excl_out, (int)n, (cudaStream_t)stream);
    if (e != cudaSuccess) return acx_fail(ACX_ERR_CUDA, "cub::DeviceScan: %s", cudaGetErrorString(e));
    k_mask_total<<<1, 1, 0, (cudaStream_t)stream>>>(mask, excl_out, n, total_out);
    ACX_CUDA(cudaGetLastError());
    return ACX_OK;
}

// ---- the same scan for callers that may not allocate (CUDA graph capture, serving loops) ------------------------
// The cub scratch (tile descriptors of the one-pass scan) lives in a file-scope device array: every GPU that loads the
// library holds ACX_MASK_SCAN_STATIC_BYTES until exit.  In host code the NAME of the array is a host-side stand-in,
// not a device address; the address comes from cudaGetSymbolAddress, asked once per device with that device current.
static __device__ __align__(256) uint8_t g_mask_scan_scratch[ACX_MASK_SCAN_STATIC_BYTES];

static int mask_scan_scratch(void** out) {
    static void* addr[ACX_MAX_DEVICES] = {nullptr};
    int dev = -1;
    ACX_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= ACX_MAX_DEVICES) return acx_fail(ACX_ERR_ARG, "acx_bfs_mask_scan_static: device %d out of range", dev);
    if (!addr[dev]) {
        void* ptr = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&ptr, g_mask_scan_scratch);
        if (e != cudaSuccess || !ptr)
            return acx_fail(ACX_ERR_CUDA, "acx_bfs_mask_scan_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
        addr[dev] = ptr;
    }
    *out = addr[dev];
    return ACX_OK;
}

extern "C" int acx_bfs_mask_scan_static(const int32_t* mask, int64_t n, int64_t* excl_out, int64_t* total_out,
                                        acx_stream_t stream) {
    if (n < 0) return acx_fail(ACX_ERR_ARG, "acx_bfs_mask_scan_static: n < 0");
    if (n > ACX_MASK_SCAN_STATIC_MAX_N)
        return acx_fail(ACX_ERR_ARG, "acx_bfs_mask_scan_static: n = %lld is beyond the static scratch (at most %lld parents per call); "
                        "use acx_bfs_mask_scan with caller scratch", (long long)n, (long long)ACX_MASK_SCAN_STATIC_MAX_N);
    if (n == 0) return ACX_OK;
    if (!mask || !excl_out || !total_out) return acx_fail(ACX_ERR_ARG, "acx_bfs_mask_scan_static: null pointer");
    cub::TransformInputIterator<int64_t, PopcountMask, const int32_t*> in(mask, PopcountMask());
    size_t need = 0;
    cudaError_t e = cub::DeviceScan::ExclusiveSum(nullptr, need, in, excl_out, (int)n, (cudaStream_t)stream);
    if (e != cudaSuccess) return acx_fail(ACX_ERR_CUDA, "cub::DeviceScan (size query): %s", cudaGetErrorString(e));
    if (need > (size_t)ACX_MASK_SCAN_STATIC_BYTES)
        return acx_fail(ACX_ERR_ARG, "acx_bfs_mask_scan_static: the scan wants %zu bytes of scratch, the static array has %d", need,
                        (int)ACX_MASK_SCAN_STATIC_BYTES);
    void* temp = nullptr;
    if (int rc = mask_scan_scratch(&temp)) return rc;
    // nothing to clear: the scan's own first kernel re-initialises every tile descriptor it is going to read
    size_t tb = (size_t)ACX_MASK_SCAN_STATIC_BYTES;
    e = cub::DeviceScan::ExclusiveSum(temp, tb, in, excl_out, (int)n, (cudaStream_t)stream);
    if (e != cudaSuccess) return acx_fail(ACX_ERR_CUDA, "cub::DeviceScan: %s", cudaGetErrorString(e));
    k_mask_total<<<1, 1, 0, (cudaStream_t)stream>>>(mask, excl_out, n, total_out);
    ACX_CUDA(cudaGetLastError());
    return ACX_OK;
}

namespace acx {
// Group the local frontier of a level by round: window w = (gid - lo) / chunk.  Inside a window the order is free,
// so a counting sort (histogram, scan on the host side, scatter) replaces a full sort.  Atomics are aggregated
// per warp with match.any: consecutive local nodes were discovered in the same round and mostly share a window.
// Block-level aggregation: a block handles WIN_ITEMS * 256 consecutive local nodes; they were discovered in one or two
// rounds, so their windows span a short range [wmin, wmax] that is counted / ranked in shared memory, with one global
// atomic per (block, window).  A block whose range exceeds the shared table falls back to per-element global atomics.
static constexpr int WIN_ITEMS = 16, WIN_BINS = 1024;

template <bool SCATTER>
__global__ void __launch_bounds__(256) k_win_part(const int64_t* gid, int64_t n, int64_t lo, int64_t chunk, int64_t n_windows, unsigned long long* counts,
                                                  const int64_t* offsets, unsigned long long* cursor, int64_t first,
                                                  int64_t* loc_out, int64_t* gid_out) {
    __shared__ unsigned s_cnt[WIN_BINS];
    __shared__ unsigned long long s_base[WIN_BINS];
    __shared__ int s_wmin, s_wmax;
    const int tid = threadIdx.x, lane = tid & 31;
    const int shift = (chunk & (chunk - 1)) == 0 ? 63 - __clzll(chunk) : -1;
    const int64_t i0 = (int64_t)blockIdx.x * (WIN_ITEMS * 256) + tid;
    if (tid == 0) {
        s_wmin = 0x7FFFFFFF;
        s_wmax = -1;
    }
    __syncthreads();
    int64_t g[WIN_ITEMS];
    int w[WIN_ITEMS];
    int mn = 0x7FFFFFFF, mx = -1;
#pragma unroll
    for (int k = 0; k < WIN_ITEMS; ++k) {
        const int64_t i = i0 + (int64_t)k * 256;
        w[k] = -1;
        g[k] = 0;
        if (i < n) {
            g[k] = gid[i];
            // outside [lo, lo + n_windows * chunk): another block of the level's windows, or gid -1 = a hole of the node
            // store (acx_pbfs_probe3), not a node
            if (g[k] >= lo) {
                // a 64-bit division per node costs more than the pass over the ids: chunks are powers of two in practice
                const int64_t wk = shift >= 0 ? ((g[k] - lo) >> shift) : (g[k] - lo) / chunk;
                if (wk < n_windows) {
                    w[k] = (int)wk;
                    mn = min(mn, w[k]);
                    mx = max(mx, w[k]);
                }
            }
        }
    }
    mn = __reduce_min_sync(0xFFFFFFFFu, mn);  // one shared-memory atomic per warp, not per thread
    mx = __reduce_max_sync(0xFFFFFFFFu, mx);
    if (lane == 0 && mx >= 0) {
        atomicMin(&s_wmin, mn);
        atomicMax(&s_wmax, mx);
    }
    __syncthreads();
    if (s_wmax < 0) return;  // nothing of this block of windows in these nodes
    const int wmin = s_wmin, range = s_wmax - s_wmin + 1;
    if (range > WIN_BINS) {  // rare: straight global atomics
#pragma unroll
        for (int k = 0; k < WIN_ITEMS; ++k) {
            if (w[k] < 0) continue;
            if (!SCATTER) {
                atomicAdd(counts + w[k], 1ull);
            } else {
                const int64_t pos = offsets[w[k]] + (int64_t)atomicAdd(cursor + w[k], 1ull);
                loc_out[pos] = first + i0 + (int64_t)k * 256;
                gid_out[pos] = g[k];
            }
        }
        return;
    }
    for (int t = tid; t < range; t += 256) s_cnt[t] = 0u;
    __syncthreads();
    unsigned rank[WIN_ITEMS];
#pragma unroll
    for (int k = 0; k < WIN_ITEMS; ++k) {
        // consecutive nodes mostly share a window: the lanes of a warp with the same window take their ranks with ONE
        // shared-memory atomic (32 same-address atomics of a warp serialize)
        const unsigned grp = __match_any_sync(0xFFFFFFFFu, w[k]);
        const int leader = __ffs(grp) - 1;
        unsigned base = 0u;
        if (lane == leader && w[k] >= 0) base = atomicAdd(&s_cnt[w[k] - wmin], (unsigned)__popc(grp));
        base = __shfl_sync(0xFFFFFFFFu, base, leader);
        rank[k] = base + __popc(grp & ((1u << lane) - 1u));
    }
    __syncthreads();
    for (int t = tid; t < range; t += 256) {
        if (!s_cnt[t]) continue;
        if (!SCATTER) atomicAdd(counts + wmin + t, (unsigned long long)s_cnt[t]);
        else s_base[t] = atomicAdd(cursor + wmin + t, (unsigned long long)s_cnt[t]);
    }
    if (!SCATTER) return;
    __syncthreads();
#pragma unroll
    for (int k = 0; k < WIN_ITEMS; ++k) {
        if (w[k] < 0) continue;
        const int64_t pos = offsets[w[k]] + (int64_t)s_base[w[k] - wmin] + rank[k];
        loc_out[pos] = first + i0 + (int64_t)k * 256;
        gid_out[pos] = g[k];
    }
}
}  // namespace acx

extern "C" int acx_bfs_window_partition(const int64_t* gid, int64_t n, int64_t lo, int64_t chunk, int64_t n_windows,
                                        uint64_t* counts, const int64_t* offsets, uint64_t* cursor, int64_t first,
                                        int64_t* loc_out, int64_t* gid_out, int phase, acx_stream_t stream) {
    if (n < 0 || chunk < 1 || n_windows < 1 || n_windows > 0x7FFFFFFF) return acx_fail(ACX_ERR_ARG, "acx_bfs_window_partition: bad argument");
    if (n == 0) return ACX_OK;
    const unsigned grid = (unsigned)((n + WIN_ITEMS * 256 - 1) / (WIN_ITEMS * 256));
    if (phase == 0) {
        if (!gid || !counts) return acx_fail(ACX_ERR_ARG, "acx_bfs_window_partition: null pointer");
        k_win_part<false><<<grid, 256, 0, (cudaStream_t)stream>>>(gid, n, lo, chunk, n_windows, reinterpret_cast<unsigned long long*>(counts), nullptr,
                                                                 nullptr, first, nullptr, nullptr);
    } else {
        if (!gid || !offsets || !cursor || !loc_out || !gid_out) return acx_fail(ACX_ERR_ARG, "acx_bfs_window_partition: null pointer");
        k_win_part<true><<<grid, 256, 0, (cudaStream_t)stream>>>(gid, n, lo, chunk, n_windows, nullptr, offsets,
                                                                reinterpret_cast<unsigned long long*>(cursor), first, loc_out, gid_out);
    }
    ACX_CUDA(cudaGetLastError());
    return ACX_OK;
}
