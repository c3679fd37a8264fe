// acx_graph.cuh -- level-synchronous BFS over UNORDERED relator pairs with full free reduction: the
// reference's native C++ programs
//     barcode_analysis/simplex_data_generation/{classic,prime}_moves/ac_bfs.cpp   (graph + filtrations)
//     barcode_analysis/5_steps_neibourhoods/neibourhoods.cpp                       (ball sizes)
// on bit planes.  A node is 16*W bytes: the sign / type planes of both relators (W 32-bit words each,
// acx_planes.cuh) with the two lengths and a 16-bit instance number folded into the top byte of the last
// word of each plane, so a relator holds at most 32*W - 8 letters.  W = 2 gives 32-byte nodes (one DRAM
// sector, relators <= 56 letters: ac_bfs.cpp is "designed for relators up to ~40", README.txt);
// W = 8 gives 128-byte nodes (248 letters: radius-5 balls around every Miller-Schupp presentation).
//
// Everything a kernel does per element is a __host__ __device__ functor here, and one level step is the
// template graph_step<W, Exec>: libacx.so instantiates it with an executor that launches kernels
// (acx_graph.cu); tests/host_harness instantiates the same code with a sequential executor so that the CPU
// suite checks the whole pipeline against the oracle before any GPU time is spent.
#pragma once
#include "../../include/acx.h"
#include "acx_planes.cuh"

#ifndef __CUDACC__
#error "acx_graph.cuh is CUDA C++ (nvcc)"
#endif

namespace acx {

// ---- the atomics the table protocol needs (host versions: sequential executor of the test harness) ----
#ifdef __CUDA_ARCH__
ACX_HD uint64_t g_cas(uint64_t* p, uint64_t cmp, uint64_t val) {
    return atomicCAS(reinterpret_cast<unsigned long long*>(p), (unsigned long long)cmp, (unsigned long long)val);
}
ACX_HD void g_min(uint64_t* p, uint64_t v) { atomicMin(reinterpret_cast<unsigned long long*>(p), (unsigned long long)v); }
ACX_HD void g_add(int64_t* p, int64_t v) { atomicAdd(reinterpret_cast<unsigned long long*>(p), (unsigned long long)v); }
ACX_HD uint64_t g_ld(const uint64_t* p) { return __ldcg(p); }
#else
ACX_HD uint64_t g_cas(uint64_t* p, uint64_t cmp, uint64_t val) {
    const uint64_t old = *p;
    if (old == cmp) *p = val;
    return old;
}
ACX_HD void g_min(uint64_t* p, uint64_t v) {
    if (v < *p) *p = v;
}
ACX_HD void g_add(int64_t* p, int64_t v) { *p += v; }
ACX_HD uint64_t g_ld(const uint64_t* p) { return *p; }
#endif

static constexpr uint64_t G_PROV = 1ull << 47;          // slot holds a claim of this round (value = candidate index)
static constexpr uint64_t G_VAL = G_PROV - 1;           // node id / candidate index
static constexpr uint64_t G_TAG = 0xFFFFull << 48;      // 16 hash bits, never zero
enum { G_MOVE_OK = 0, G_MOVE_OVERFLOW = 1, G_MOVE_EMPTY = 2 };

template <int W>
struct GPres {
    Rel<W> r[2];
    uint32_t inst;
};

template <int W>
struct GNode {
    uint32_t w[4 * W];
};

// Node / candidate access: 16*W bytes, 32-byte aligned (W is even).  On the device every 32 bytes move as ONE
// 256-bit global access (a whole DRAM sector per request, sm_100).
template <int W>
ACX_HD GNode<W> g_load(const uint32_t* base, int64_t id) {
    GNode<W> o;
#ifdef __CUDA_ARCH__
    const uint32_t* p = base + id * (4 * W);
#pragma unroll
    for (int i = 0; i < W / 2; ++i)
        asm volatile("ld.global.v8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                     : "=r"(o.w[8 * i]), "=r"(o.w[8 * i + 1]), "=r"(o.w[8 * i + 2]), "=r"(o.w[8 * i + 3]), "=r"(o.w[8 * i + 4]),
                       "=r"(o.w[8 * i + 5]), "=r"(o.w[8 * i + 6]), "=r"(o.w[8 * i + 7])
                     : "l"(p + 8 * i)
                     : "memory");
#else
    const uint4* p = reinterpret_cast<const uint4*>(base) + id * W;
    for (int i = 0; i < W; ++i) {
        const uint4 v = p[i];
        o.w[4 * i] = v.x;
        o.w[4 * i + 1] = v.y;
        o.w[4 * i + 2] = v.z;
        o.w[4 * i + 3] = v.w;
    }
#endif
    return o;
}

template <int W>
ACX_HD void g_store(uint32_t* base, int64_t id, const GNode<W>& n) {
#ifdef __CUDA_ARCH__
    uint32_t* p = base + id * (4 * W);
#pragma unroll
    for (int i = 0; i < W / 2; ++i)
        asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(p + 8 * i), "r"(n.w[8 * i]), "r"(n.w[8 * i + 1]),
                     "r"(n.w[8 * i + 2]), "r"(n.w[8 * i + 3]), "r"(n.w[8 * i + 4]), "r"(n.w[8 * i + 5]), "r"(n.w[8 * i + 6]),
                     "r"(n.w[8 * i + 7])
                     : "memory");
#else
    uint4* p = reinterpret_cast<uint4*>(base) + id * W;
    for (int i = 0; i < W; ++i) {
        uint4 v;
        v.x = n.w[4 * i];
        v.y = n.w[4 * i + 1];
        v.z = n.w[4 * i + 2];
        v.w = n.w[4 * i + 3];
        p[i] = v;
    }
#endif
}

template <int W>
ACX_HD bool g_equal(const GNode<W>& a, const GNode<W>& b) {
    uint32_t d = 0;
#pragma unroll
    for (int i = 0; i < 4 * W; ++i) d |= a.w[i] ^ b.w[i];
    return d == 0u;
}

template <int W>
ACX_HD uint64_t g_hash(const GNode<W>& n) {
    uint64_t h = 0x9E3779B97F4A7C15ull;
#pragma unroll
    for (int k = 0; k < 2 * W; ++k) {
        h ^= (uint64_t)n.w[2 * k] | ((uint64_t)n.w[2 * k + 1] << 32);
        h *= 0xFF51AFD7ED558CCDull;
        h ^= h >> 32;
    }
    h *= 0xC4CEB9FE1A85EC53ull;
    h ^= h >> 29;
    return h;
}
ACX_HD uint64_t g_tag(uint64_t h) { return ((h >> 48) | 1ull) << 48; }
ACX_HD uint64_t g_slot0(uint64_t h, uint64_t mask) { return (h ^ (h >> 23)) & mask; }

template <int W>
ACX_HD GNode<W> g_pack(const GPres<W>& p) {
    GNode<W> o;
#pragma unroll
    for (int i = 0; i < W; ++i) {
        o.w[i] = p.r[0].S.w[i];
        o.w[W + i] = p.r[0].T.w[i];
        o.w[2 * W + i] = p.r[1].S.w[i];
        o.w[3 * W + i] = p.r[1].T.w[i];
    }
    o.w[W - 1] |= (uint32_t)p.r[0].n << 24;
    o.w[2 * W - 1] |= (uint32_t)p.r[1].n << 24;
    o.w[3 * W - 1] |= (p.inst & 0xFFu) << 24;
    o.w[4 * W - 1] |= ((p.inst >> 8) & 0xFFu) << 24;
    return o;
}

template <int W>
ACX_HD GPres<W> g_unpack(const GNode<W>& n) {
    GPres<W> p;
#pragma unroll
    for (int i = 0; i < W; ++i) {
        const uint32_t keep = (i == W - 1) ? 0x00FFFFFFu : 0xFFFFFFFFu;
        p.r[0].S.w[i] = n.w[i] & keep;
        p.r[0].T.w[i] = n.w[W + i] & keep;
        p.r[1].S.w[i] = n.w[2 * W + i] & keep;
        p.r[1].T.w[i] = n.w[3 * W + i] & keep;
    }
    p.r[0].n = (int)(n.w[W - 1] >> 24);
    p.r[1].n = (int)(n.w[2 * W - 1] >> 24);
    p.inst = (n.w[3 * W - 1] >> 24) | ((n.w[4 * W - 1] >> 24) << 8);
    return p;
}

template <int W>
ACX_HD int g_total(const GNode<W>& n) {
    return (int)(n.w[W - 1] >> 24) + (int)(n.w[2 * W - 1] >> 24);
}

// operator<(Relator, Relator) (AC_UTILS_as_sets.h:18-36): shorter first, then lexicographic on the letter
// VALUES -2 < -1 < 1 < 2, i.e. B < A < a < b
template <int W>
ACX_HD bool g_less(const Rel<W>& a, const Rel<W>& b) {
    if (a.n != b.n) return a.n < b.n;
    const BV<W> d = bv_or<W>(bv_xor<W>(a.S, b.S), bv_xor<W>(a.T, b.T));
    const int pos = bv_ctz<W>(d);
    if (pos >= a.n) return false;  // equal words
    const uint32_t sa = bv_bit<W>(a.S, pos), ta = bv_bit<W>(a.T, pos), sb = bv_bit<W>(b.S, pos), tb = bv_bit<W>(b.T, pos);
    const uint32_t ra = sa ? (ta ? 0u : 1u) : (ta ? 3u : 2u), rb = sb ? (tb ? 0u : 1u) : (tb ? 3u : 2u);
    return ra < rb;
}

// Presentation(rel1, rel2) (AC_UTILS_as_sets.h:283-289) / sort_ (AC_UTILS_no_hash.cpp:120-136)
template <int W>
ACX_HD void g_sort(GPres<W>& p) {
    if (!g_less<W>(p.r[0], p.r[1])) {
        const Rel<W> t = p.r[0];
        p.r[0] = p.r[1];
        p.r[1] = t;
    }
}

// Presentation::move (AC_UTILS_as_sets.h:298-363; AC_UTILS_no_hash.cpp:144-210) on a SORTED parent whose
// relators are freely reduced and non-empty.  concat_ / conj0_ end in reduce_ (full free reduction,
// :242-276); for reduced operands that is the junction cancellation of rel_concat and the one-letter test at
// each end of rel_conjugate (conj0_(rel, x) = x^-1 rel x, i.e. g rel g^-1 with g = x^-1).
// An EMPTY relator (r1 = r2^+-1 was concatenated: the pair does not present the trivial group) is legal in
// neibourhoods.cpp, whose words are plain vectors -- conj0_ and inv0_ of the empty word are empty, concat_ with it
// is the other word, and it sorts first -- while ac_bfs.cpp would index out of bounds (Hash(), :118-139).
template <int W>
ACX_HD int g_move(const GPres<W>& p, int k, bool classic, int cap, bool allow_empty, GPres<W>& c) {
    int tgt;
    Rel<W> x;
    bool ok = true;
    const Rel<W>&r0 = p.r[0], &r1 = p.r[1];  // selects instead of indexed access: the relators stay in registers
    if (classic) {
        if (k < 4) {  // 0: (r1 r2, r2)  1: (r2 r1, r2)  2: (r1, r1 r2)  3: (r1, r2 r1)
            tgt = k >> 1;
            const bool sw = (k & 1) != 0;
            x = sw ? r1 : r0;
            ok = rel_concat<W>(x, sw ? r0 : r1, cap);
        } else if (k < 12) {  // 4..7: conj0_(r2, a b A B)   8..11: conj0_(r1, a b A B)
            tgt = k < 8 ? 1 : 0;
            const int g = (k - 4) & 3;
            x = tgt ? r1 : r0;
            if (x.n > 0) ok = rel_conjugate<W>(x, g >= 2 ? 0u : 1u, (uint32_t)(g & 1), cap);
        } else {  // 12: inv0_(r1)   13: inv0_(r2)
            tgt = k - 12;
            x = rel_inverse<W>(tgt ? r1 : r0);
        }
    } else {
        if (k < 4) {  // 0: (r1 r2, r2)  1: (r1, r2 r1)  2: (r1 r2^-1, r2)  3: (r1, r2 r1^-1)
            tgt = k & 1;
            x = tgt ? r1 : r0;
            Rel<W> o = tgt ? r0 : r1;
            if (k >= 2) o = rel_inverse<W>(o);
            ok = rel_concat<W>(x, o, cap);
        } else {  // 4..7: conj0_(r1, B A a b)   8..11: conj0_(r2, B A a b)
            tgt = k < 8 ? 0 : 1;
            const int g = (k - 4) & 3;
            x = tgt ? r1 : r0;
            if (x.n > 0) ok = rel_conjugate<W>(x, g < 2 ? 0u : 1u, (g == 0 || g == 3) ? 1u : 0u, cap);
        }
    }
    if (!ok) return G_MOVE_OVERFLOW;
    if (x.n == 0 && !allow_empty) return G_MOVE_EMPTY;
    c.r[0] = tgt ? r0 : x;
    c.r[1] = tgt ? x : r1;
    c.inst = p.inst;
    g_sort<W>(c);
    return G_MOVE_OK;
}

// ---------------------------------------------------------------------------------------------------
// functors: one call per element

template <int W>
struct GPackRows {  // int8 rows [r1 | r2] (non-zero entries of each half, neibourhoods.cpp:74-86) -> sorted nodes
    const int8_t* rows;
    int m;
    const int32_t* inst;
    uint32_t* nodes;
    uint8_t* ok;
    ACX_HD void operator()(int64_t j) const {
        GPres<W> p;
        bool good = true;
        const int cap = 32 * W - 8;
        for (int h = 0; h < 2; ++h) {
            Rel<W> r;
            r.S = bv_zero<W>();
            r.T = bv_zero<W>();
            int n = 0;
            int prev = 0;
            for (int k = 0; k < m; ++k) {
                const int v = rows[j * 2 * m + h * m + k];
                if (v == 0) continue;
                if (v != 1 && v != -1 && v != 2 && v != -2) good = false;
                if (v == -prev) good = false;  // not freely reduced: outside the plane domain
                if (n >= cap) {
                    good = false;
                    break;
                }
                if (v < 0) r.S.w[n >> 5] |= 1u << (n & 31);
                if (v == 2 || v == -2) r.T.w[n >> 5] |= 1u << (n & 31);
                prev = v;
                ++n;
            }
            if (n == 0) good = false;
            r.n = n;
            p.r[h] = r;
        }
        p.inst = inst ? (uint32_t)inst[j] : 0u;
        if (p.inst > 0xFFFFu) good = false;
        GNode<W> node;
        if (good) {
            g_sort<W>(p);
            node = g_pack<W>(p);
        } else {
#pragma unroll
            for (int i = 0; i < 4 * W; ++i) node.w[i] = 0u;
        }
        g_store<W>(nodes, j, node);
        ok[j] = good ? 1 : 0;
    }
};

template <int W>
struct GUnpackRows {  // nodes -> zero-padded int8 rows of 2*m (relators longer than m are truncated: caller sizes m)
    const uint32_t* nodes;
    int m;
    int8_t* rows;
    int32_t* inst;
    ACX_HD void operator()(int64_t j) const {
        const GPres<W> p = g_unpack<W>(g_load<W>(nodes, j));
        for (int h = 0; h < 2; ++h)
            for (int k = 0; k < m; ++k) {
                int v = 0;
                if (k < p.r[h].n) {
                    v = ((p.r[h].T.w[k >> 5] >> (k & 31)) & 1u) ? 2 : 1;
                    if ((p.r[h].S.w[k >> 5] >> (k & 31)) & 1u) v = -v;
                }
                rows[j * 2 * m + h * m + k] = (int8_t)v;
            }
        if (inst) inst[j] = (int32_t)p.inst;
    }
};

template <int W>
struct GRehash {  // insert node ids (all states distinct) into a zeroed table
    const uint32_t* nodes;
    uint64_t* table;
    int64_t tcap;
    ACX_HD void operator()(int64_t id) const {
        const uint64_t h = g_hash<W>(g_load<W>(nodes, id));
        const uint64_t mask = (uint64_t)tcap - 1, mine = g_tag(h) | (uint64_t)id;
        uint64_t s = g_slot0(h, mask);
        while (g_cas(table + s, 0ull, mine) != 0ull) s = (s + 1) & mask;
    }
};

template <int W>
struct GCountInst {
    const uint32_t* nodes;
    int64_t* counts;
    int64_t n_inst;
    ACX_HD void operator()(int64_t id) const {
        const GNode<W> n = g_load<W>(nodes, id);
        const int64_t inst = (int64_t)((n.w[3 * W - 1] >> 24) | ((n.w[4 * W - 1] >> 24) << 8));
        if (inst < n_inst) g_add(counts + inst, 1);
    }
};

// candidate c = (parent - parent_first) * n_moves + move: the position at which the reference's FIFO loop
// reaches the child (ac_bfs.cpp:52-57), so "lowest c wins" reproduces its discovery order.
//
// Expand pass: one block of 32 * n_moves threads per 32 consecutive parents.  Phase 1: warp k applies move k to the 32
// parents (one move per warp: no divergence between the concat / conjugate / inverse paths; coalesced parent loads) and
// parks the child in shared memory at its logical position lane * n_moves + k.  Phase 2: the block's candidates are one
// contiguous range of the per-candidate arrays, written back with consecutive threads on consecutive entries (whole
// sectors instead of 32 scattered 32-byte stores per instruction).
static constexpr uint64_t G_FOUND = 1ull << 62;  // ref[c] = G_FOUND | node id: the child is a committed node

template <int W>
struct GExpand {
    acx_graph_step_t a;
    int nm;
    ACX_HD static size_t smem_bytes(int nm) { return (size_t)32 * nm * (16 * W + 2); }
    ACX_HD void phase1(int64_t block, int tid, uint8_t* smem) const {
        const int lane = tid & 31, k = tid >> 5;
        const int64_t local = block * 32 + lane;
        uint32_t* scand = reinterpret_cast<uint32_t*>(smem);
        uint16_t* ssize = reinterpret_cast<uint16_t*>(smem + (size_t)32 * nm * 16 * W);
        const int j = lane * nm + k;
        if (local >= a.n_parents) {
            ssize[j] = 0;
            return;
        }
        const int cap = 32 * W - 8;
        const GPres<W> p = g_unpack<W>(g_load<W>(a.nodes, a.parent_first + local));
        GPres<W> ch;
        const int st = g_move<W>(p, k, a.classic != 0, cap, a.allow_empty != 0, ch);
        bool skip = false, err = false;
        int total = 0;
        if (st == G_MOVE_OVERFLOW) {
            // a relator longer than cap: with a size limit <= cap + 1 the child is over the limit anyway
            if (a.max_total > 0 && a.max_total <= cap + 1) skip = true;
            else err = true;
        } else if (st == G_MOVE_EMPTY) {
            err = true;  // a relator cancelled completely and the caller follows ac_bfs.cpp (see g_move)
        } else {
            total = ch.r[0].n + ch.r[1].n;
            if (a.max_total > 0 && total > a.max_total) skip = true;  // ac_bfs.cpp:59,72
        }
        if (err) {
            a.counts[2] = 1;
            skip = true;
        }
        ssize[j] = skip ? (uint16_t)0 : (uint16_t)total;
        if (!skip) {
            const GNode<W> n = g_pack<W>(ch);
#pragma unroll
            for (int i = 0; i < 4 * W; ++i) scand[(size_t)j * 4 * W + i] = n.w[i];
        }
    }
    ACX_HD void phase2(int64_t block, int tid, const uint8_t* smem) const {
        const uint32_t* scand = reinterpret_cast<const uint32_t*>(smem);
        const uint16_t* ssize = reinterpret_cast<const uint16_t*>(smem + (size_t)32 * nm * 16 * W);
        const int64_t c = block * 32 * nm + tid;
        if (c >= a.n_parents * nm) return;
        const uint16_t sz = ssize[tid];
        a.ref[c] = -1;
        a.csize[c] = sz;
        if (sz) {
            GNode<W> n;
#pragma unroll
            for (int i = 0; i < 4 * W; ++i) n.w[i] = scand[(size_t)tid * 4 * W + i];
            g_store<W>(a.cand, c, n);
        }
    }
};

// Visited-set probe (ac_bfs.cpp:59 `visited.find(child)`).  Slot = tag | PROV | candidate for a claim of this
// step, tag | node id for a committed node.  Equal candidates meet in one slot; the lowest index stays.
template <int W>
struct GProbe {
    acx_graph_step_t a;
    ACX_HD void operator()(int64_t c) const {
        if (a.csize[c] == 0) return;
        const GNode<W> s = g_load<W>(a.cand, c);
        const uint64_t h = g_hash<W>(s);
        const uint64_t tag = g_tag(h), mine = tag | G_PROV | (uint64_t)c, mask = (uint64_t)a.tcap - 1;
        uint64_t slot = g_slot0(h, mask);
        for (;;) {
            uint64_t e = g_ld(a.table + slot);
            if (e == 0ull) {
                e = g_cas(a.table + slot, 0ull, mine);
                if (e == 0ull) {
                    a.ref[c] = (int64_t)slot;
                    return;
                }
            }
            if ((e & G_TAG) == tag) {
                if (e & G_PROV) {
                    // a claim is only ever replaced by a claim of an equal state, so one comparison settles it
                    if (g_equal<W>(s, g_load<W>(a.cand, (int64_t)(e & G_VAL)))) {
                        g_min(a.table + slot, mine);
                        a.ref[c] = (int64_t)slot;
                        return;
                    }
                } else {
                    const int64_t id = (int64_t)(e & G_VAL);
                    if (g_equal<W>(s, g_load<W>(a.nodes, id))) {
                        a.ref[c] = (int64_t)(G_FOUND | (uint64_t)id);
                        return;
                    }
                }
            }
            slot = (slot + 1) & mask;
        }
    }
};

template <int W>
struct GResolve {  // flag = this candidate is the first discovery of a new state
    acx_graph_step_t a;
    ACX_HD void operator()(int64_t c) const {
        const int64_t s = a.ref[c];
        a.flag[c] = (s >= 0 && !((uint64_t)s & G_FOUND) && (a.table[s] & (G_PROV | G_VAL)) == (G_PROV | (uint64_t)c)) ? 1 : 0;
    }
};

struct GTotal {  // out[0] = number of flagged candidates
    const uint8_t* flag;
    const int32_t* excl;
    int64_t n;
    int64_t* out;
    ACX_HD void operator()(int64_t) const { out[0] = (int64_t)excl[n - 1] + (int64_t)flag[n - 1]; }
};

// Winners become nodes in candidate order (ac_bfs.cpp:60-69: id = n_nodes + exclusive count of winners before c);
// in the same pass every in-range child learns its name and flags its edge (ac_bfs.cpp:72-79: an edge whenever the
// child's name is larger than the parent's).  A loser may read the slot before or after the winner has replaced the
// claim by the node id: both forms name the same node (claim -> n_nodes + excl[winner]).
template <int W>
struct GCommitEdge {
    acx_graph_step_t a;
    int nm;
    ACX_HD void operator()(int64_t c) const {
        const bool win = a.flag[c] != 0;
        if (!win && !a.want_edges) return;
        uint8_t f = 0;
        if (a.csize[c] != 0) {
            const int64_t r = a.ref[c];
            int64_t cc;
            if (win) {
                cc = a.n_nodes + (int64_t)a.excl[c];
                g_store<W>(a.nodes, cc, g_load<W>(a.cand, c));
                a.node_size[cc] = a.csize[c];
                a.table[r] = (g_ld(a.table + r) & G_TAG) | (uint64_t)cc;
            } else if ((uint64_t)r & G_FOUND) {
                cc = (int64_t)((uint64_t)r & G_VAL);
            } else {
                const uint64_t e = g_ld(a.table + r);
                cc = (e & G_PROV) ? a.n_nodes + (int64_t)a.excl[(int64_t)(e & G_VAL)] : (int64_t)(e & G_VAL);
            }
            a.ref[c] = cc;
            f = (a.parent_first + c / nm) < cc ? 1 : 0;
        }
        if (a.want_edges) a.flag[c] = f;
    }
};

template <int W>
struct GEdgeWrite {
    acx_graph_step_t a;
    int nm;
    ACX_HD void operator()(int64_t c) const {
        if (!a.flag[c]) return;
        const int64_t pos = (int64_t)a.excl[c], par = a.parent_first + c / nm;
        a.edges[2 * pos] = par;
        a.edges[2 * pos + 1] = a.ref[c];
        const uint16_t ps = a.node_size[par], cs = a.csize[c];
        a.edge_filt[pos] = ps > cs ? ps : cs;  // ac_bfs.cpp:76
    }
};

// One step: expand parents [parent_first, parent_first + n_parents), dedup against the table, append the new
// nodes in discovery order and (optionally) emit the edge list of the step.
// Exec::run(n, functor), Exec::run_shuffled(n, functor) (order-free functors; the host executor permutes them),
// Exec::run_blocks(blocks, threads, smem bytes, functor with phase1 / phase2 separated by a block barrier),
// Exec::scan(flag, excl, n, tmp, bytes).
template <int W, class Exec>
int graph_step(Exec& ex, const acx_graph_step_t& a) {
    const int nm = a.classic ? 14 : 12;
    const int64_t C = a.n_parents * nm;
    if (C == 0) return 0;
    ex.run_blocks((a.n_parents + 31) / 32, 32 * nm, GExpand<W>::smem_bytes(nm), GExpand<W>{a, nm});
    ex.run_shuffled(C, GProbe<W>{a});
    ex.run_shuffled(C, GResolve<W>{a});
    if (int rc = ex.scan(a.flag, a.excl, C, a.scan_tmp, a.scan_tmp_bytes)) return rc;
    ex.run(1, GTotal{a.flag, a.excl, C, a.counts});
    ex.run_shuffled(C, GCommitEdge<W>{a, nm});
    if (a.want_edges) {
        if (int rc = ex.scan(a.flag, a.excl, C, a.scan_tmp, a.scan_tmp_bytes)) return rc;
        ex.run(1, GTotal{a.flag, a.excl, C, a.counts + 1});
        ex.run_shuffled(C, GEdgeWrite<W>{a, nm});
    }
    return 0;
}

}  // namespace acx
