/*
 * oracle/barcode_oracle.c -- CPU restatement of the reference's native C++ programs
 * (barcode_analysis/simplex_data_generation/{classic,prime}_moves/ac_bfs.cpp and
 * barcode_analysis/5_steps_neibourhoods/neibourhoods.cpp with their AC_UTILS headers).
 *
 * THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, bench legs that time the CPU
 * baseline and __graft_entry__.smoke() may load it, and only as the checker.
 *
 * Parity is PINNED: oracle/Makefile compiles the reference programs themselves into oracle/_ref/
 * (from the sources where they lie under /root/reference) and tests/test_barcode_oracle.py /
 * tests/golden/make_barcode_golden.py compare this restatement with their output files; the
 * known answers of 5_steps_neibourhoods/README.txt:33-37 (28631, 49668, 72392) are checked too.
 *
 * Words are int8 letters a=1 A=-1 b=2 B=-2 (AC_UTILS_as_sets.h:11-14).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define BCO_MAXLEN 1024

typedef struct {
    int n;
    int8_t w[BCO_MAXLEN];
} word_t;

typedef struct {
    word_t r[2];
} pres_t;

/* operator<(Relator, Relator): AC_UTILS_as_sets.h:18-36 / AC_UTILS_no_hash.cpp:20-38 */
static int word_less(const word_t *l, const word_t *r) {
    if (l->n != r->n) return l->n < r->n;
    int i = 0;
    while (i < l->n && l->w[i] == r->w[i]) i++;
    if (i == l->n) return 0;
    return l->w[i] < r->w[i];
}

/* reduce0_ (AC_UTILS_as_sets.h:220-240): ONE left-to-right pass, kept literal (including the test of the
 * last letter against its left neighbour whether or not that neighbour was consumed) */
static void reduce0(const word_t *rel, word_t *out) {
    if (rel->n < 2) {
        *out = *rel;
        return;
    }
    out->n = 0;
    int j = 0;
    while (j < rel->n - 1) {
        if (rel->w[j] + rel->w[j + 1] == 0) j += 2;
        else out->w[out->n++] = rel->w[j++];
    }
    if (rel->w[rel->n - 1] + rel->w[rel->n - 2] != 0) out->w[out->n++] = rel->w[rel->n - 1];
}

/* reduce_ (:242-249): passes until a fixed point */
static void reduce(word_t *rel) {
    word_t out;
    reduce0(rel, &out);
    while (out.n != rel->n || memcmp(out.w, rel->w, (size_t)rel->n) != 0) {
        *rel = out;
        reduce0(rel, &out);
    }
    *rel = out;
}

/* conj0_ (:251-259): x^-1 rel x, reduced */
static void conj0(const word_t *rel, int x, word_t *out) {
    out->n = rel->n + 2;
    out->w[0] = (int8_t)(-x);
    memcpy(out->w + 1, rel->w, (size_t)rel->n);
    out->w[out->n - 1] = (int8_t)x;
    reduce(out);
}

/* inv0_ (:261-267) */
static void inv0(const word_t *rel, word_t *out) {
    out->n = rel->n;
    for (int i = 0; i < rel->n; i++) out->w[i] = (int8_t)(-rel->w[rel->n - i - 1]);
}

/* concat_ (:269-276) */
static void concat(const word_t *a, const word_t *b, word_t *out) {
    out->n = a->n + b->n;
    memcpy(out->w, a->w, (size_t)a->n);
    memcpy(out->w + a->n, b->w, (size_t)b->n);
    reduce(out);
}

/* Presentation(rel1, rel2) (:283-289) / sort_ (AC_UTILS_no_hash.cpp:120-136) */
static void sorted_pair(const word_t *x, const word_t *y, pres_t *out) {
    if (word_less(x, y)) {
        out->r[0] = *x;
        out->r[1] = *y;
    } else {
        out->r[0] = *y;
        out->r[1] = *x;
    }
}

/* Presentation::move (AC_UTILS_as_sets.h:298-363), move (AC_UTILS_no_hash.cpp:144-210).
 * Returns 0, or -1 when a word would exceed BCO_MAXLEN. */
static int pres_move(const pres_t *p, int t, int classic, pres_t *out) {
    static const int8_t CL[4] = {1, 2, -1, -2}; /* a b A B */
    static const int8_t PR[4] = {-2, -1, 1, 2}; /* B A a b */
    const word_t *r1 = &p->r[0], *r2 = &p->r[1];
    word_t x, y;
    if (r1->n + r2->n + 2 > BCO_MAXLEN) return -1;
    if (classic) {
        if (t == 0) { concat(r1, r2, &x); sorted_pair(&x, r2, out); }
        else if (t == 1) { concat(r2, r1, &x); sorted_pair(&x, r2, out); }
        else if (t == 2) { concat(r1, r2, &x); sorted_pair(r1, &x, out); }
        else if (t == 3) { concat(r2, r1, &x); sorted_pair(r1, &x, out); }
        else if (t < 8) { conj0(r2, CL[t - 4], &x); sorted_pair(r1, &x, out); }
        else if (t < 12) { conj0(r1, CL[t - 8], &x); sorted_pair(&x, r2, out); }
        else if (t == 12) { inv0(r1, &x); sorted_pair(&x, r2, out); }
        else { inv0(r2, &x); sorted_pair(r1, &x, out); }
    } else {
        if (t == 0) { concat(r1, r2, &x); sorted_pair(&x, r2, out); }
        else if (t == 1) { concat(r2, r1, &x); sorted_pair(r1, &x, out); }
        else if (t == 2) { inv0(r2, &y); concat(r1, &y, &x); sorted_pair(&x, r2, out); }
        else if (t == 3) { inv0(r1, &y); concat(r2, &y, &x); sorted_pair(r1, &x, out); }
        else if (t < 8) { conj0(r1, PR[t - 4], &x); sorted_pair(&x, r2, out); }
        else { conj0(r2, PR[t - 8], &x); sorted_pair(r1, &x, out); }
    }
    return 0;
}

/* ---- visited set: open addressing over an arena of variable-length keys (std::set stand-in) ---- */
typedef struct {
    int64_t *slot; /* node id + 1, 0 = empty */
    int64_t cap, count;
    int8_t *arena; /* per node: n1, n2 as 2x int16, then the letters */
    int64_t arena_len, arena_cap;
    int64_t *off; /* arena offset per node */
    int64_t off_cap;
} vset_t;

static uint64_t key_hash(const pres_t *p) {
    uint64_t h = 1469598103934665603ull;
    for (int k = 0; k < 2; k++) {
        h = (h ^ (uint64_t)p->r[k].n) * 1099511628211ull;
        for (int i = 0; i < p->r[k].n; i++) h = (h ^ (uint8_t)p->r[k].w[i]) * 1099511628211ull;
    }
    h ^= h >> 31;
    return h;
}

static int key_equal(const vset_t *s, int64_t id, const pres_t *p) {
    const int8_t *a = s->arena + s->off[id];
    int16_t n[2];
    memcpy(n, a, 4);
    if (n[0] != p->r[0].n || n[1] != p->r[1].n) return 0;
    return memcmp(a + 4, p->r[0].w, (size_t)n[0]) == 0 && memcmp(a + 4 + n[0], p->r[1].w, (size_t)n[1]) == 0;
}

static void vset_get(const vset_t *s, int64_t id, pres_t *p) {
    const int8_t *a = s->arena + s->off[id];
    int16_t n[2];
    memcpy(n, a, 4);
    p->r[0].n = n[0];
    p->r[1].n = n[1];
    memcpy(p->r[0].w, a + 4, (size_t)n[0]);
    memcpy(p->r[1].w, a + 4 + n[0], (size_t)n[1]);
}

static void vset_init(vset_t *s) {
    memset(s, 0, sizeof(*s));
    s->cap = 1 << 16;
    s->slot = (int64_t *)calloc((size_t)s->cap, sizeof(int64_t));
    s->arena_cap = 1 << 20;
    s->arena = (int8_t *)malloc((size_t)s->arena_cap);
    s->off_cap = 1 << 14;
    s->off = (int64_t *)malloc((size_t)s->off_cap * sizeof(int64_t));
}

static void vset_free(vset_t *s) {
    free(s->slot);
    free(s->arena);
    free(s->off);
}

static void vset_grow(vset_t *s) {
    int64_t ncap = s->cap * 2;
    int64_t *ns = (int64_t *)calloc((size_t)ncap, sizeof(int64_t));
    pres_t p;
    for (int64_t id = 0; id < s->count; id++) {
        vset_get(s, id, &p);
        uint64_t h = key_hash(&p) & (uint64_t)(ncap - 1);
        while (ns[h]) h = (h + 1) & (uint64_t)(ncap - 1);
        ns[h] = id + 1;
    }
    free(s->slot);
    s->slot = ns;
    s->cap = ncap;
}

/* returns the id of p; *is_new = 1 when it was inserted (id = insertion order) */
static int64_t vset_find_or_insert(vset_t *s, const pres_t *p, int insert, int *is_new) {
    uint64_t h = key_hash(p) & (uint64_t)(s->cap - 1);
    *is_new = 0;
    while (s->slot[h]) {
        if (key_equal(s, s->slot[h] - 1, p)) return s->slot[h] - 1;
        h = (h + 1) & (uint64_t)(s->cap - 1);
    }
    if (!insert) return -1;
    const int64_t need = 4 + p->r[0].n + p->r[1].n;
    if (s->arena_len + need > s->arena_cap) {
        while (s->arena_len + need > s->arena_cap) s->arena_cap *= 2;
        s->arena = (int8_t *)realloc(s->arena, (size_t)s->arena_cap);
    }
    if (s->count == s->off_cap) {
        s->off_cap *= 2;
        s->off = (int64_t *)realloc(s->off, (size_t)s->off_cap * sizeof(int64_t));
    }
    const int64_t id = s->count++;
    s->off[id] = s->arena_len;
    int16_t n[2] = {(int16_t)p->r[0].n, (int16_t)p->r[1].n};
    memcpy(s->arena + s->arena_len, n, 4);
    memcpy(s->arena + s->arena_len + 4, p->r[0].w, (size_t)n[0]);
    memcpy(s->arena + s->arena_len + 4 + n[0], p->r[1].w, (size_t)n[1]);
    s->arena_len += need;
    s->slot[h] = id + 1;
    *is_new = 1;
    if (s->count * 2 > s->cap) vset_grow(s);
    return id;
}

static int row_to_pres(const int8_t *row, int m, pres_t *p) {
    word_t a, b;
    a.n = b.n = 0;
    for (int i = 0; i < m; i++)
        if (row[i] != 0) a.w[a.n++] = row[i];
    for (int i = m; i < 2 * m; i++)
        if (row[i] != 0) b.w[b.n++] = row[i];
    if (a.n == 0 || b.n == 0) return -1;
    sorted_pair(&a, &b, p); /* neibourhoods.cpp:88 */
    return 0;
}

/* One move on a zero-padded row [r1 | r2] (m slots each); out gets the sorted child.  Returns the total
 * length, or -1 when a relator does not fit m slots / is empty. */
int bco_move(const int8_t *row, int m, int t, int classic, int8_t *out) {
    pres_t p, c;
    if (row_to_pres(row, m, &p)) return -1;
    if (pres_move(&p, t, classic, &c)) return -1;
    if (c.r[0].n > m || c.r[1].n > m || c.r[0].n == 0 || c.r[1].n == 0) return -1;
    memset(out, 0, (size_t)(2 * m));
    memcpy(out, c.r[0].w, (size_t)c.r[0].n);
    memcpy(out + m, c.r[1].w, (size_t)c.r[1].n);
    return c.r[0].n + c.r[1].n;
}

typedef struct {
    int64_t n_nodes, n_edges;
    uint16_t *node_size;  /* zero_filtrations: size of node id */
    int64_t *edge;        /* one_simplices: (cn, cc) pairs in emission order */
    uint16_t *edge_filt;  /* one_filtrations */
    int8_t *rows;         /* node presentations, [n_nodes, 2*row_m] zero padded */
    int row_m;
} bco_graph_t;

/* ac_bfs.cpp:13-97 main(): BFS from ({1},{2}) over presentations of total size <= n_max. */
int bco_graph_bfs(int n_max, int classic, bco_graph_t *g) {
    const int nm = classic ? 14 : 12; /* ac_bfs.cpp:56 */
    vset_t vs;
    vset_init(&vs);
    int64_t ecap = 1 << 16, ncap = 1 << 14;
    memset(g, 0, sizeof(*g));
    g->edge = (int64_t *)malloc((size_t)ecap * 2 * sizeof(int64_t));
    g->edge_filt = (uint16_t *)malloc((size_t)ecap * sizeof(uint16_t));
    g->node_size = (uint16_t *)malloc((size_t)ncap * sizeof(uint16_t));
    pres_t cur, child;
    word_t ra = {1, {1}}, rb = {1, {2}};
    sorted_pair(&ra, &rb, &cur); /* :40 */
    int is_new;
    vset_find_or_insert(&vs, &cur, 1, &is_new);
    g->node_size[0] = 2;
    /* the FIFO is the id order itself: ids are handed out in push order (:61-69) */
    for (int64_t head = 0; head < vs.count; head++) {
        vset_get(&vs, head, &cur);
        const int csize = cur.r[0].n + cur.r[1].n;
        for (int k = 0; k < nm; k++) {
            if (pres_move(&cur, k, classic, &child)) return -1;
            const int sz = child.r[0].n + child.r[1].n;
            if (sz > n_max) continue; /* :59 and :72 */
            if (child.r[0].n == 0) return -2; /* Hash() would read out of bounds */
            const int64_t cc = vset_find_or_insert(&vs, &child, 1, &is_new);
            if (is_new) {
                if (cc == ncap) {
                    ncap *= 2;
                    g->node_size = (uint16_t *)realloc(g->node_size, (size_t)ncap * sizeof(uint16_t));
                }
                g->node_size[cc] = (uint16_t)sz;
            }
            if (head < cc) { /* :75-78 */
                if (g->n_edges == ecap) {
                    ecap *= 2;
                    g->edge = (int64_t *)realloc(g->edge, (size_t)ecap * 2 * sizeof(int64_t));
                    g->edge_filt = (uint16_t *)realloc(g->edge_filt, (size_t)ecap * sizeof(uint16_t));
                }
                g->edge[2 * g->n_edges] = head;
                g->edge[2 * g->n_edges + 1] = cc;
                g->edge_filt[g->n_edges] = (uint16_t)(csize > sz ? csize : sz);
                g->n_edges++;
            }
        }
    }
    g->n_nodes = vs.count;
    g->row_m = n_max > 1 ? n_max - 1 : 1;
    g->rows = (int8_t *)calloc((size_t)g->n_nodes * 2 * (size_t)g->row_m, 1);
    for (int64_t id = 0; id < vs.count; id++) {
        vset_get(&vs, id, &cur);
        memcpy(g->rows + id * 2 * g->row_m, cur.r[0].w, (size_t)cur.r[0].n);
        memcpy(g->rows + id * 2 * g->row_m + g->row_m, cur.r[1].w, (size_t)cur.r[1].n);
    }
    vset_free(&vs);
    return 0;
}

void bco_graph_free(bco_graph_t *g) {
    free(g->node_size);
    free(g->edge);
    free(g->edge_filt);
    free(g->rows);
    memset(g, 0, sizeof(*g));
}

/* neibourhood() (neibourhoods.cpp:18-54): number of presentations within `radius` moves of the row. */
int64_t bco_neighbourhood(const int8_t *row, int m, int radius, int classic) {
    const int nm = classic ? 14 : 12;
    vset_t vs;
    vset_init(&vs);
    pres_t cur, child;
    if (row_to_pres(row, m, &cur)) return -1;
    int is_new;
    vset_find_or_insert(&vs, &cur, 1, &is_new);
    int64_t level_end = 1, head = 0;
    for (int dist = 0; dist < radius; dist++) { /* only nodes with dist < radius add children (:41) */
        for (; head < level_end; head++) {
            vset_get(&vs, head, &cur);
            for (int k = 0; k < nm; k++) {
                if (pres_move(&cur, k, classic, &child)) {
                    vset_free(&vs);
                    return -1;
                }
                vset_find_or_insert(&vs, &child, 1, &is_new);
            }
        }
        level_end = vs.count;
    }
    const int64_t size = vs.count;
    vset_free(&vs);
    return size;
}
