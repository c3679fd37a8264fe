/*
 * oracle/ac_search_oracle.c -- greedy_search and bfs restated in plain C.
 * TEST INFRASTRUCTURE ONLY (see ac_oracle.h).
 *
 * Reference behaviour reproduced (paths relative to /root/reference):
 *   greedy: ac_solver/search/greedy.py:15-121   heap order = (total_len, depth,
 *           state compared element-wise as signed ints); success test
 *           `new_length == 2` before the visited test (:91-100); budget test on
 *           len(tree_nodes) once per popped node (:115-119); failure returns the
 *           last popped node's path + (11, last child's length) (:121).
 *   bfs:    ac_solver/search/breadth_first.py:15-97   FIFO; failure returns None.
 */
#include "ac_oracle.h"
#include <stdlib.h>
#include <string.h>

typedef struct {
    int m, row;          /* row = 2m bytes */
    int64_t cap, n;      /* node store */
    int8_t *states;      /* [cap][row] */
    int32_t *parent;     /* [cap] */
    int32_t *depth;      /* [cap] */
    uint8_t *action;     /* [cap] */
    int16_t *total;      /* [cap] */
    int64_t tcap;        /* hash set (power of two) of node ids + 1 */
    int32_t *table;
} store_t;

static uint64_t hash_row(const int8_t *s, int row) {
    uint64_t h = 1469598103934665603ull;
    for (int k = 0; k < row; ++k) {
        h ^= (uint8_t)s[k];
        h *= 1099511628211ull;
    }
    return h ^ (h >> 29);
}

static int store_init(store_t *st, int m, int64_t budget) {
    memset(st, 0, sizeof *st);
    st->m = m;
    st->row = 2 * m;
    st->cap = budget + 16;
    st->tcap = 64;
    while (st->tcap < 2 * st->cap) st->tcap <<= 1;
    st->states = malloc((size_t)st->cap * st->row);
    st->parent = malloc((size_t)st->cap * sizeof(int32_t));
    st->depth = malloc((size_t)st->cap * sizeof(int32_t));
    st->action = malloc((size_t)st->cap);
    st->total = malloc((size_t)st->cap * sizeof(int16_t));
    st->table = calloc((size_t)st->tcap, sizeof(int32_t));
    return st->states && st->parent && st->depth && st->action && st->total && st->table;
}

static void store_free(store_t *st) {
    free(st->states); free(st->parent); free(st->depth);
    free(st->action); free(st->total); free(st->table);
}

/* returns the new node id, or -1 when the state was already present */
static int64_t store_add(store_t *st, const int8_t *s, int parent, int action, int total, int depth) {
    uint64_t mask = (uint64_t)st->tcap - 1, h = hash_row(s, st->row) & mask;
    while (st->table[h]) {
        if (memcmp(st->states + (int64_t)(st->table[h] - 1) * st->row, s, (size_t)st->row) == 0)
            return -1;
        h = (h + 1) & mask;
    }
    int64_t id = st->n++;
    memcpy(st->states + id * st->row, s, (size_t)st->row);
    st->parent[id] = parent;
    st->action[id] = (uint8_t)action;
    st->total[id] = (int16_t)total;
    st->depth[id] = depth;
    st->table[h] = (int32_t)(id + 1);
    return id;
}

/* path of node `id` followed by one extra (action, len) entry */
static int emit_path(const store_t *st, int64_t id, int last_action, int last_len,
                     int32_t *pa, int32_t *pl, int cap) {
    int d = st->depth[id];
    int n = d + 2;
    if (n > cap) return -n;
    pa[n - 1] = last_action;
    pl[n - 1] = last_len;
    for (int k = d; k >= 0; --k) {
        pa[k] = k == 0 ? -1 : st->action[id];
        pl[k] = st->total[id];
        id = st->parent[id];
    }
    return n;
}

/* heap of node ids ordered like Python's tuple compare in greedy.py:104-113 */
static int node_less(const store_t *st, int32_t a, int32_t b) {
    if (st->total[a] != st->total[b]) return st->total[a] < st->total[b];
    if (st->depth[a] != st->depth[b]) return st->depth[a] < st->depth[b];
    const int8_t *x = st->states + (int64_t)a * st->row, *y = st->states + (int64_t)b * st->row;
    for (int k = 0; k < st->row; ++k)
        if (x[k] != y[k]) return x[k] < y[k];
    return 0;
}

static void heap_push(const store_t *st, int32_t *heap, int64_t *hn, int32_t id) {
    int64_t k = (*hn)++;
    heap[k] = id;
    while (k > 0) {
        int64_t p = (k - 1) / 2;
        if (!node_less(st, heap[k], heap[p])) break;
        int32_t t = heap[k]; heap[k] = heap[p]; heap[p] = t;
        k = p;
    }
}

static int32_t heap_pop(const store_t *st, int32_t *heap, int64_t *hn) {
    int32_t top = heap[0];
    int64_t n = --(*hn);
    heap[0] = heap[n];
    int64_t k = 0;
    for (;;) {
        int64_t l = 2 * k + 1, r = l + 1, s = k;
        if (l < n && node_less(st, heap[l], heap[s])) s = l;
        if (r < n && node_less(st, heap[r], heap[s])) s = r;
        if (s == k) break;
        int32_t t = heap[k]; heap[k] = heap[s]; heap[s] = t;
        k = s;
    }
    return top;
}

static int search(const int8_t *root, int m, int64_t budget, int cyclical, int greedy,
                  int32_t *pa, int32_t *pl, int cap, int *path_n,
                  int64_t *n_seen, int64_t *n_expanded) {
    store_t st;
    if (m <= 0 || m > 1024 || !orc_is_valid(root, m)) return -2;
    if (budget < 1) budget = 1;
    if (!store_init(&st, m, budget)) { store_free(&st); return -3; }
    int32_t *order = malloc((size_t)st.cap * sizeof(int32_t)); /* heap or FIFO */
    int8_t *child = malloc((size_t)st.row);
    int64_t hn = 0, head = 0, tail = 0, expanded = 0;
    int rc = 0, l[2];
    *path_n = 0;

    int len0 = 0;
    for (int k = 0; k < st.row; ++k) len0 += root[k] != 0;
    store_add(&st, root, -1, 0xff, len0, 0);
    if (greedy) heap_push(&st, order, &hn, 0); else order[tail++] = 0;

    int32_t cur = 0;
    int last_len = len0;
    while (greedy ? hn > 0 : head < tail) {
        cur = greedy ? heap_pop(&st, order, &hn) : order[head++];
        ++expanded;
        const int8_t *s = st.states + (int64_t)cur * st.row;
        for (int a = 0; a < 12; ++a) {
            int status = orc_acmove(a, s, child, m, cyclical, l);
            if (status != ORC_OK) { rc = -1; goto out; } /* reference raises */
            last_len = l[0] + l[1];
            if (last_len == 2) {
                int n = emit_path(&st, cur, a, 2, pa, pl, cap);
                if (n < 0) { rc = -4; goto out; }
                *path_n = n;
                rc = 1;
                goto out;
            }
            int64_t id = store_add(&st, child, cur, a, last_len, st.depth[cur] + 1);
            if (id >= 0) {
                if (greedy) heap_push(&st, order, &hn, (int32_t)id); else order[tail++] = (int32_t)id;
            }
        }
        if (st.n >= budget) break;
    }
    if (greedy) { /* greedy.py:121 */
        int n = emit_path(&st, cur, 11, last_len, pa, pl, cap);
        if (n < 0) { rc = -4; goto out; }
        *path_n = n;
    }
out:
    if (n_seen) *n_seen = st.n;
    if (n_expanded) *n_expanded = expanded;
    free(order); free(child);
    store_free(&st);
    return rc;
}

int orc_greedy(const int8_t *root, int m, int64_t budget, int cyclical,
               int32_t *pa, int32_t *pl, int cap, int *path_n, int64_t *n_seen, int64_t *n_expanded) {
    return search(root, m, budget, cyclical, 1, pa, pl, cap, path_n, n_seen, n_expanded);
}

int orc_bfs(const int8_t *root, int m, int64_t budget, int cyclical,
            int32_t *pa, int32_t *pl, int cap, int *path_n, int64_t *n_seen, int64_t *n_expanded) {
    return search(root, m, budget, cyclical, 0, pa, pl, cap, path_n, n_seen, n_expanded);
}
