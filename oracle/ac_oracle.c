/*
 * oracle/ac_oracle.c -- word arithmetic, ACMove and ACEnv.step restated in plain C.
 * TEST INFRASTRUCTURE ONLY (see ac_oracle.h).  Build: `make -C oracle`.
 *
 * The structure here is a fresh statement of the algorithm (explicit word
 * buffers + a stack for free reduction); it is not a line translation of the
 * numpy code, but each routine cites the reference lines whose observable
 * behaviour it must reproduce.
 */
#include "ac_oracle.h"
#include <string.h>
#include <stdlib.h>
#include <pthread.h>

#define ORC_MAX_M 1024

static int word_len(const int8_t *r, int m) {
    int n = 0;
    for (int k = 0; k < m; ++k) n += (r[k] != 0);
    return n;
}

/* utils.py:13-54 -- both halves non-empty and zeros only on the right */
int orc_is_valid(const int8_t *p, int m) {
    for (int h = 0; h < 2; ++h) {
        const int8_t *r = p + h * m;
        int n = word_len(r, m);
        if (n == 0) return 0;
        for (int k = n; k < m; ++k)
            if (r[k] != 0) return 0;
    }
    return 1;
}

/* utils.py:57-87 -- valid, both lengths 1, sorted |letters| == [1,2] */
int orc_is_trivial(const int8_t *p, int m) {
    if (!orc_is_valid(p, m)) return 0;
    if (word_len(p, m) != 1 || word_len(p + m, m) != 1) return 0;
    int a = abs(p[0]), b = abs(p[m]);
    return (a == 1 && b == 2) || (a == 2 && b == 1);
}

/* utils.py:211-220 -- the unique freely reduced form, via a stack */
static int free_reduce(int8_t *w, int n) {
    int top = 0;
    for (int k = 0; k < n; ++k) {
        if (top > 0 && w[top - 1] == -w[k]) --top;
        else w[top++] = w[k];
    }
    return top;
}

/* utils.py:223-232 -- number of letters cancelling across the two ends */
static int cyclic_depth(const int8_t *w, int n) {
    int k = 0;
    while (2 * k + 1 < n && w[k] == -w[n - 1 - k]) ++k;
    return k;
}

int orc_simplify_relator(int8_t *r, int m, int cyclical) {
    int n = word_len(r, m);
    n = free_reduce(r, n);
    if (cyclical && n > 0) {
        int k = cyclic_depth(r, n);
        if (k) {
            memmove(r, r + k, (size_t)(n - 2 * k));
            n -= 2 * k;
        }
    }
    memset(r + n, 0, (size_t)(m - n));
    return n;
}

int orc_simplify_presentation(int8_t *p, int m, int cyclical, int *lens) {
    if (!orc_is_valid(p, m)) return ORC_INVALID_INPUT;
    lens[0] = orc_simplify_relator(p, m, cyclical);
    lens[1] = orc_simplify_relator(p + m, m, cyclical);
    return ORC_OK;
}

/* gather the non-zero letters of a half-row (ac_moves.py:50-51, :110) */
static int gather(const int8_t *r, int m, int8_t *dst) {
    int n = 0;
    for (int k = 0; k < m; ++k)
        if (r[k] != 0) dst[n++] = r[k];
    return n;
}

static void store_word(int8_t *r, int m, const int8_t *w, int n) {
    memcpy(r, w, (size_t)n);
    memset(r + n, 0, (size_t)(m - n));
}

/* ac_moves.py:4-76 */
void orc_concatenate(int8_t *p, int m, int i, int sign) {
    int8_t a[ORC_MAX_M], b[ORC_MAX_M], out[2 * ORC_MAX_M];
    int j = 1 - i;
    int la = gather(p + i * m, m, a);
    int lb = gather(p + j * m, m, b);
    if (sign < 0) { /* r_j^-1 : reversed and negated (ac_moves.py:41-48) */
        for (int k = 0; k < lb / 2; ++k) {
            int8_t t = b[k];
            b[k] = b[lb - 1 - k];
            b[lb - 1 - k] = t;
        }
        for (int k = 0; k < lb; ++k) b[k] = (int8_t)-b[k];
    }
    /* junction-only cancellation (ac_moves.py:56-60) */
    int lim = la < lb ? la : lb, acc = 0;
    while (acc < lim && a[la - 1 - acc] == -b[acc]) ++acc;
    int n = la + lb - 2 * acc;
    if (n > m) return; /* guard (ac_moves.py:64): row unchanged */
    memcpy(out, a, (size_t)(la - acc));
    memcpy(out + (la - acc), b + acc, (size_t)(lb - acc));
    store_word(p + i * m, m, out, n);
}

/* ac_moves.py:79-156 */
void orc_conjugate(int8_t *p, int m, int i, int j, int sign) {
    int8_t a[ORC_MAX_M], out[ORC_MAX_M + 2];
    int la = gather(p + i * m, m, a);
    if (la == 0) return; /* reference would raise IndexError; callers validate first */
    int8_t g = (int8_t)(sign * j);
    int sc = (a[0] == -g), ec = (a[la - 1] == g); /* one letter each (:119-120) */
    int n = la + 2 - 2 * (sc + ec);
    if (n > m) return; /* guard (:126) */
    if (n < 0) n = 0;  /* la==1 cannot have sc&&ec; kept for safety */
    int w = 0;
    if (!sc) out[w++] = g;
    for (int k = sc; k < la - ec; ++k) out[w++] = a[k];
    if (!ec) out[w++] = (int8_t)-g;
    store_word(p + i * m, m, out, w);
}

/* decode table implied by ac_moves.py:192-206 */
static void decode_move(int id, int *is_conj, int *i, int *j, int *sign) {
    int t = id + 1;
    *i = t % 2;
    if (id < 4) {
        *is_conj = 0;
        *j = 1 - *i;
        *sign = (((t - *i) / 2) % 2) ? -1 : 1;
    } else {
        *is_conj = 1;
        int jp = ((t - *i) / 2) % 2;
        *sign = (((t - *i - 2 * jp) / 4) % 2) ? -1 : 1;
        *j = jp + 1;
    }
}

/* ac_moves.py:159-231.  Order: move (junction cancel + guard) -> validity
 * assert -> full free reduction of BOTH relators -> optional cyclic reduction. */
int orc_acmove(int move_id, const int8_t *in, int8_t *out, int m, int cyclical, int *lens) {
    int8_t tmp[2 * ORC_MAX_M];
    memcpy(out, in, (size_t)(2 * m));
    lens[0] = word_len(in, m);
    lens[1] = word_len(in + m, m);
    if (!orc_is_valid(in, m)) return ORC_INVALID_INPUT;
    memcpy(tmp, in, (size_t)(2 * m));
    int is_conj, i, j, sign;
    decode_move(move_id, &is_conj, &i, &j, &sign);
    if (is_conj) orc_conjugate(tmp, m, i, j, sign);
    else orc_concatenate(tmp, m, i, sign);
    int l2[2];
    if (orc_simplify_presentation(tmp, m, cyclical, l2) != ORC_OK)
        return ORC_EMPTY_RELATOR; /* utils.py:264-266 raises; row left unchanged */
    /* If free reduction (not the junction cancel) emptied a relator the
     * reference still returns the row: its assert runs before the reduction. */
    memcpy(out, tmp, (size_t)(2 * m));
    lens[0] = l2[0];
    lens[1] = l2[1];
    return ORC_OK;
}

/* ---- lockstep batches: rows split evenly over `threads` pthreads ---------- */
typedef struct {
    int kind; /* 0 = acmove, 1 = expand12, 2 = env step */
    int64_t lo, hi;
    const int8_t *in; const uint8_t *moves; int8_t *out; int32_t *lens; uint8_t *status;
    int m, cyclical, horizon;
    int8_t *state; int32_t *steps; const int8_t *pool; const int32_t *next_idx;
    int32_t *reward; uint8_t *done, *truncated, *lens8;
} job_t;

static void run_rows(const job_t *jb) {
    const int m = jb->m;
    for (int64_t r = jb->lo; r < jb->hi; ++r) {
        int l[2];
        if (jb->kind == 0) {
            int st = orc_acmove(jb->moves[r], jb->in + r * 2 * m, jb->out + r * 2 * m, m, jb->cyclical, l);
            jb->lens[2 * r] = l[0];
            jb->lens[2 * r + 1] = l[1];
            if (jb->status) jb->status[r] = (uint8_t)st;
        } else if (jb->kind == 1) {
            for (int a = 0; a < 12; ++a) {
                int64_t c = r * 12 + a;
                int st = orc_acmove(a, jb->in + r * 2 * m, jb->out + c * 2 * m, m, jb->cyclical, l);
                jb->lens[2 * c] = l[0];
                jb->lens[2 * c + 1] = l[1];
                if (jb->status) jb->status[c] = (uint8_t)st;
            }
        } else { /* ac_env.py:91-111 (+ :113-129 for the reset) */
            int8_t nxt[2 * ORC_MAX_M];
            int8_t *s = jb->state + r * 2 * m;
            const int32_t max_reward = jb->horizon * m * 2;          /* :76  */
            int st = orc_acmove(jb->moves[r], s, nxt, m, /*cyclical=*/1, l);
            memcpy(s, nxt, (size_t)(2 * m));
            int total = l[0] + l[1];
            int d = (total == 2) && orc_is_trivial(s, m);            /* :99  */
            jb->reward[r] = d ? max_reward : -total;                 /* :100 */
            jb->steps[r] += 1;                                       /* :102 */
            int tr = jb->steps[r] >= jb->horizon;                    /* :103 */
            jb->done[r] = (uint8_t)d;
            jb->truncated[r] = (uint8_t)tr;
            if (jb->status) jb->status[r] = (uint8_t)st;
            if (jb->next_idx && (d || tr)) {                         /* :113-129 */
                memcpy(s, jb->pool + (int64_t)jb->next_idx[r] * 2 * m, (size_t)(2 * m));
                jb->steps[r] = 0;
                l[0] = word_len(s, m);
                l[1] = word_len(s + m, m);
            }
            if (jb->lens8) {
                jb->lens8[2 * r] = (uint8_t)l[0];
                jb->lens8[2 * r + 1] = (uint8_t)l[1];
            }
        }
    }
}

static void *thread_main(void *p) { run_rows((const job_t *)p); return NULL; }

static void parallel_rows(job_t *proto, int64_t n, int threads) {
    if (threads < 1) threads = 1;
    if (threads > 256) threads = 256;
    if (threads == 1 || n < 2 * threads) {
        proto->lo = 0; proto->hi = n;
        run_rows(proto);
        return;
    }
    pthread_t tid[256];
    job_t jobs[256];
    for (int t = 0; t < threads; ++t) {
        jobs[t] = *proto;
        jobs[t].lo = n * t / threads;
        jobs[t].hi = n * (t + 1) / threads;
        pthread_create(&tid[t], NULL, thread_main, &jobs[t]);
    }
    for (int t = 0; t < threads; ++t) pthread_join(tid[t], NULL);
}

void orc_acmove_batch(const int8_t *in, const uint8_t *moves, int8_t *out,
                      int32_t *lens, uint8_t *status, int64_t n, int m,
                      int cyclical, int threads) {
    job_t jb; memset(&jb, 0, sizeof jb);
    jb.kind = 0; jb.in = in; jb.moves = moves; jb.out = out; jb.lens = lens;
    jb.status = status; jb.m = m; jb.cyclical = cyclical;
    parallel_rows(&jb, n, threads);
}

void orc_expand12_batch(const int8_t *in, int8_t *out, int32_t *lens,
                        uint8_t *status, int64_t n, int m, int cyclical, int threads) {
    job_t jb; memset(&jb, 0, sizeof jb);
    jb.kind = 1; jb.in = in; jb.out = out; jb.lens = lens;
    jb.status = status; jb.m = m; jb.cyclical = cyclical;
    parallel_rows(&jb, n, threads);
}

void orc_env_step_batch(int8_t *state, int32_t *steps, const uint8_t *actions,
                        const int8_t *pool, const int32_t *next_idx,
                        int32_t *reward, uint8_t *done, uint8_t *truncated,
                        uint8_t *lens_out, uint8_t *status,
                        int64_t n, int m, int horizon, int threads) {
    job_t jb; memset(&jb, 0, sizeof jb);
    jb.kind = 2; jb.state = state; jb.steps = steps; jb.moves = actions;
    jb.pool = pool; jb.next_idx = next_idx; jb.reward = reward; jb.done = done;
    jb.truncated = truncated; jb.lens8 = lens_out; jb.status = status;
    jb.m = m; jb.horizon = horizon;
    parallel_rows(&jb, n, threads);
}
