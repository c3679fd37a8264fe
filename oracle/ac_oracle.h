/*
 * oracle/ac_oracle.h -- CPU restatement of AC-Solver's transition function.
 *
 * THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, bench.py's
 * cpu_baseline / --impl reference legs and __graft_entry__.smoke() may load it,
 * and only as the checker.  The product (libacx.so) never links or calls it.
 *
 * Every function states which reference lines (relative to /root/reference)
 * it follows.  Parity of this restatement is PINNED: oracle/validate_vs_reference.py
 * runs it against the reference Python (imported in the build container) and
 * tests/golden/ holds vectors produced by the reference itself.
 *
 * Row format (ac_solver/envs/utils.py:4-7): a presentation is 2*m int8,
 * [r_0 | r_1], letters x=1 x^-1=-1 y=2 y^-1=-2, zeros padded on the right of
 * each half.  The word arithmetic is symbol-agnostic (any non-zero int8).
 */
#ifndef AC_ORACLE_H
#define AC_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* per-row status, shared with the product's row_status convention */
enum {
    ORC_OK = 0,
    ORC_EMPTY_RELATOR = 1, /* reference raises AssertionError (utils.py:264-266) */
    ORC_INVALID_INPUT = 2  /* input is not a valid presentation (utils.py:13-54) */
};

int orc_is_valid(const int8_t *p, int m);
int orc_is_trivial(const int8_t *p, int m);

/* simplify_relator (utils.py:178-243): in place on m slots; returns the length */
int orc_simplify_relator(int8_t *r, int m, int cyclical);
/* simplify_presentation (utils.py:246-283) */
int orc_simplify_presentation(int8_t *p, int m, int cyclical, int *lens);
/* concatenate_relators (ac_moves.py:4-76): r_i <- r_i r_{1-i}^{sign}; in place */
void orc_concatenate(int8_t *p, int m, int i, int sign);
/* conjugate (ac_moves.py:79-156): r_i <- g r_i g^-1, g = sign*j; in place */
void orc_conjugate(int8_t *p, int m, int i, int j, int sign);
/* ACMove (ac_moves.py:159-231) */
int orc_acmove(int move_id, const int8_t *in, int8_t *out, int m, int cyclical, int *lens);

/* lockstep batches (threads > 1 uses OpenMP over rows) */
void orc_acmove_batch(const int8_t *in, const uint8_t *moves, int8_t *out,
                      int32_t *lens, uint8_t *status, int64_t n, int m,
                      int cyclical, int threads);
/* all 12 successors of each parent: out[n][12][2m], lens[n][12][2], status[n][12] */
void orc_expand12_batch(const int8_t *in, int8_t *out, int32_t *lens,
                        uint8_t *status, int64_t n, int m, int cyclical, int threads);
/* ACEnv.step (ac_env.py:91-111) + reset-on-done/truncated from a pool
 * (ac_env.py:113-129; agents/training.py:351-352).  state/steps updated in place.
 * next_idx may be NULL (no auto-reset). reward is the raw integer reward. */
void orc_env_step_batch(int8_t *state, int32_t *steps, const uint8_t *actions,
                        const int8_t *pool, const int32_t *next_idx,
                        int32_t *reward, uint8_t *done, uint8_t *truncated,
                        uint8_t *lens_out, uint8_t *status,
                        int64_t n, int m, int horizon, int threads);

/* greedy_search (search/greedy.py:15-121) and bfs (search/breadth_first.py:15-97).
 * path_actions/path_lens receive up to path_cap entries incl. the (-1,len0) head.
 * returns: 1 solved, 0 not solved, <0 error.  *path_n = entries written
 * (bfs failure: 0, the reference returns None).  *n_seen = len(tree_nodes) at exit. */
int orc_greedy(const int8_t *root, int m, int64_t budget, int cyclical,
               int32_t *path_actions, int32_t *path_lens, int path_cap,
               int *path_n, int64_t *n_seen, int64_t *n_expanded);
int orc_bfs(const int8_t *root, int m, int64_t budget, int cyclical,
            int32_t *path_actions, int32_t *path_lens, int path_cap,
            int *path_n, int64_t *n_seen, int64_t *n_expanded);

#ifdef __cplusplus
}
#endif
#endif
