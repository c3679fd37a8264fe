/*
 * include/acx.h -- C ABI of libacx.so, the sm_100a Andrews-Curtis move engine.
 *
 * The reference (Avi161/AC-Solver-Caltech) has no FFI: its boundary for this path is a set of
 * Python call signatures.  Each entry point below names the reference function(s) whose work it
 * does for a whole batch in one launch; ac-solver-caltech_b200/ac_solver/ re-exports the reference
 * names on top of these.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer owned by the caller unless the name ends in `_host`;
 *   - a presentation row is 2*m int8: [r_0 | r_1], letters 1,-1,2,-2, zero padded on the right
 *     of each half (ac_solver/envs/utils.py:4-7);  1 <= m <= 128;
 *   - calls are asynchronous on `stream` (a cudaStream_t passed as void*); nothing is allocated
 *     or freed across the boundary except opaque handles;
 *   - return value: 0 ok, < 0 error (acx_last_error() has the text, thread-local);
 *   - per-row anomalies never fail the call: they go to row_status[] (ACX_ROW_*), and the row is
 *     handed back unchanged.  The Python wrappers raise AssertionError for them exactly where the
 *     reference does (utils.py:264-266).
 *   - there is no CPU implementation behind any of these: without a CUDA device they fail.
 */
#ifndef ACX_H
#define ACX_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* acx_stream_t;

enum {
    ACX_OK = 0,
    ACX_ERR_ARG = -1,  /* bad argument (null pointer, m out of range, ...) */
    ACX_ERR_CUDA = -2, /* a CUDA runtime call failed */
    ACX_ERR_STATE = -3 /* handle in the wrong state */
};

enum {
    ACX_ROW_STATUS_OK = 0,
    ACX_ROW_STATUS_EMPTY_RELATOR = 1, /* reference raises AssertionError (utils.py:264-266) */
    ACX_ROW_STATUS_INVALID_INPUT = 2  /* input is not a valid presentation (utils.py:13-54) */
};

#define ACX_MAX_RELATOR_LENGTH 128

const char* acx_last_error(void);
int acx_version(void);
/* number of CUDA devices visible; < 0 on error */
int acx_device_count(void);
/* make `dev` the current device of the calling thread for this library's runtime */
int acx_set_device(int dev);

/* ACMove in lockstep -- ac_solver/envs/ac_moves.py:159-231 (decode :188-206, concatenate_relators
 * :4-76, conjugate :79-156) followed by simplify_presentation, ac_solver/envs/utils.py:246-283.
 * states_out may alias states_in.  lens_out [n,2] and row_status [n] may be NULL. */
int acx_move_batch(const int8_t* states_in, const uint8_t* moves, int8_t* states_out, uint8_t* lens_out,
                   uint8_t* row_status, int64_t n, int m, int cyclical, acx_stream_t stream);

/* The word-arithmetic primitives the reference exposes next to ACMove, for its scalar API and tests
 * (symbol-agnostic byte path, any non-zero int8 letters, 1 <= m <= 4096):
 *   mode = ACX_MODE_NO_MOVE [| ACX_MODE_CYCLICAL]   simplify_presentation (utils.py:246-283)
 *   mode = ACX_MODE_RAW_MOVE                        concatenate_relators / conjugate only
 *                                                   (ac_moves.py:4-156; moves[] = the id of the
 *                                                   ACMove that uses that primitive), no reduction
 *   mode = 0 | ACX_MODE_CYCLICAL                    full ACMove, byte path
 * states_out may alias states_in. */
#define ACX_MODE_CYCLICAL 1
#define ACX_MODE_NO_MOVE 2
#define ACX_MODE_RAW_MOVE 4
int acx_rows_primitive(const int8_t* states_in, const uint8_t* moves, int8_t* states_out, uint8_t* lens_out,
                       uint8_t* row_status, int64_t n, int m, int mode, acx_stream_t stream);

/* ACEnv.step in lockstep, in place -- ac_solver/envs/ac_env.py:91-111 (cyclical=True move,
 * done = total length 2 and trivial, reward = max_reward*done - total*(1-done) with
 * max_reward = horizon*m*2 (:76), count_steps += 1, truncated = count_steps >= horizon), fused
 * with the reset the PPO loop applies to finished envs (ACEnv.reset, ac_env.py:113-129;
 * agents/training.py:351-352): when pool != NULL and done|truncated, state <- pool[next_idx[e]],
 * count_steps <- 0.
 *
 * Persistent state of env e: state[e] (2m int8, the reference's own row) and ONE 32-bit word
 *     meta[e] = count_steps (bits 0..15) | len(r_0) << 16 | len(r_1) << 24        (horizon <= 65535)
 * (ACEnv.count_steps / ACEnv.lengths, ac_env.py:84-88).  acx_rows_certify / acx_env_meta_init fill it.
 * step_out [n] is one 8-byte record per env (16-byte aligned array), written every step:
 *     int32 reward | uint8 done | uint8 truncated | uint8 len(r_0) | uint8 len(r_1)
 * (lengths are those of the state left in `state`, i.e. after a reset if one happened).
 * Optional outputs (NULL = off):
 *   reward_clipped float [n]   clip(reward, clip_lo, clip_hi): TransformReward, agents/environment.py:48-52
 *   row_status uint8 [n]       ACX_ROW_STATUS_* of every row, written every step
 *   err_count int32 [1]        += number of rows with a non-zero status (the reference raises for them,
 *                              utils.py:264-266; here the row keeps its state) -- one word to poll
 *   action_ring uint8 [ring_len, n]  row (t % ring_len) <- actions: the device-side action log
 *                              (ACEnv.actions, ac_env.py:92); with ring_len >= horizon the actions of a
 *                              finished episode of length L are ring[(t-L+1 .. t) % ring_len, e]
 *   fin_*                      finished-episode side buffer: for every env that is done|truncated in this
 *                              step, k = (*fin_count)++ and, if k < fin_cap, fin_env[k] = e, fin_steps[k] =
 *                              its episode length, fin_state[k] = its final state BEFORE the reset
 *                              (gymnasium's final_observation / final_info, agents/training.py:273-280)
 * flags: ACX_ENV_CERTIFIED promises that every row of `state` and of `pool` is certified
 * (acx_rows_certify said so, or the row is the output of an earlier acx_env_step on certified rows)
 * and that meta holds the current lengths; the kernel then skips the per-row input checks (m = 36:
 * kernel k_env_tiles, every array through bulk copies; arrays 16-byte aligned).  Results are identical
 * either way.
 * chain (optional, ACX_ENV_CHAIN_SLOTS zeroed uint32, one set per env batch): every launch of the certified
 * kernel keeps a parity counter per CTA there.  ACX_ENV_CHAINED additionally promises that every input of this
 * step other than state/meta (actions, pool, next_idx) was complete before the PREVIOUS acx_env_step on these
 * buffers was launched and that nothing else touched state/meta in between; CTA b then waits for CTA b of the
 * previous launch only (the exact row dependency) instead of the whole grid, so consecutive steps overlap
 * head-to-tail.  Without the flag the step is ordered after everything earlier in the stream, as usual. */
#define ACX_ENV_CERTIFIED 1
#define ACX_ENV_CHAINED 2
#define ACX_ENV_CHAIN_SLOTS 4096
typedef struct acx_env_io {
    int8_t* state;           /* [n, 2m] current states (rewritten in place when state_out is NULL) */
    int8_t* state_out;       /* [n, 2m] or NULL: write the next states here instead and leave `state` untouched
                                (e.g. row t+1 of a rollout buffer whose row t is `state`: no observation copy) */
    uint32_t* meta;          /* [n] in place */
    const uint8_t* actions;  /* [n] move ids 0..11 */
    uint8_t* step_out;       /* [n, 8] */
    float* reward_clipped;   /* [n] or NULL */
    uint8_t* row_status;     /* [n] or NULL */
    int32_t* err_count;      /* [1] or NULL */
    const int8_t* pool;      /* [P, 2m] or NULL */
    const int32_t* next_idx; /* [n] or NULL (goes with pool) */
    uint8_t* action_ring;    /* [ring_len, n] or NULL */
    int64_t t;               /* index of this step (selects the ring row) */
    int32_t ring_len;
    int32_t fin_cap;
    int32_t* fin_count;      /* [1] or NULL */
    int32_t* fin_env;        /* [fin_cap] */
    int32_t* fin_steps;      /* [fin_cap] */
    int8_t* fin_state;       /* [fin_cap, 2m] */
    uint32_t* chain;         /* [ACX_ENV_CHAIN_SLOTS] or NULL */
    int64_t n;
    int32_t m, horizon;
    float clip_lo, clip_hi;
    int32_t flags;
} acx_env_io;
int acx_env_step(const acx_env_io* io /* host struct */, acx_stream_t stream);
/* meta[e] = steps[e] (NULL = 0) | lengths of state[e] (non-zero counts per half): after writing states by hand */
int acx_env_meta_init(const int8_t* state, const int32_t* steps, uint32_t* meta, int64_t n, int m, acx_stream_t stream);

/* Row certification for the env fast path: certified_out[r] = 1 iff row r has letters in {0,+-1,+-2},
 * zeros only as right padding, and both relators freely and cyclically reduced (every output of
 * ACMove(cyclical=True) is).  lens_out [n,2] = relator lengths.  Either output may be NULL. */
int acx_rows_certify(const int8_t* rows, int64_t n, int m, uint8_t* lens_out, uint8_t* certified_out,
                     acx_stream_t stream);

/* All 12 successors of every parent in one launch -- the `for action in range(12)` loops of
 * ac_solver/search/greedy.py:76-83 and ac_solver/search/breadth_first.py:69-76.
 * children [f,12,2m], child_lens [f,12,2], child_status [f,12] (the last two may be NULL). */
int acx_expand12(const int8_t* parents, int8_t* children, uint8_t* child_lens, uint8_t* child_status,
                 int64_t f, int m, int cyclical, acx_stream_t stream);

/* greedy_search / bfs for a batch of independent instances (one root each, same m) in ONE launch --
 * ac_solver/search/greedy.py:15-121 (greedy = 1) and ac_solver/search/breadth_first.py:15-97
 * (greedy = 0).  budget = max_nodes_to_explore, cyclical = cyclically_reduce_after_moves.
 * workspace: device scratch of at least acx_search_workspace_bytes(n_instances, m, budget) bytes.
 * result [n_instances, 8] int32: solved, len(tree_nodes) at exit, nodes expanded, path length
 *   (0 = no path: bfs failure returns None; < 0 = path_cap too small, |value| needed), error (1 = a
 *   relator became empty: the reference raises AssertionError), last popped node, last action, last
 *   child length.
 * path [n_instances, path_cap, 2] int32: (action, total length) pairs starting with (-1, len(root)),
 *   exactly the list the reference returns (for a failed greedy search: its last expansion). */
int64_t acx_search_workspace_bytes(int n_instances, int m, int64_t budget);
int acx_search_batch(const int8_t* roots, int n_instances, int m, int64_t budget, int cyclical, int greedy,
                     void* workspace, int64_t workspace_bytes, int32_t* result, int32_t* path, int path_cap,
                     acx_stream_t stream);

/* Level-synchronous bfs building blocks (ac_solver/search/breadth_first.py:15-97 for frontiers too large
 * for one warp; driver: acx/bfs.py, sharded driver: acx/bfs_sharded.py).  m must be even.
 * The visited table is an open-addressing array of tcap (power of two) 64-bit slots, zero = empty.
 *   acx_bfs_seed    puts the root (node `node_id`) into the table.
 *   acx_bfs_probe   for the n candidate rows of one round (children of the frontier in discovery order,
 *                   candidate j = parent_rank*12 + action, or order[j] when given): slot_out[j] = its slot,
 *                   -1 = already visited, -2 = not inserted (total length 2 -> *min_success = min order;
 *                   non-zero status -> *min_error = min order); is_new_out[j] = 1 iff j is the first
 *                   candidate (smallest order) carrying a state that was not visited before;
 *                   hash_out (optional) = the 64-bit row hash (the owner key of the sharded bfs).
 *   acx_bfs_commit  turns the new candidates j < limit into nodes node_base + new_rank[j] (new_rank =
 *                   exclusive scan of is_new), recording parent id = parent_base + order/12, action =
 *                   order % 12 (order[j] = j when order is NULL) and total length.
 *   acx_rows_hash   the 64-bit state hash the table uses; the sharded bfs partitions states by it. */
int acx_bfs_seed(const int8_t* row, int m, uint64_t* table, int64_t tcap, int64_t node_id, acx_stream_t stream);
int acx_bfs_probe(const int8_t* cand, const uint8_t* lens, const uint8_t* status, const int64_t* order, int64_t n,
                  int m, uint64_t* table, int64_t tcap, const int8_t* nodes, int64_t* slot_out,
                  uint64_t* min_success, uint64_t* min_error, uint64_t* hash_out, uint8_t* is_new_out,
                  acx_stream_t stream);
int acx_bfs_commit(const int8_t* cand, const uint8_t* lens, const int64_t* slot, const uint8_t* is_new,
                   const int64_t* new_rank, const int64_t* order, int64_t parent_base, int64_t n, int64_t limit,
                   int m, uint64_t* table, int8_t* nodes, int64_t* parent, uint8_t* action, uint16_t* total,
                   int64_t node_base, acx_stream_t stream);
int acx_rows_hash(const int8_t* rows, int64_t n, int m, uint64_t* hash_out, acx_stream_t stream);

/* Hash-partitioned (multi-GPU) bfs: candidates travel to their owner rank as packed records
 *     row (2m bytes, padded to 8) | int64 order | uint8 len0 | uint8 len1 | uint8 status | pad to 16   (acx_bfs_record_bytes)
 * with owner(state) = ((hash >> 20) & 0x7fffffff) % world.
 *   acx_bfs_bucket        phase 0: owner of each candidate + per-owner counts; phase 1: scatter the records into
 *                         `send`, grouped by owner at offsets[] (exclusive scan of counts), ready for the all-to-all-v.
 *                         order of candidate j = parent_ord[j / 12] * 12 + j % 12.
 *   acx_bfs_probe_packed  acx_bfs_probe on received records; new candidates OR their bit into mask[order / 12].
 *   acx_bfs_commit_packed new candidates become local nodes; global id = gid_base + excl_par[p] +
 *                         popcount(gmask[p] & bits below the action), from the all-reduced masks.
 *   acx_bfs_expand_bucket the fused frontier step for max_relator_length 36: parents nodes[loc[k]] (k < f, global
 *                         discovery ids gids[k]) -> 12 children each -> packed records written straight to
 *                         dst_regions[owner] (host array of `world` device pointers, each a region of region_cap
 *                         records: a local send buffer, or this rank's region inside the owner's peer receive
 *                         buffer); cursor[o] counts region o, cursor[world] != 0 flags a region overflow.
 *                         order of child (k, a) = (gids[k] - ord_base) * 12 + a.  Children equal to their parent
 *                         are dropped (already visited).
 *   probe/commit_packed take the records as n_regions regions of region_cap records, region s holding
 *                         region_counts[s] (host array) records from its start; record index = s * region_cap + i.
 */
int acx_bfs_record_bytes(int m);
int acx_bfs_expand_bucket(const int8_t* nodes, const int64_t* loc, const int64_t* gids, int64_t ord_base, int64_t f, int m,
                          int cyclical, int world, int64_t region_cap, uint64_t* cursor, uint8_t* const* dst_regions,
                          acx_stream_t stream);
int acx_bfs_bucket(const int8_t* cand, const uint8_t* lens, const uint8_t* status, const int64_t* parent_ord, int64_t n,
                   int m, int world, uint8_t* owner_scratch, uint64_t* counts, int64_t* offsets, uint64_t* cursor,
                   uint8_t* send, int phase, acx_stream_t stream);
int acx_bfs_probe_packed(const uint8_t* recs, const int64_t* region_counts, int n_regions, int64_t region_cap, int m,
                         uint64_t* table, int64_t tcap, const int8_t* nodes, int64_t* slot, uint64_t* mins, int32_t* mask,
                         acx_stream_t stream);
int acx_bfs_commit_packed(const uint8_t* recs, const int64_t* region_counts, int n_regions, int64_t region_cap, int m,
                          const int64_t* slot, const int64_t* excl_par, const int32_t* gmask, int64_t parent_base,
                          int64_t gid_base, uint64_t* n_local, uint64_t* table, int8_t* nodes, int64_t* gid, int64_t* parent,
                          uint8_t* action, uint16_t* total, acx_stream_t stream);

/* Bit-plane bfs (max_relator_length <= 36, presentations over {x, y} with a freely reduced root -- cyclically
 * reduced too when cyclical): a node is four 64-bit words
 *     w0 = S(r0) | len(r0) << 40    w1 = T(r0) | len(r1) << 40    w2 = S(r1)    w3 = T(r1)
 * (S: letter is an inverse, T: letter is y; bit k = letter k), 32 bytes instead of 2m; a candidate record is the
 * same 32 bytes with its order (48 bits) and row status (2 bits) in the spare top bits of w2 / w3.  Same search
 * semantics and the same region / order / mask protocol as acx_bfs_expand_bucket, acx_bfs_probe_packed and
 * acx_bfs_commit_packed (breadth_first.py:52-95); `total` of a node is len(r0) + len(r1).
 *   acx_pbfs_pack    byte rows -> nodes; ok_out[j] = 0 when row j is outside the plane domain (use the byte kernels)
 *   acx_pbfs_unpack  nodes -> byte rows */
int acx_pbfs_pack(const int8_t* rows, int64_t n, int m, int cyclical, uint64_t* nodes_out, uint8_t* ok_out,
                  acx_stream_t stream);
int acx_pbfs_unpack(const uint64_t* nodes, int64_t n, int m, int8_t* rows_out, acx_stream_t stream);
int acx_pbfs_hash(const uint64_t* nodes, int64_t n, uint64_t* hash_out, acx_stream_t stream);
int acx_pbfs_seed(const uint64_t* node, uint64_t* table, int64_t tcap, int64_t node_id, acx_stream_t stream);
int acx_pbfs_expand(const uint64_t* nodes, const int64_t* loc, const int64_t* gids, int64_t ord_base, int64_t f, int m,
                    int cyclical, int world, int64_t region_cap, uint64_t* cursor, uint8_t* const* dst_regions,
                    int ctas_per_sm_limit /* 0 = fill the GPU; k = leave room for a kernel on another stream */,
                    acx_stream_t stream);
int acx_pbfs_probe(const uint64_t* recs, const int64_t* region_counts, int n_regions, int64_t region_cap, uint64_t* table,
                   int64_t tcap, const uint64_t* nodes, int64_t* slot, uint64_t* mins, int32_t* mask, acx_stream_t stream);
int acx_pbfs_commit(const uint64_t* recs, const int64_t* region_counts, int n_regions, int64_t region_cap,
                    const int64_t* slot, const int64_t* excl_par, const int32_t* gmask, int64_t parent_base,
                    int64_t gid_base, uint64_t* n_local, uint64_t* table, uint64_t* nodes, int64_t* gid, int64_t* parent,
                    uint8_t* action, acx_stream_t stream);

/* The device-driven round of the plane bfs (csrc/acx_pbfs2.cu; driver acx/pbfs.py): same records, table and search
 * semantics as above, but nothing comes back to the host inside a level.  `ctl` is a control block of int64 words in
 * device memory, identical on every rank: [0] stop code (0 none, 1 solved, 2 error, 3 budget), [1] len(tree_nodes),
 * [2] nodes expanded, [3] first global id of the round being committed, [4] order of the deciding candidate
 * (parent * 12 + action inside the round), [5] global id of the round's first parent, [6] a region overflowed, [7] rounds,
 * [8] records this rank staged, [9] of which for itself (at least 10 words); every kernel returns at once when ctl[0] != 0.
 * counts = the all-gathered cursors, int64 [world, world + 1]: counts[s][o] records rank s staged for owner o, counts[s][world]
 * its overflow flag; the regions a rank received hold region_counts_dev[s * count_stride] records (= counts[s][rank]).
 *   acx_pbfs_expand2  as acx_pbfs_expand, plus: the records of a CTA are sorted by owner in shared memory and written as
 *                     contiguous runs padded to 4 records (128 bytes) with null records (region_cap % 4 == 0, regions
 *                     128-byte aligned); with cyclical == 0 the child that undoes the move `action[parent]` (the
 *                     grandparent; 255 = no such move) is not sent.  cursor[o] counts records incl. padding.
 *   acx_pbfs_probe3   region counts are a DEVICE array.  A claim is final: the candidate first reserves a place in the
 *                     node store (n_local, one atomic per warp), then claims the empty table slot with tag | place and
 *                     writes its state to nodes[place] -- the slot is never written again.  Equal candidates of one round
 *                     settle the reference's first discovery beside the node: gid[place] (caller: -1 beyond n_local) holds
 *                     2^62 | the smallest order until the commit, and mask32 (uint32 view of one uint16 per parent, caller
 *                     zeroes) ends with bit a of parent p set iff child (p, a) is the first discovery of a new state.
 *                     round_start_dev = n_local before the call (a place at or beyond it belongs to this round); a place
 *                     of an earlier round is a visited state whether or not its commit has run, so this probe may run
 *                     next to the previous round's commit.  mins = {min success order, min error order} (caller sets
 *                     both to 2^62).  A reserved place whose claim lost its race is reused by the same thread; what is
 *                     left over stays a hole (gid -1, zero node): acx_pbfs_expand2 and acx_bfs_window_partition skip it.
 *                     A full node store sets ctl[6].
 *   acx_pbfs_decide   exclusive scan of popcount(mask16) -> excl_out (int32), then the decision of the round
 *                     (breadth_first.py:84-95 in the reference's order) into ctl.  temp == NULL returns the scratch size.
 *   acx_pbfs_commit3  the places [round_start, round_end) get global id ctl[3] + excl[p] + popcount(mask16[p] & bits below a),
 *                     parent and action from their winning order (p, a); max_new bounds round_end - round_start.
 *                     round_log (may be NULL): four words {round_start, round_end, first global id, new nodes} of the
 *                     round, from which the driver finds the places that hold a window of the next level. */
int acx_pbfs_expand2(const uint64_t* nodes, const uint8_t* action, const int64_t* loc, const int64_t* gids, int64_t ord_base,
                     int64_t f, int m, int cyclical, int world, int64_t region_cap, uint64_t* cursor,
                     uint8_t* const* dst_regions, const int64_t* ctl, acx_stream_t stream);
int acx_pbfs_probe3(const uint64_t* recs, const int64_t* region_counts_dev, int64_t count_stride, int n_regions, int64_t region_cap,
                    uint64_t* table, int64_t tcap, uint64_t* nodes, int64_t node_cap, int64_t* gid, uint64_t* n_local,
                    const int64_t* round_start_dev, uint64_t* mins, uint32_t* mask32, int64_t* ctl, acx_stream_t stream);
int64_t acx_pbfs_decide(const uint16_t* mask16, int32_t* excl_out, int64_t n_parents, const uint64_t* mins, int64_t parent_base,
                        int64_t budget, int64_t* ctl, const int64_t* counts, int world, int rank, void* temp, int64_t temp_bytes,
                        acx_stream_t stream);
int acx_pbfs_commit3(const int64_t* round_start_dev, const int64_t* round_end_dev, int64_t max_new, int64_t node_cap,
                     const int32_t* excl, const uint32_t* mask32, int64_t parent_base, const int64_t* ctl, int64_t* gid,
                     int64_t* parent, uint8_t* action, int64_t* round_log, acx_stream_t stream);

/* Exclusive scan of popcount(mask[p]) over the parents of a round: excl_out[p] = new nodes discovered by the
 * parents before p (reference order), total_out[0] = new nodes of the round.  temp == NULL returns the scratch
 * size in bytes; otherwise 0 / a negative error. */
int64_t acx_bfs_mask_scan(const int32_t* mask, int64_t n, int64_t* excl_out, int64_t* total_out, void* temp,
                          int64_t temp_bytes, acx_stream_t stream);

/* acx_bfs_mask_scan without any allocation and without caller scratch.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  Same inputs, outputs and results as acx_bfs_mask_scan; the scan's scratch is a
 * static device array of ACX_MASK_SCAN_STATIC_BYTES that every GPU which loads the library holds until exit.
 * n > ACX_MASK_SCAN_STATIC_MAX_N is rejected (ACX_ERR_ARG, message in acx_last_error()).  The entry point calls no
 * cudaMalloc / cudaFree / allocating routine.  All calls on a device share the one array, which keeps the last call's
 * contents (nothing has to start at zero: the scan re-initialises what it reads): two calls that overlap on the GPU,
 * e.g. on different streams, RACE -- order them on one stream or with events. */
#define ACX_MASK_SCAN_STATIC_MAX_N (1 << 22)
#define ACX_MASK_SCAN_STATIC_BYTES (256 * 1024)
int acx_bfs_mask_scan_static(const int32_t* mask, int64_t n, int64_t* excl_out, int64_t* total_out, acx_stream_t stream);

/* Group the local frontier of a level by round (window w = (gid - lo) / chunk of the global discovery ids; the
 * order inside a window is free): phase 0 counts[w] += 1; phase 1 scatters loc_out / gid_out to
 * offsets[w] + (running cursor[w]), offsets = exclusive scan of counts.  loc_out = first + index into gid. */
int acx_bfs_window_partition(const int64_t* gid, int64_t n, int64_t lo, int64_t chunk, int64_t n_windows, uint64_t* counts,
                             const int64_t* offsets, uint64_t* cursor, int64_t first, int64_t* loc_out, int64_t* gid_out,
                             int phase, acx_stream_t stream);

/* Batched path replay: ACMove along a time axis, every intermediate state emitted (replaces the Python loops of
 * agents/training.py:74-135 replay_final_state / actions_to_path, value_search/replay_path.py:107-160,
 * value_search/data_extraction.py:53-190).  Path p applies actions[p, 0 .. path_len[p]) (0-based move ids,
 * path_len NULL = max_steps for all) to roots[p]:
 *   states_out int8 [n_paths, max_steps, 2m] and lens_out uint8 [n_paths, max_steps, 2] after every step (nullable),
 *   final_out int8 [n_paths, 2m], final_lens uint8 [n_paths, 2], status uint8 [n_paths] (ACX_ROW_STATUS_*),
 *   fail_step int32 [n_paths] = the step that stopped the path (-1: none).  A step that empties a relator -- where
 *   the reference raises AssertionError (utils.py:264-266) -- stops the path with the state before that step. */
int acx_path_replay(const int8_t* roots, const uint8_t* actions, const int32_t* path_len, int64_t n_paths, int max_steps, int m,
                    int cyclical, int8_t* states_out, uint8_t* lens_out, int8_t* final_out, uint8_t* final_lens,
                    int32_t* fail_step, uint8_t* status, acx_stream_t stream);

/* expand_path_with_cyclic_reductions (value_search/value_guided_search.py:581-635) for a batch of compact paths found with
 * cyclical=True: replay with free reduction only and insert the conjugation moves (:563-578) that perform the cyclic
 * reductions.  out_actions uint8 / out_totals uint16 [n_paths, out_cap] receive the expanded (action, total_length) pairs,
 * out_len[p] their number.  status: ACX_ROW_STATUS_*, or 3 = out_cap too small (retry with more room); fail_step as in
 * acx_path_replay. */
int acx_path_expand_cyclic(const int8_t* roots, const uint8_t* actions, const int32_t* path_len, int64_t n_paths, int max_steps,
                           int m, uint8_t* out_actions, uint16_t* out_totals, int out_cap, int32_t* out_len, int32_t* fail_step,
                           uint8_t* status, acx_stream_t stream);

/* The 14 handcrafted value-network features of value_search/feature_extraction.py:11-91, float32 [n, 14], computed
 * where the rows already are (removes the numpy hop of value_guided_search.py:49-63). */
int acx_rows_features(const int8_t* rows, int64_t n, int m, float* features_out, acx_stream_t stream);

/* Peer memory (CUDA IPC): a receive buffer the other ranks of one NVLink domain write into directly, so that
 * acx_bfs_expand_bucket IS the all-to-all (stores over NVLink from inside the kernel, no staging, no collective). */
int acx_peer_alloc(int64_t bytes, void** ptr);
int acx_peer_free(void* ptr);
int acx_peer_export(const void* ptr, uint8_t* handle64);
int acx_peer_open(const uint8_t* handle64, void** ptr);
int acx_peer_close(void* ptr);

/* ---- Graph engine: the reference's native C++ programs ---------------------------------
 * barcode_analysis/simplex_data_generation/{classic,prime}_moves/ac_bfs.cpp:13-97 (BFS from (a, b) over UNORDERED
 * relator pairs of total size <= n with full free reduction; node list, edge list and both filtrations) and
 * barcode_analysis/5_steps_neibourhoods/neibourhoods.cpp:18-54 (size of the radius-r ball around a presentation).
 * A node is 16*W bytes (W = width_words in {2, 4, 8}): sign / type planes of both relators, lengths and a 16-bit
 * instance number in the spare top bytes; a relator holds at most 32*W - 8 letters.  Node ids are the reference's
 * names (discovery order of its FIFO loop, ac_bfs.cpp:44-69). */
typedef struct acx_graph_step {
    int32_t width_words; /* W */
    int32_t classic;     /* 1: the 14 classic moves, 0: the 12 prime moves (AC_UTILS_as_sets.h:298-363) */
    int32_t max_total;   /* drop children with len(r1)+len(r2) > max_total (ac_bfs.cpp:59); <= 0: no limit */
    int32_t want_edges;  /* emit the edge list of the step (ac_bfs.cpp:72-79) */
    int32_t allow_empty; /* 1: a relator may cancel completely (neibourhoods.cpp keeps plain vectors); 0: that sets
                            counts[2] (ac_bfs.cpp's Hash() cannot represent it, AC_UTILS_as_sets.h:118-139) */
    uint32_t* nodes;     /* [capacity, 4W] node store in id order; capacity >= n_nodes + n_parents * n_moves */
    int64_t n_nodes;     /* nodes committed before this step */
    int64_t parent_first, n_parents; /* ids [parent_first, parent_first + n_parents) are expanded */
    uint64_t* table;     /* open-addressing visited set, tcap slots (power of two, >= 2 * capacity in use) */
    int64_t tcap;
    /* scratch per candidate, C = n_parents * n_moves (14 or 12) */
    uint32_t* cand;      /* [C, 4W] */
    int64_t* ref;        /* [C] claimed table slot / committed node id of every candidate */
    uint16_t* csize;     /* [C] */
    uint8_t* flag;       /* [C] */
    int32_t* excl;       /* [C] exclusive counts of the two stream compactions (C < 2^31) */
    void* scan_tmp;      /* acx_graph_scan_bytes(C) bytes */
    int64_t scan_tmp_bytes;
    /* outputs */
    uint16_t* node_size; /* [capacity] total length per node id (zero_filtrations) */
    int64_t* edges;      /* [C, 2] (parent id, child id) of this step in the reference's emission order */
    uint16_t* edge_filt; /* [C] max(size(parent), size(child)) (one_filtrations) */
    int64_t* counts;     /* int64[4]: [0] new nodes, [1] edges of the step, [2] != 0: a relator overflowed the
                            planes or cancelled completely (caller raises); caller zeroes [2] once */
} acx_graph_step_t;

int64_t acx_graph_scan_bytes(int64_t n_candidates);
int acx_graph_step(const acx_graph_step_t* step, acx_stream_t stream);
/* int8 rows [r1 | r2] (non-zero entries of each half, neibourhoods.cpp:74-86) -> sorted nodes (sort_,
 * AC_UTILS_no_hash.cpp:120-136); inst may be NULL (0).  ok_out[j] = 0: letters outside {+-1,+-2}, not freely
 * reduced, empty or too long for the planes. */
int acx_graph_pack(const int8_t* rows, int64_t n, int m, const int32_t* inst, int width_words, uint32_t* nodes_out,
                   uint8_t* ok_out, acx_stream_t stream);
int acx_graph_unpack(const uint32_t* nodes, int64_t n, int width_words, int m, int8_t* rows_out, int32_t* inst_out,
                     acx_stream_t stream);
/* insert node ids [0, n_nodes) into a zeroed table (after growing it, or for the roots) */
int acx_graph_rehash(const uint32_t* nodes, int64_t n_nodes, int width_words, uint64_t* table, int64_t tcap,
                     acx_stream_t stream);
/* counts[i] += number of nodes of instance i (ball sizes, neibourhoods.cpp:30-48) */
int acx_graph_count_instances(const uint32_t* nodes, int64_t n_nodes, int width_words, int64_t* counts, int64_t n_inst,
                              acx_stream_t stream);

/* Enumeration step of generate_miller_schupp_presentations (ac_solver/search/miller_schupp/miller_schupp.py:20-83):
 * for the words [first, first + count) of itertools.product([1, 2, -1, -2], repeat=w_len) (:53) apply the x-exponent
 * filter (:56-57), reduce x^-1 w freely and cyclically (:60-63), drop x^-1 (:66-67).  rel_out = reduced relator, 2 bits
 * per letter (0..3 = 1, 2, -1, -2) | length << 32; key_out = its minimum rotation | length << 32 (the class the
 * `seen` set of :69-75 identifies), ~0 for a word that was filtered out.  1 <= w_len <= 15. */
int acx_ms_enumerate(int w_len, int64_t first, int64_t count, uint64_t* key_out, uint64_t* rel_out, acx_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* ACX_H */
