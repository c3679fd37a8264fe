"""beam_search on the GPU: value_search/value_guided_search.py:417-561.

The reference expands the beam one ACMove at a time in Python, moves every child to numpy features / token ids and
back to torch for scoring.  Here one step is

    gather beam rows -> acx_expand12 (beam x 12 children) -> acx_bfs_probe (success test before the visited test,
    first discovery in (beam rank, action) order) -> budget cut per child (:519-522) -> acx_bfs_commit ->
    acx_rows_features / token ids on the device -> the caller's torch model -> stable top-k

with the visited set, the node store and the scores resident in HBM.  Same signature and return value as the
reference (`solved, path, stats`).  `solution_cache` (a Python dict probed with rotation variants, :457-461,
:500-509) is honoured: the children of a step are looked up on the host in the reference's order, between the
trivial test and the visited test (acx.solution_cache).
"""
import time

import numpy as np
import torch

from . import _lib
from .batch import _prep, _ptr
from .bfs import U64_MAX, _next_pow2
from .solution_cache import _check_cache_with_rotations


def features(rows):
    """compute_features (value_search/feature_extraction.py:11-91) for int8 CUDA rows [N, 2m] -> float32 [N, 14]"""
    n, rowb = rows.shape
    stream = _prep(rows.device)
    out = torch.empty((n, 14), dtype=torch.float32, device=rows.device)
    _lib.check(_lib.lib().acx_rows_features(_ptr(rows.contiguous()), n, rowb // 2, _ptr(out), stream), "acx_rows_features")
    return out


def score_states_mlp(rows, model, feat_mean, feat_std):
    """_score_states_mlp (value_guided_search.py:49-63) without the numpy hop; float32 scores on the device"""
    f = features(rows)
    mean = torch.as_tensor(np.asarray(feat_mean, dtype=np.float32), device=rows.device)
    std = torch.as_tensor(np.asarray(feat_std, dtype=np.float32), device=rows.device)
    with torch.no_grad():
        return torch.expm1(model((f - mean) / std).squeeze(-1))


def score_states_seq(rows, model, max_state_dim):
    """_score_states_seq (value_guided_search.py:66-83): token ids = letter + 2, padding token 2"""
    n, rowb = rows.shape
    tok = torch.full((n, max_state_dim), 2, dtype=torch.long, device=rows.device)
    tok[:, :rowb] = rows.to(torch.long) + 2
    with torch.no_grad():
        return torch.expm1(model(tok).squeeze(-1))


def beam_search(presentation, model, architecture="mlp", feat_mean=None, feat_std=None, beam_width=50,
                max_nodes_to_explore=10000, device="cuda:0", cyclically_reduce_after_moves=False, verbose=False,
                solution_cache=None, time_limit=None):
    """Returns (solved, path, stats): path = [(action, total_length), ...] without a root sentinel (:89-96),
    stats = {'nodes_explored', 'steps'} (:496-497, :560)."""
    root = np.ascontiguousarray(np.asarray(presentation), dtype=np.int8)
    rowb = len(root)
    m = rowb // 2
    if solution_cache is not None:  # :459-465
        cached = _check_cache_with_rotations(tuple(int(v) for v in root), m, solution_cache)
        if cached is not None:
            return True, cached, {"nodes_explored": 1, "steps": 0, "cache_hit": True}
    dev = torch.device(device)
    stream = _prep(dev)
    L = _lib.lib()
    budget = int(max_nodes_to_explore)
    cyc = int(bool(cyclically_reduce_after_moves))
    W = int(beam_width)
    cap = max(budget, 1) + 16
    tcap = _next_pow2(2 * (cap + 12 * W))
    nodes = torch.zeros((cap, rowb), dtype=torch.int8, device=dev)
    parent = torch.full((cap,), -1, dtype=torch.int64, device=dev)
    action = torch.zeros((cap,), dtype=torch.uint8, device=dev)
    total = torch.zeros((cap,), dtype=torch.int16, device=dev)
    table = torch.zeros((tcap,), dtype=torch.int64, device=dev)
    nodes[0].copy_(torch.from_numpy(root))
    total[0] = int(np.count_nonzero(root))
    _lib.check(L.acx_bfs_seed(_ptr(nodes), m, _ptr(table), tcap, 0, stream), "acx_bfs_seed")

    def path_to(node, last_action, last_total=2):
        chain = []
        while node > 0:
            chain.append((int(action[node]), int(total[node])))
            node = int(parent[node])
        chain.reverse()
        return chain + [(int(last_action), int(last_total))]

    beam = torch.zeros((1,), dtype=torch.int64, device=dev)
    total_nodes, step = 1, 0
    t_start = time.time()
    while beam.numel() and total_nodes < budget:
        if time_limit is not None and (time.time() - t_start) > time_limit:
            if verbose:
                print(f"Time limit ({time_limit}s) exceeded")
            break
        step += 1
        F = beam.numel()
        n = 12 * F
        rows = nodes[beam]
        ch = torch.empty((n, rowb), dtype=torch.int8, device=dev)
        ln = torch.empty((n, 2), dtype=torch.uint8, device=dev)
        st = torch.empty((n,), dtype=torch.uint8, device=dev)
        slot = torch.empty((n,), dtype=torch.int64, device=dev)
        is_new = torch.empty((n,), dtype=torch.uint8, device=dev)
        mins = torch.full((2,), -1, dtype=torch.int64, device=dev)
        _lib.check(L.acx_expand12(_ptr(rows), _ptr(ch), _ptr(ln), _ptr(st), F, m, cyc, stream), "acx_expand12")
        _lib.check(L.acx_bfs_probe(_ptr(ch), _ptr(ln), _ptr(st), 0, n, m, _ptr(table), tcap, _ptr(nodes), _ptr(slot),
                                   mins[0:].data_ptr(), mins[1:].data_ptr(), 0, _ptr(is_new), stream), "acx_bfs_probe")
        cum = torch.cumsum(is_new, dim=0, dtype=torch.int64)
        need = budget - total_nodes  # the child that brings total_nodes to the budget ends the expansion (:519-522)
        head = torch.cat([mins, cum[-1:], torch.searchsorted(cum, torch.tensor([need], device=dev))]).cpu().numpy()
        c_s, c_e, n_new = int(head[0]) & U64_MAX, int(head[1]) & U64_MAX, int(head[2])
        c_b = int(head[3]) if n_new >= need else U64_MAX
        first = min(c_s, c_e, c_b)
        if solution_cache is not None:
            # the cache is consulted per child between the trivial test and the visited test (:489-498): a hit at child c
            # counts when no earlier child solved / raised and the budget was not reached before c
            bound = min(n, c_s, c_e, c_b + 1 if c_b != U64_MAX else n)
            rows_h = ch[:bound].cpu().numpy()
            tots_h = ln[:bound].to(torch.int64).sum(dim=1).cpu().numpy()
            for c in range(bound):
                cached = _check_cache_with_rotations(tuple(rows_h[c].tolist()), m, solution_cache)
                if cached is not None:
                    before = int(cum[c - 1]) if c > 0 else 0
                    path = path_to(int(beam[c // 12]), c % 12, int(tots_h[c])) + list(cached)
                    return True, path, {"nodes_explored": total_nodes + before, "steps": step, "cache_hit": True}
        if first != U64_MAX and first == c_e:
            raise AssertionError("a relator became empty during the search: not a valid presentation")
        if first != U64_MAX and first == c_s:
            before = int(cum[c_s - 1]) if c_s > 0 else 0
            # the nodes added before the solving child exist in the reference's tree; they only matter for the count
            stats = {"nodes_explored": total_nodes + before, "steps": step}
            return True, path_to(int(beam[c_s // 12]), c_s % 12), stats
        limit = n if first == U64_MAX else c_b + 1
        excl = cum - is_new
        _lib.check(L.acx_bfs_commit(_ptr(ch), _ptr(ln), _ptr(slot), _ptr(is_new), _ptr(excl), 0, 0, n, limit, m, _ptr(table),
                                    _ptr(nodes), _ptr(parent), _ptr(action), total.data_ptr(), total_nodes, stream), "acx_bfs_commit")
        added = n_new if first == U64_MAX else need
        new_ids = torch.arange(total_nodes, total_nodes + added, device=dev)
        parent[new_ids] = beam[parent[new_ids]]  # commit stored the beam rank; the parent is that beam entry's node
        total_nodes += added
        if added == 0:
            break
        if total_nodes >= budget:
            break  # the reference scores this last batch and then leaves the loop without using it (:421, :526-541)
        child_rows = nodes[new_ids]
        if architecture == "mlp":
            scores = score_states_mlp(child_rows, model, feat_mean, feat_std)
        else:
            scores = score_states_seq(child_rows, model, rowb)
        order = torch.sort(scores.to(torch.float64), stable=True).indices  # ties keep discovery order (:540)
        beam = new_ids[order[:W]]
        if verbose and step % 10 == 0:
            print(f"  Step {step}: beam_size={beam.numel()}, nodes={total_nodes}, best_score={float(scores[order[0]]):.1f}, "
                  f"min_len={int(total[beam].min())}")
    return False, [], {"nodes_explored": total_nodes, "steps": step}


def value_guided_greedy_search(presentation, model=None, architecture="mlp", feat_mean=None, feat_std=None,
                               max_nodes_to_explore=10000, device="cuda:0", use_length_priority=False,
                               cyclically_reduce_after_moves=False, verbose=False, solution_cache=None, time_limit=None):
    """value_search/value_guided_search.py:253-414: best-first search on V(P) (or on the total length).

    The pops are sequential (the heap lives on the host, as in the reference); each pop is one `acx_expand12` of the
    popped row, one `acx_bfs_probe` against the resident visited set, one device-side scoring batch of the children
    and one `acx_bfs_commit`.  Quirks kept: a child is "new" when its state was not in the map BEFORE this pop, so
    two equal children of one pop are both pushed (:375-376, :392-401); the budget test runs after each pop (:403).
    Returns (solved, path, {'nodes_explored', 'min_length'})."""
    import heapq

    root = np.ascontiguousarray(np.asarray(presentation), dtype=np.int8)
    rowb = len(root)
    m = rowb // 2
    dev = torch.device(device)
    stream = _prep(dev)
    L = _lib.lib()
    budget = int(max_nodes_to_explore)
    cyc = int(bool(cyclically_reduce_after_moves))
    cap = 2 * (max(budget, 1) + 16)  # equal children of one pop are stored twice
    tcap = _next_pow2(2 * (cap + 12))
    nodes = torch.zeros((cap, rowb), dtype=torch.int8, device=dev)
    parent_t = torch.full((cap,), -1, dtype=torch.int64, device=dev)  # scratch for acx_bfs_commit (the tree lives on the host)
    action_t = torch.zeros((cap,), dtype=torch.uint8, device=dev)
    total_t = torch.zeros((cap,), dtype=torch.int16, device=dev)
    table = torch.zeros((tcap,), dtype=torch.int64, device=dev)
    nodes[0].copy_(torch.from_numpy(root))
    _lib.check(L.acx_bfs_seed(_ptr(nodes), m, _ptr(table), tcap, 0, stream), "acx_bfs_seed")
    total0 = int(np.count_nonzero(root))
    parent_map = {0: (None, total0, None)}

    def score(rows):
        if architecture == "mlp":
            return score_states_mlp(rows, model, feat_mean, feat_std).cpu().numpy().tolist()
        return score_states_seq(rows, model, rowb).cpu().numpy().tolist()

    def path_to(nid):
        path = []
        while nid is not None:
            a, ln, par = parent_map[nid]
            if par is None:
                break
            path.append((a, ln))
            nid = par
        path.reverse()
        return path

    init_priority = float(total0) if use_length_priority else score(nodes[0:1])[0]
    if solution_cache is not None:  # :325-330
        cached = _check_cache_with_rotations(tuple(int(v) for v in root), m, solution_cache)
        if cached is not None:
            return True, cached, {"nodes_explored": 1, "min_length": 2, "cache_hit": True}
    heap = [(init_priority, 0, 0)]
    next_id, n_states, min_length = 1, 1, total0
    ch = torch.empty((12, rowb), dtype=torch.int8, device=dev)
    ln = torch.empty((12, 2), dtype=torch.uint8, device=dev)
    st = torch.empty((12,), dtype=torch.uint8, device=dev)
    slot = torch.empty((12,), dtype=torch.int64, device=dev)
    is_new = torch.empty((12,), dtype=torch.uint8, device=dev)
    mins = torch.empty((2,), dtype=torch.int64, device=dev)
    t_start = time.time()
    while heap:
        if time_limit is not None and (time.time() - t_start) > time_limit:
            if verbose:
                print(f"Time limit ({time_limit}s) exceeded")
            break
        _, path_length, cur = heapq.heappop(heap)
        mins.fill_(-1)
        _lib.check(L.acx_expand12(nodes[cur].data_ptr(), _ptr(ch), _ptr(ln), _ptr(st), 1, m, cyc, stream), "acx_expand12")
        _lib.check(L.acx_bfs_probe(_ptr(ch), _ptr(ln), _ptr(st), 0, 12, m, _ptr(table), tcap, _ptr(nodes), _ptr(slot),
                                   mins[0:].data_ptr(), mins[1:].data_ptr(), 0, _ptr(is_new), stream), "acx_bfs_probe")
        h = torch.cat([ln.to(torch.int64).sum(dim=1), st.to(torch.int64), slot, is_new.to(torch.int64)]).tolist()
        tot, stat, slots, news = h[0:12], h[12:24], h[24:36], h[36:48]
        ch_host = ch.cpu().numpy() if solution_cache is not None else None
        kids = []
        for a in range(12):
            if stat[a]:
                raise AssertionError("a relator became empty during the search: not a valid presentation")
            if tot[a] < min_length:
                min_length = tot[a]
                if verbose:
                    print(f"New min length: {min_length} (nodes: {n_states})")
            if tot[a] == 2:
                return True, path_to(cur) + [(a, 2)], {"nodes_explored": n_states, "min_length": min_length}
            if solution_cache is not None:  # :365-373
                cached = _check_cache_with_rotations(tuple(ch_host[a].tolist()), m, solution_cache)
                if cached is not None:
                    return True, path_to(cur) + [(a, tot[a])] + list(cached), {"nodes_explored": n_states, "min_length": 2,
                                                                              "cache_hit": True}
            if slots[a] >= 0:
                kids.append(a)
        if kids:
            idx = torch.tensor(kids, device=dev)
            ids = torch.arange(next_id, next_id + len(kids), device=dev)
            nodes[ids] = ch[idx]
            scores = [tot[a] for a in kids] if use_length_priority else score(nodes[next_id: next_id + len(kids)])
            rank = torch.zeros((12,), dtype=torch.int64, device=dev)
            rank[idx] = torch.arange(len(kids), device=dev)
            _lib.check(L.acx_bfs_commit(_ptr(ch), _ptr(ln), _ptr(slot), _ptr(is_new), _ptr(rank), 0, 0, 12, 12, m, _ptr(table),
                                        _ptr(nodes), _ptr(parent_t), _ptr(action_t), total_t.data_ptr(), next_id, stream),
                       "acx_bfs_commit")
            for k, a in enumerate(kids):
                nid = next_id + k
                parent_map[nid] = (a, tot[a], cur)
                heapq.heappush(heap, (scores[k], path_length + 1, nid))
                n_states += news[a]
            next_id += len(kids)
        if n_states >= budget:
            if verbose:
                print(f"Budget exhausted: {n_states} nodes explored")
            break
    return False, [], {"nodes_explored": n_states, "min_length": min_length}
