"""Level-synchronous bfs on one GPU with the reference's exact discovery order, budget rule and result
(ac_solver/search/breadth_first.py:15-97).  For frontiers of millions of nodes; small searches use the
one-warp-per-instance engine (acx.search)."""
import numpy as np
import torch

from . import _lib
from .batch import _prep, _ptr

U64_MAX = (1 << 64) - 1


def _next_pow2(x):
    p = 64
    while p < x:
        p <<= 1
    return p


class LevelBFS:
    def __init__(self, root, budget, cyclical=False, device="cuda:0", chunk=1 << 20):
        root = np.ascontiguousarray(np.asarray(root), dtype=np.int8)
        self.m = len(root) // 2
        self.rowb = 2 * self.m
        self.budget = max(1, int(budget))
        self.cyc = bool(cyclical)
        self.dev = torch.device(device)
        self.stream = _prep(self.dev)
        self.chunk = int(min(chunk, max(1024, self.budget)))
        self.cap = self.budget + 16
        self.tcap = _next_pow2(2 * (self.cap + 12 * self.chunk))
        d = self.dev
        self.nodes = torch.zeros((self.cap, self.rowb), dtype=torch.int8, device=d)
        self.parent = torch.full((self.cap,), -1, dtype=torch.int64, device=d)
        self.action = torch.zeros((self.cap,), dtype=torch.uint8, device=d)
        self.total = torch.zeros((self.cap,), dtype=torch.int16, device=d)
        self.table = torch.zeros((self.tcap,), dtype=torch.int64, device=d)
        nc = 12 * self.chunk
        self.children = torch.empty((self.chunk, 12, self.rowb), dtype=torch.int8, device=d)
        self.lens = torch.empty((self.chunk, 12, 2), dtype=torch.uint8, device=d)
        self.status = torch.empty((self.chunk, 12), dtype=torch.uint8, device=d)
        self.slot = torch.empty((nc,), dtype=torch.int64, device=d)
        self.is_new = torch.empty((nc,), dtype=torch.uint8, device=d)
        self.mins = torch.empty((2,), dtype=torch.int64, device=d)  # min_success, min_error (as uint64 bit patterns)
        self.nodes[0].copy_(torch.from_numpy(root))
        self.total[0] = int(np.count_nonzero(root))
        L = _lib.lib()
        _lib.check(L.acx_bfs_seed(_ptr(self.nodes), self.m, _ptr(self.table), self.tcap, 0, self.stream), "acx_bfs_seed")
        self.n_nodes = 1
        self.n_expanded = 0
        self.levels = []

    def _path(self, node, action):
        ids = []
        while node >= 0:
            ids.append(node)
            node = int(self.parent[node])
        ids.reverse()
        idx = torch.tensor(ids, device=self.dev)
        acts = self.action[idx].cpu().numpy()
        tots = self.total[idx].cpu().numpy()
        path = [(-1, int(tots[0]))] + [(int(a), int(t)) for a, t in zip(acts[1:], tots[1:])]
        return path + [(int(action), 2)]

    def run(self):
        """Returns (solved, path or None).  self.n_nodes = len(tree_nodes) at exit, self.n_expanded = popped nodes."""
        L = _lib.lib()
        lo, hi = 0, 1
        while lo < hi:
            self.levels.append(hi - lo)
            for off in range(lo, hi, self.chunk):
                F = min(self.chunk, hi - off)
                n = 12 * F
                _lib.check(L.acx_expand12(self.nodes[off].data_ptr(), _ptr(self.children), _ptr(self.lens), _ptr(self.status),
                                          F, self.m, int(self.cyc), self.stream), "acx_expand12")
                self.mins.fill_(-1)
                _lib.check(L.acx_bfs_probe(_ptr(self.children), _ptr(self.lens), _ptr(self.status), 0, n, self.m,
                                           _ptr(self.table), self.tcap, _ptr(self.nodes), _ptr(self.slot),
                                           self.mins[0:].data_ptr(), self.mins[1:].data_ptr(), 0, _ptr(self.is_new),
                                           self.stream), "acx_bfs_probe")
                cum = torch.cumsum(self.is_new[:n], dim=0, dtype=torch.int64)
                seen_after = cum[11::12] + self.n_nodes  # len(tree_nodes) after each parent (breadth_first.py:91)
                over = seen_after >= self.budget
                any_over = over.any()
                pstar = torch.argmax(over.to(torch.uint8))
                head = torch.stack([self.mins[0], self.mins[1], any_over.to(torch.int64), pstar, seen_after[pstar], cum[-1]]).cpu().numpy()
                c_s, c_e = int(head[0]) & U64_MAX, int(head[1]) & U64_MAX
                p_star = int(head[3]) if head[2] else None
                # the first event in the reference's processing order decides
                events = []
                if c_s != U64_MAX:
                    events.append((c_s, 0, "solved"))
                if c_e != U64_MAX:
                    events.append((c_e, 0, "error"))
                if p_star is not None:
                    events.append((12 * p_star + 11, 1, "budget"))
                if events:
                    pos, _, what = min(events)
                    if what == "solved":
                        self.n_expanded += pos // 12 + 1
                        self.n_nodes += int(cum[pos - 1]) if pos > 0 else 0
                        return True, self._path(off + pos // 12, pos % 12)
                    if what == "error":
                        raise AssertionError("a relator became empty during the search: not a valid presentation")
                    self.n_expanded += p_star + 1
                    self.n_nodes = int(head[4])
                    return False, None
                excl = cum - self.is_new[:n]
                _lib.check(L.acx_bfs_commit(_ptr(self.children), _ptr(self.lens), _ptr(self.slot), _ptr(self.is_new),
                                            excl.data_ptr(), 0, off, n, n, self.m, _ptr(self.table), _ptr(self.nodes),
                                            _ptr(self.parent), _ptr(self.action), self.total.data_ptr(), self.n_nodes,
                                            self.stream), "acx_bfs_commit")
                self.n_nodes += int(head[5])
                self.n_expanded += F
            lo, hi = hi, self.n_nodes
        return False, None


def bfs_level(presentation, max_nodes_to_explore=10000, cyclically_reduce_after_moves=False, device="cuda:0",
              chunk=1 << 20, stats=None):
    eng = LevelBFS(presentation, max_nodes_to_explore, cyclically_reduce_after_moves, device, chunk)
    ok, path = eng.run()
    if stats is not None:
        stats.update({"n_seen": eng.n_nodes, "n_expanded": eng.n_expanded, "levels": eng.levels})
    return ok, path
