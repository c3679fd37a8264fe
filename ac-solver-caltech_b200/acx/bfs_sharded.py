"""Hash-partitioned level-synchronous bfs over G GPUs (BASELINE configs[3]).

Every state is owned by rank `owner(hash(state)) = (hash >> 20) % G`; a rank keeps the visited-table
shard, the rows and the (parent, action, total) records of the states it owns.  Per round (a window of
`chunk` frontier parents in GLOBAL discovery order):

    expand local parents (acx_expand12)  ->  hash, bucket by owner into packed records (acx_bfs_bucket)
    ->  all-to-all-v (NCCL over NVLink)  ->  insert into the local table shard with first-discovery
        semantics (acx_bfs_probe_packed: order numbers travel with the rows)  ->  all-reduce of one 12-bit
        "new" mask per parent (SUM == OR: bits are disjoint)  ->  every rank derives the same global
        discovery ids, the same budget cut (breadth_first.py:91-95) and the same early exit (:84-85)
    ->  commit (acx_bfs_commit_packed).

Global node ids therefore equal the reference's FIFO discovery order, and (solved, path, len(tree_nodes))
are identical to the single-GPU engine and to the reference.  Host logic is plain torch + torch.distributed
and is device agnostic: tests/ drive it on CPU with gloo and a stand-in shard backend; the product backend
is GpuShard below (libacx kernels only, no CPU path)."""
import time

import numpy as np
import torch
import torch.distributed as dist

from . import _lib
from .batch import _prep, _ptr
from .bfs import _next_pow2

SENT = 1 << 62


def popcount12(x):
    out = torch.zeros_like(x)
    for a in range(12):
        out += (x >> a) & 1
    return out


def owner_of(h, world):
    """partition key: bits above the table-slot bits of the 64-bit row hash"""
    return ((h >> 20) & 0x7FFFFFFF) % world


def record_bytes(rowb):
    """row | int64 order | len0 | len1 | status | pad  (acx_bfs_record_bytes)"""
    return (rowb + 7) // 8 * 8 + 16


_POP_LUT = {}


def popcount_lut(mask):
    """popcount of 12-bit masks through a 4096-entry table (one gather)"""
    key = str(mask.device)
    if key not in _POP_LUT:
        _POP_LUT[key] = popcount12(torch.arange(4096, dtype=torch.int64, device=mask.device))
    return _POP_LUT[key][mask.to(torch.int64)]


class GpuShard:
    """One rank's shard on its GPU: rows, records and table of the states it owns (libacx kernels)."""

    def __init__(self, m, cyclical, cap_local, max_cand, device):
        self.m, self.rowb, self.cyc = m, 2 * m, bool(cyclical)
        self.rec = record_bytes(self.rowb)
        self.dev = torch.device(device)
        self.stream = _prep(self.dev)
        d = self.dev
        self.cap = int(cap_local)
        self.tcap = _next_pow2(2 * (self.cap + max_cand))
        self.nodes = torch.zeros((self.cap, self.rowb), dtype=torch.int8, device=d)
        self.gid = torch.full((self.cap,), -1, dtype=torch.int64, device=d)
        self.parent = torch.full((self.cap,), -1, dtype=torch.int64, device=d)
        self.action = torch.zeros((self.cap,), dtype=torch.uint8, device=d)
        self.total = torch.zeros((self.cap,), dtype=torch.int16, device=d)
        self.table = torch.zeros((self.tcap,), dtype=torch.int64, device=d)
        self.n_local_dev = torch.zeros((1,), dtype=torch.int64, device=d)
        self.L = _lib.lib()
        assert self.L.acx_bfs_record_bytes(m) == self.rec

    def owner_of_row(self, row, world):
        r = torch.from_numpy(np.ascontiguousarray(row[None, :])).to(self.dev)
        h = torch.empty((1,), dtype=torch.int64, device=self.dev)
        _lib.check(self.L.acx_rows_hash(_ptr(r), 1, self.m, _ptr(h), self.stream), "acx_rows_hash")
        return int(owner_of(h, world)[0])

    def seed(self, root):
        self.nodes[0].copy_(torch.from_numpy(root))
        self.total[0] = int(np.count_nonzero(root))
        self.gid[0] = 0
        _lib.check(self.L.acx_bfs_seed(_ptr(self.nodes), self.m, _ptr(self.table), self.tcap, 0, self.stream), "acx_bfs_seed")
        self.n_local_dev.fill_(1)

    def n_local(self):
        return int(self.n_local_dev.item())

    def bucket(self, rows, parent_ord, world):
        """expand `rows` (12 children each) and group the candidate records by owner rank"""
        f = rows.shape[0]
        n = 12 * f
        d = self.dev
        send = torch.empty((n, self.rec), dtype=torch.uint8, device=d)
        counts = torch.zeros((world,), dtype=torch.int64, device=d)
        if f == 0:
            return send, counts
        ch = torch.empty((n, self.rowb), dtype=torch.int8, device=d)
        ln = torch.empty((n, 2), dtype=torch.uint8, device=d)
        st = torch.empty((n,), dtype=torch.uint8, device=d)
        _lib.check(self.L.acx_expand12(_ptr(rows), _ptr(ch), _ptr(ln), _ptr(st), f, self.m, int(self.cyc), self.stream), "acx_expand12")
        own = torch.empty((n,), dtype=torch.uint8, device=d)
        po = parent_ord.contiguous()
        _lib.check(self.L.acx_bfs_bucket(_ptr(ch), 0, 0, 0, n, self.m, world, _ptr(own), _ptr(counts), 0, 0, 0, 0, self.stream),
                   "acx_bfs_bucket")
        offsets = torch.cumsum(counts, 0) - counts
        cursor = torch.zeros((world,), dtype=torch.int64, device=d)
        _lib.check(self.L.acx_bfs_bucket(_ptr(ch), _ptr(ln), _ptr(st), _ptr(po), n, self.m, world, _ptr(own), _ptr(counts),
                                         _ptr(offsets), _ptr(cursor), _ptr(send), 1, self.stream), "acx_bfs_bucket")
        return send, counts

    def probe(self, recv, n_parents):
        n = recv.shape[0]
        slot = torch.empty((n,), dtype=torch.int64, device=self.dev)
        mask = torch.zeros((n_parents,), dtype=torch.int32, device=self.dev)
        mins = torch.full((2,), -1, dtype=torch.int64, device=self.dev)
        if n:
            _lib.check(self.L.acx_bfs_probe_packed(_ptr(recv), n, self.m, _ptr(self.table), self.tcap, _ptr(self.nodes), _ptr(slot),
                                                   _ptr(mins), _ptr(mask), self.stream), "acx_bfs_probe_packed")
        mins = torch.where(mins < 0, torch.full_like(mins, SENT), mins)
        return slot, mask, mins

    def commit(self, recv, slot, excl_par, gmask, parent_base, gid_base):
        n = recv.shape[0]
        if n:
            _lib.check(self.L.acx_bfs_commit_packed(_ptr(recv), n, self.m, _ptr(slot), _ptr(excl_par), _ptr(gmask), parent_base,
                                                    gid_base, _ptr(self.n_local_dev), _ptr(self.table), _ptr(self.nodes),
                                                    _ptr(self.gid), _ptr(self.parent), _ptr(self.action), self.total.data_ptr(),
                                                    self.stream), "acx_bfs_commit_packed")


class ShardedBFS:
    def __init__(self, root, budget, shard, chunk=1 << 20, group=None):
        self.root = np.ascontiguousarray(np.asarray(root), dtype=np.int8)
        self.budget = max(1, int(budget))
        self.shard = shard
        self.group = group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.chunk = int(chunk)
        self.dev = shard.dev
        self.n_nodes, self.n_expanded, self.levels = 1, 0, []
        self.comm_bytes = 0
        self.timers = None  # set to {} to collect a per-phase wall-clock breakdown (adds a device sync per phase)
        self._t0 = 0.0

    def _tick(self, name):
        if self.timers is None:
            return
        if self.dev.type == "cuda":
            torch.cuda.synchronize(self.dev)
        now = time.perf_counter()
        if name is not None:
            self.timers[name] = self.timers.get(name, 0.0) + now - self._t0
        self._t0 = now

    def _lookup(self, gid):
        """(parent gid, action, total) of node `gid`, from whichever rank owns it"""
        s = self.shard
        hit = torch.nonzero(s.gid[: s.n_local()] == gid).flatten()
        rec = torch.zeros((3,), dtype=torch.int64, device=self.dev)
        if hit.numel():
            k = hit[0]
            rec = torch.stack([s.parent[k] + 1, s.action[k].to(torch.int64), s.total[k].to(torch.int64)])
        dist.all_reduce(rec, group=self.group)
        return int(rec[0]) - 1, int(rec[1]), int(rec[2])

    def _path(self, node, action):
        chain = []
        while node >= 0:
            par, act, tot = self._lookup(node)
            chain.append((act, tot))
            node = par
        chain.reverse()
        return [(-1, chain[0][1])] + chain[1:] + [(int(action), 2)]

    def run(self):
        s, G, dev = self.shard, self.world, self.dev
        if s.owner_of_row(self.root, G) == self.rank:
            s.seed(self.root)
        lo, hi = 0, 1
        first = 0  # local index of the first node this rank discovered in the current frontier level
        while lo < hi:
            self.levels.append(hi - lo)
            self._tick(None)
            n_local = s.n_local()
            gids, perm = torch.sort(s.gid[first:n_local])  # local frontier in global discovery order
            loc = perm + first
            first = n_local
            bounds = torch.arange(lo, hi + self.chunk, self.chunk, device=dev).clamp_(max=hi)
            cuts = torch.searchsorted(gids, bounds).tolist()
            self._tick("level_sort")
            for k, off in enumerate(range(lo, hi, self.chunk)):
                C = min(self.chunk, hi - off)
                a, b = cuts[k], cuts[k + 1]
                rows = s.nodes[loc[a:b]]
                send, counts = s.bucket(rows, gids[a:b] - off, G)
                self._tick("expand_bucket")
                rc = torch.empty_like(counts)
                dist.all_to_all_single(rc, counts, group=self.group)
                rcl, scl = [int(v) for v in rc.tolist()], [int(v) for v in counts.tolist()]
                recv = torch.empty((sum(rcl), send.shape[1]), dtype=torch.uint8, device=dev)
                dist.all_to_all_single(recv, send, output_split_sizes=rcl, input_split_sizes=scl, group=self.group)
                self.comm_bytes += send.shape[1] * (send.shape[0] - scl[self.rank])
                self._tick("all_to_all")
                slot, mask, mins = s.probe(recv, C)
                self._tick("probe")
                dist.all_reduce(mask, group=self.group)
                dist.all_reduce(mins, op=dist.ReduceOp.MIN, group=self.group)
                cnt = popcount_lut(mask)
                cum = torch.cumsum(cnt, dim=0)
                c_s, c_e, n_new = torch.cat([mins, cum[-1:]]).tolist()
                self._tick("reduce_scan")
                events = []
                if c_s < SENT:
                    events.append((c_s, 0, "solved"))
                if c_e < SENT:
                    events.append((c_e, 0, "error"))
                p_star = None
                if self.n_nodes + n_new >= self.budget:
                    # first parent after which len(tree_nodes) >= max_nodes_to_explore (breadth_first.py:91-95)
                    p_star = int(torch.searchsorted(cum, torch.tensor([self.budget - self.n_nodes], device=dev))[0])
                    events.append((12 * p_star + 11, 1, "budget"))
                if events:
                    pos, _, what = min(events)
                    if what == "solved":
                        self.n_expanded += pos // 12 + 1
                        p, aa = pos // 12, pos % 12
                        before = int(cum[p] - cnt[p]) + int(popcount12(mask[p].to(torch.int64) & ((1 << aa) - 1)))
                        self.n_nodes += before
                        return True, self._path(off + p, aa)
                    if what == "error":
                        raise AssertionError("a relator became empty during the search: not a valid presentation")
                    self.n_expanded += p_star + 1
                    self.n_nodes += int(cum[p_star])
                    return False, None
                s.commit(recv, slot, cum - cnt, mask, off, self.n_nodes)
                self._tick("commit")
                self.n_nodes += n_new
                self.n_expanded += C
            lo, hi = hi, self.n_nodes
        return False, None
