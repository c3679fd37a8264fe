"""Hash-partitioned level-synchronous bfs over G GPUs (BASELINE configs[3]).

Every state is owned by rank `owner(hash(state)) = (hash >> 20) % G`; a rank keeps the visited-table
shard, the rows and the (parent, action, total) records of the states it owns.  Per round (a window of
`chunk` frontier parents in GLOBAL discovery order):

    expand local parents (acx_expand12)  ->  hash, bucket by owner  ->  all-to-all-v (NCCL over NVLink)
    ->  insert into the local table shard with first-discovery semantics (acx_bfs_probe, order numbers
        travel with the rows)  ->  all-reduce of one 12-bit "new" mask per parent (SUM == OR: bits are
        disjoint)  ->  every rank derives the same global discovery ids, the same budget cut
        (breadth_first.py:91-95) and the same early exit (:84-85)  ->  commit (acx_bfs_commit).

Global node ids therefore equal the reference's FIFO discovery order, and (solved, path, len(tree_nodes))
are identical to the single-GPU engine and to the reference.  Host logic is plain torch + torch.distributed
and is device agnostic: tests/ drive it on CPU with gloo and a stand-in shard backend; the product backend
is GpuShard below (libacx kernels only, no CPU path)."""
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


def exchange(rows, order, lens, status, own, world, group=None):
    """all-to-all-v of candidate records to their owners.  One record = row | order (8) | lens (2) | status (1),
    padded to a multiple of 8 bytes; ONE payload collective per round (plus the count exchange)."""
    n, rowb = rows.shape
    rec = (rowb + 11 + 7) // 8 * 8
    perm = torch.argsort(own, stable=True)
    counts = torch.bincount(own, minlength=world)
    packed = torch.zeros((n, rec), dtype=torch.uint8, device=rows.device)
    packed[:, :rowb] = rows.view(torch.uint8)
    packed[:, rowb:rowb + 8] = order.view(torch.uint8).reshape(n, 8)
    packed[:, rowb + 8:rowb + 10] = lens
    packed[:, rowb + 10] = status
    send = packed[perm]
    rc = torch.empty_like(counts)
    dist.all_to_all_single(rc, counts, group=group)
    rcl, scl = [int(v) for v in rc.tolist()], [int(v) for v in counts.tolist()]
    recv = torch.empty((sum(rcl), rec), dtype=torch.uint8, device=rows.device)
    dist.all_to_all_single(recv, send, output_split_sizes=rcl, input_split_sizes=scl, group=group)
    r_rows = recv[:, :rowb].contiguous().view(torch.int8)
    r_ord = recv[:, rowb:rowb + 8].contiguous().view(torch.int64).reshape(-1)
    r_len = recv[:, rowb + 8:rowb + 10].contiguous()
    r_st = recv[:, rowb + 10].contiguous()
    return r_rows, r_ord, r_len, r_st, rec * (n - scl[dist.get_rank(group)])


class GpuShard:
    """One rank's shard on its GPU: rows, records and table of the states it owns (libacx kernels)."""

    def __init__(self, m, cyclical, cap_local, max_cand, device):
        self.m, self.rowb, self.cyc = m, 2 * m, bool(cyclical)
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
        self.n_local = 0
        self.L = _lib.lib()

    def seed(self, root):
        self.nodes[0].copy_(torch.from_numpy(root))
        self.total[0] = int(np.count_nonzero(root))
        self.gid[0] = 0
        _lib.check(self.L.acx_bfs_seed(_ptr(self.nodes), self.m, _ptr(self.table), self.tcap, 0, self.stream), "acx_bfs_seed")
        self.n_local = 1

    def hashes(self, rows):
        h = torch.empty((rows.shape[0],), dtype=torch.int64, device=self.dev)
        _lib.check(self.L.acx_rows_hash(_ptr(rows), rows.shape[0], self.m, _ptr(h), self.stream), "acx_rows_hash")
        return h

    def expand(self, rows):
        f = rows.shape[0]
        ch = torch.empty((f * 12, self.rowb), dtype=torch.int8, device=self.dev)
        ln = torch.empty((f * 12, 2), dtype=torch.uint8, device=self.dev)
        st = torch.empty((f * 12,), dtype=torch.uint8, device=self.dev)
        if f:
            _lib.check(self.L.acx_expand12(_ptr(rows), _ptr(ch), _ptr(ln), _ptr(st), f, self.m, int(self.cyc), self.stream),
                       "acx_expand12")
        return ch, ln, st

    def probe(self, cand, lens, status, order):
        n = cand.shape[0]
        slot = torch.empty((n,), dtype=torch.int64, device=self.dev)
        is_new = torch.zeros((n,), dtype=torch.uint8, device=self.dev)
        mins = torch.full((2,), -1, dtype=torch.int64, device=self.dev)
        if n:
            _lib.check(self.L.acx_bfs_probe(_ptr(cand), _ptr(lens), _ptr(status), _ptr(order), n, self.m, _ptr(self.table),
                                            self.tcap, _ptr(self.nodes), _ptr(slot), mins[0:].data_ptr(), mins[1:].data_ptr(),
                                            0, _ptr(is_new), self.stream), "acx_bfs_probe")
        mins = torch.where(mins < 0, torch.full_like(mins, SENT), mins)
        return slot, is_new, mins

    def commit(self, cand, lens, slot, is_new, excl, order, parent_base):
        n = cand.shape[0]
        if n:
            _lib.check(self.L.acx_bfs_commit(_ptr(cand), _ptr(lens), _ptr(slot), _ptr(is_new), _ptr(excl), _ptr(order),
                                             parent_base, n, n, self.m, _ptr(self.table), _ptr(self.nodes), _ptr(self.parent),
                                             _ptr(self.action), self.total.data_ptr(), self.n_local, self.stream),
                       "acx_bfs_commit")


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

    def _lookup(self, gid):
        """(parent gid, action, total) of node `gid`, from whichever rank owns it"""
        s = self.shard
        hit = torch.nonzero(s.gid[: s.n_local] == gid).flatten()
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
        h0 = s.hashes(torch.from_numpy(self.root[None, :].copy()).to(dev))
        if int(owner_of(h0, G)[0]) == self.rank:
            s.seed(self.root)
        lo, hi = 0, 1
        level_first_local = 0  # local index of the first node this rank discovered in the current frontier level
        while lo < hi:
            self.levels.append(hi - lo)
            loc = torch.arange(level_first_local, s.n_local, device=dev)
            gids = s.gid[level_first_local: s.n_local]
            gids, perm = torch.sort(gids)  # local frontier in global discovery order
            loc = loc[perm]
            level_first_local = s.n_local
            for off in range(lo, hi, self.chunk):
                C = min(self.chunk, hi - off)
                a, b = torch.searchsorted(gids, torch.tensor([off, off + C], device=dev)).tolist()
                rows = s.nodes[loc[a:b]]
                ch, ln, st = s.expand(rows)
                order = ((gids[a:b] - off) * 12)[:, None] + torch.arange(12, device=dev)[None, :]
                order = order.reshape(-1)
                own = owner_of(s.hashes(ch), G)
                r_rows, r_ord, r_len, r_st, sent = exchange(ch, order, ln, st, own, G, self.group)
                self.comm_bytes += sent
                slot, is_new, mins = s.probe(r_rows, r_len, r_st, r_ord)
                par_idx, act = r_ord // 12, r_ord % 12
                mask = torch.zeros((C,), dtype=torch.int32, device=dev)
                mask.scatter_add_(0, par_idx, (is_new.to(torch.int64) << act).to(torch.int32))
                dist.all_reduce(mask, group=self.group)
                dist.all_reduce(mins, op=dist.ReduceOp.MIN, group=self.group)
                mask = mask.to(torch.int64)
                cnt = popcount12(mask)
                cum = torch.cumsum(cnt, dim=0)
                seen_after = cum + self.n_nodes
                over = seen_after >= self.budget
                pstar = torch.argmax(over.to(torch.uint8))
                head = torch.stack([mins[0], mins[1], over.any().to(torch.int64), pstar, seen_after[pstar], cum[-1]]).tolist()
                c_s, c_e = int(head[0]), int(head[1])
                p_star = int(head[3]) if head[2] else None
                events = []
                if c_s < SENT:
                    events.append((c_s, 0, "solved"))
                if c_e < SENT:
                    events.append((c_e, 0, "error"))
                if p_star is not None:
                    events.append((12 * p_star + 11, 1, "budget"))
                if events:
                    pos, _, what = min(events)
                    if what == "solved":
                        self.n_expanded += pos // 12 + 1
                        p, aa = pos // 12, pos % 12
                        before = int(cum[p] - cnt[p]) + int(popcount12(mask[p] & ((1 << aa) - 1)))
                        self.n_nodes += before
                        return True, self._path(off + p, aa)
                    if what == "error":
                        raise AssertionError("a relator became empty during the search: not a valid presentation")
                    self.n_expanded += p_star + 1
                    self.n_nodes = int(head[4])
                    return False, None
                excl_par = cum - cnt
                grank = excl_par[par_idx] + popcount12(mask[par_idx] & ((1 << act) - 1))
                lex = torch.cumsum(is_new.to(torch.int64), dim=0) - is_new
                s.commit(r_rows, r_len, slot, is_new, lex, r_ord, off)
                sel = is_new.bool()
                n_new_local = int(sel.sum())
                s.gid[s.n_local + lex[sel]] = self.n_nodes + grank[sel]
                s.n_local += n_new_local
                self.n_nodes += int(head[5])
                self.n_expanded += C
            lo, hi = hi, self.n_nodes
        return False, None
