"""PlaneBFS -- the hash-partitioned level-synchronous bfs of acx.bfs_sharded with DEVICE-DRIVEN rounds
(csrc/acx_pbfs2.cu), for roots inside the plane domain (every Miller-Schupp / AK(n) presentation, m <= 36).

Same partition (owner = hash of the state), same first-discovery-by-order semantics and therefore the same
(solved, path, len(tree_nodes), nodes expanded) as ShardedBFS and as the reference's bfs (breadth_first.py:52-95).
What changed is who drives a round.  ShardedBFS reads the record counts and the round's outcome back to the host
twice per round; here the host enqueues every round of a level back to back:

    expand2 (records sorted by owner, whole 128-byte lines to the peers)  ->  all-gather of the G x G counts (also the
    barrier behind which the peer stores are complete)  ->  probe2 (winner resolution fused in)  ->  all-reduce of
    16 bits per parent (SUM == OR, as an int32 view)  +  all-reduce MIN of the two event orders  ->  decide (scan + the
    reference's success / budget decision, into a control block on the device)  ->  commit2

and looks at the control block once per level.  After a stop code every later kernel returns at once.
"""
import ctypes

import torch
import torch.distributed as dist

from . import _lib
from .batch import _ptr
from .bfs_sharded import PlaneShard, ShardedBFS

SENT = 1 << 62
STOP_SOLVED, STOP_ERROR, STOP_BUDGET = 1, 2, 3


class PlaneShard2(PlaneShard):
    """PlaneShard whose regions hold runs padded to whole 128-byte lines (four records)"""

    def _region_cap(self, G, f):
        # every CTA of 256 parents pads each of its 3 x G runs by at most 3 null records
        pad = 9 * (f // 256 + 2)
        cap = 12 * f + pad if G == 1 else min(12 * f + pad, int(12 * f / G * 1.15) + 8192 + pad)
        return (cap + 3) // 4 * 4

    def seed(self, root):
        super().seed(root)
        self.action[0] = 255  # the root was not discovered by a move


def make_plane_shard(root, cyclical, cap_local, max_parents, device, transport="auto"):
    """PlaneShard2 when the root is inside the plane domain, else None (use acx.bfs_sharded.make_shard)"""
    import numpy as np

    root = np.ascontiguousarray(np.asarray(root), dtype=np.int8)
    if len(root) // 2 > 36:
        return None
    sh = PlaneShard2(len(root) // 2, cyclical, cap_local, max_parents, device, transport)
    return sh if sh.accepts(root) else None


class PlaneBFS(ShardedBFS):
    def __init__(self, root, budget, shard, chunk=1 << 20, group=None):
        super().__init__(root, budget, shard, chunk=chunk, group=group)
        if self.world > 1 and shard.transport != "p2p":
            raise ValueError("PlaneBFS on more than one rank needs the peer-store transport")
        if self.chunk > (1 << 31) // 12 - 1:
            raise ValueError("chunk: the new nodes of a round are counted in 32 bits")

    def run(self):
        s, G, dev, L = self.shard, self.world, self.dev, self.shard.L
        rank, C_max = self.rank, self.chunk
        stream = s.stream
        if s.owner_of_row(self.root, G) == rank:
            s.seed(self.root)
        ctl = torch.zeros(16, dtype=torch.int64, device=dev)
        ctl[1] = 1
        c2 = (C_max + 1) // 2 * 2
        mask16 = torch.zeros(c2, dtype=torch.int16, device=dev)
        mask32 = mask16.view(torch.int32)
        excl = torch.empty(C_max, dtype=torch.int32, device=dev)
        mins = torch.empty(2, dtype=torch.int64, device=dev)
        cursor = torch.zeros(G + 1, dtype=torch.int64, device=dev)
        allc = torch.empty(G * (G + 1), dtype=torch.int64, device=dev)
        slot = torch.empty(G * s.rcap, dtype=torch.int64, device=dev)
        need = int(L.acx_pbfs_decide(0, 0, C_max, 0, 0, 0, 0, 0, 1, 0, 0, 0, stream))
        if need < 0:
            _lib.check(need, "acx_pbfs_decide")
        temp = torch.empty(max(need, 1), dtype=torch.uint8, device=dev)
        recv = s.recv_bufs if s.transport == "p2p" else s.sends
        dsts = [(ctypes.c_void_p * G)(*s.dst[b]) for b in range(2)]
        # the host runs at most `ahead` rounds in front of the device: it waits for the event of round k - ahead, looks
        # at that round's copy of the stop code (pinned memory) and stops enqueueing when the search has ended
        ahead = 4
        mirror = torch.zeros((ahead + 1, 16), dtype=torch.int64).pin_memory()
        events = [None] * (ahead + 1)
        lo, hi, first = 0, 1, 0
        while lo < hi:
            self.levels.append(hi - lo)
            self._tick(None)
            n_local = s.n_local()
            gl = s.gid[first:n_local]
            nwin = (hi - lo + C_max - 1) // C_max
            loc, gids, cuts = s.window_partition(gl, first, lo, C_max, nwin)
            first = n_local
            self._tick("level_sort")
            for i, off in enumerate(range(lo, hi, C_max)):
                C = min(C_max, hi - off)
                a, b = cuts[i], cuts[i + 1]
                k = self.rounds
                ring = k % (ahead + 1)
                if events[ring] is not None:
                    events[ring].synchronize()
                    if int(mirror[ring, 0]) != 0:
                        break
                buf = recv[k % 2]
                cursor.zero_()
                if b > a:
                    _lib.check(L.acx_pbfs_expand2(_ptr(s.nodes), _ptr(s.action), loc[a:b].data_ptr(), gids[a:b].data_ptr(), off, b - a,
                                                  s.m, int(s.cyc), G, s.rcap, _ptr(cursor), dsts[k % 2], _ptr(ctl), stream), "acx_pbfs_expand2")
                self._tick("expand_exchange")
                if G > 1:
                    dist.all_gather_into_tensor(allc, cursor, group=self.group)
                    counts = allc
                else:
                    counts = cursor
                rc_ptr = counts.data_ptr() + 8 * rank  # records rank s staged for this rank: counts[s * (G + 1) + rank]
                self._tick("count_allgather")
                mask16[: (C + 1) // 2 * 2].zero_()
                mins.fill_(SENT)
                _lib.check(L.acx_pbfs_probe2(_ptr(buf), rc_ptr, G + 1, G, s.rcap, _ptr(s.table), s.tcap, _ptr(s.nodes), _ptr(slot), _ptr(mins),
                                             _ptr(mask32), _ptr(ctl), stream), "acx_pbfs_probe2")
                self._tick("probe")
                if G > 1:
                    if self.timers is not None:
                        dist.barrier(group=self.group)
                        self._tick("wait_slowest_rank")
                    dist.all_reduce(mask32[: (C + 1) // 2], group=self.group)
                    dist.all_reduce(mins, op=dist.ReduceOp.MIN, group=self.group)
                rcd = int(L.acx_pbfs_decide(_ptr(mask16), _ptr(excl), C, _ptr(mins), off, self.budget, _ptr(ctl), _ptr(counts), G, rank,
                                            _ptr(temp), need, stream))
                if rcd:
                    _lib.check(rcd, "acx_pbfs_decide")
                self._tick("reduce_scan_decide")
                _lib.check(L.acx_pbfs_commit2(_ptr(buf), rc_ptr, G + 1, G, s.rcap, _ptr(slot), _ptr(excl), _ptr(mask32), off, _ptr(ctl),
                                              _ptr(s.n_local_dev), _ptr(s.table), _ptr(s.nodes), _ptr(s.gid), _ptr(s.parent), _ptr(s.action),
                                              stream), "acx_pbfs_commit2")
                self._tick("commit")
                mirror[ring].copy_(ctl, non_blocking=True)
                events[ring] = torch.cuda.Event()
                events[ring].record()
                self.rounds += 1
            c = ctl.tolist()  # the one read-back of the level
            self.n_nodes, self.n_expanded = c[1], c[2]
            if c[6]:
                raise RuntimeError("plane bfs: a send region overflowed (hash imbalance above the 15% head room)")
            self.n_candidates, self.comm_bytes = c[8], 32 * (c[8] - c[9])
            if c[0]:
                if c[0] == STOP_SOLVED:
                    return True, self._path(c[5] + c[4] // 12, c[4] % 12)
                if c[0] == STOP_ERROR:
                    raise AssertionError("a relator became empty during the search: not a valid presentation")
                return False, None
            lo, hi = hi, self.n_nodes
        return False, None
