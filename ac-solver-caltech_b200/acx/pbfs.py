"""PlaneBFS -- the hash-partitioned level-synchronous bfs of acx.bfs_sharded with DEVICE-DRIVEN rounds
(csrc/acx_pbfs2.cu), for roots inside the plane domain (every Miller-Schupp / AK(n) presentation, m <= 36).

Same partition (owner = hash of the state), same first-discovery-by-order semantics and therefore the same
(solved, path, len(tree_nodes), nodes expanded) as ShardedBFS and as the reference's bfs (breadth_first.py:52-95).
What changed is who drives a round.  ShardedBFS reads the record counts and the round's outcome back to the host
twice per round; here the host enqueues every round of a level back to back:

    expand2 (records sorted by owner, whole 128-byte lines to the peers)  ->  all-gather of the G x G counts (also the
    barrier behind which the peer stores are complete)  ->  probe3 (a claim is final: the candidate reserves its place in
    the node store, claims the slot with it and writes its state there; first discovery settled beside the node)  ->
    all-reduce of 16 bits per parent (SUM == OR, as an int32 view)  +  all-reduce MIN of the two event orders  ->  decide
    (scan + the reference's success / budget decision, into a control block on the device)  ->  commit3 (global id, parent
    and action of the round's nodes: a streaming pass, the table is not touched again)

and looks at the control block once per level.  After a stop code every later kernel returns at once.
"""
import ctypes
import os

import numpy as np
import torch
import torch.distributed as dist

from . import _lib
from .batch import _ptr
from .bfs_sharded import PlaneShard, ShardedBFS

SENT = 1 << 62
STOP_SOLVED, STOP_ERROR, STOP_BUDGET = 1, 2, 3
WB = 64  # rounds (windows of a level) grouped per pass over the level's nodes


_count_groups = {}


def _count_group(group):
    """A second communicator over the same ranks for the count all-gathers.  NCCL runs the collectives of ONE communicator
    in issue order, so on the search's own communicator the all-gather of round k+1 would queue behind the mask all-reduce
    of round k -- which waits for probe k -- and its latency would be exposed in every round; on a communicator of its own
    it completes while probe k runs.  Collective call: every rank creates it at the same point (PlaneBFS.__init__)."""
    key = id(group) if group is not None else 0
    if key not in _count_groups:
        ranks = dist.get_process_group_ranks(group if group is not None else dist.group.WORLD)
        _count_groups[key] = dist.new_group(ranks=ranks, backend="nccl")
    return _count_groups[key]


def block_places(round_log, blo, bhi, first, n_local):
    """The places [a, b) of the local node store that can hold nodes with a global id in [blo, bhi).  round_log: one row
    {first place, end place, first global id, new nodes} per round of the level that discovered those nodes (numpy
    [rounds, 4], written by acx_pbfs_commit3), or None = unknown: every place of that level, [first, n_local).  The rounds
    of a level append their nodes to consecutive places and give them consecutive global ids, so the answer is the span
    of the rounds whose id range meets [blo, bhi); rounds that appended nothing on this rank do not count."""
    if round_log is None:
        return first, n_local
    log = np.asarray(round_log, dtype=np.int64).reshape(-1, 4)
    hit = (log[:, 2] < bhi) & (log[:, 2] + log[:, 3] > blo) & (log[:, 1] > log[:, 0])
    if not hit.any():
        return first, first
    return max(first, int(log[hit, 0].min())), min(n_local, int(log[hit, 1].max()))


class PlaneShard2(PlaneShard):
    """PlaneShard whose regions hold runs padded to whole 128-byte lines (four records), with a node store that
    acx_pbfs_probe3 appends to (head room for the rounds around a stop and for holes).  Three exchange buffers: the
    records of round k are read by probe k while round k+1 lands and round k+2 may already be staged."""

    n_bufs = 3

    def _region_cap(self, G, f):
        # every CTA of 256 parents pads each of its 3 x G runs by at most 3 null records
        pad = 9 * (f // 256 + 2)
        cap = 12 * f + pad if G == 1 else min(12 * f + pad, int(12 * f / G * 1.15) + 8192 + pad)
        return (cap + 3) // 4 * 4

    def __init__(self, m, cyclical, cap_local, max_parents, device, transport="auto"):
        # acx_pbfs_probe3 appends a node the moment it claims a state: the round that the budget cuts may append up to
        # 12 children per parent beyond the budget, and so may the round behind it (its probe runs next to the decision
        # of the round that stops the search); a reserved place whose claim lost its race can stay a hole
        cap = int(cap_local) + 24 * int(max_parents) + 5 * int(cap_local) // 100 + (1 << 18)  # holes: 1.4 - 2.2 % measured from AK(3)
        super().__init__(m, cyclical, cap, max_parents, device, transport)

    def _alloc_part(self):
        """PlaneBFS groups a level WB rounds at a time: the index buffers hold one block of rounds, not the level
        (2 x 8 bytes per node of the shard otherwise: 17 GB at 1e9 nodes)"""
        n = min(self.cap, WB * self.fmax + 4096)
        return (torch.empty((n,), dtype=torch.int64, device=self.dev), torch.empty((n,), dtype=torch.int64, device=self.dev))

    def seed(self, root):
        super().seed(root)
        self.action[0] = 255  # the root was not discovered by a move

    def probe(self, n_parents):
        raise NotImplementedError("PlaneShard2 is driven by PlaneBFS (device-driven rounds)")

    commit = probe


def make_plane_shard(root, cyclical, cap_local, max_parents, device, transport="auto"):
    """PlaneShard2 when the root is inside the plane domain, else None (use acx.bfs_sharded.make_shard)"""
    import numpy as np

    root = np.ascontiguousarray(np.asarray(root), dtype=np.int8)
    if len(root) // 2 > 36:
        return None
    sh = PlaneShard2(len(root) // 2, cyclical, cap_local, max_parents, device, transport)
    return sh if sh.accepts(root) else None


class PlaneBFS(ShardedBFS):
    """overlap=True (default): three streams form a pipeline over the rounds of a level --
        stage   expand / exchange of round k+1 and its count all-gather
        probe   probe of round k+1                      (follows probe k; needs the records of k+1 and the ring slot of k-1)
        commit  mask all-reduce, scan, decision, commit of round k
    The dependencies that remain: the all-gather of round k waits for this rank's probe of round k-1 and commit of round
    k-2, so its completion means that every rank is there and the receive buffer that round k+1 lands in (one of three)
    is free everywhere.  A probe needs nothing from the previous round's commit: a table slot names a place in the node
    store that was written by the probe which claimed it, and a place below the round's start is a visited state whether
    or not its global id has been written yet; decisions and commits stay in round order on one stream."""

    def __init__(self, root, budget, shard, chunk=1 << 20, group=None, overlap=True):
        super().__init__(root, budget, shard, chunk=chunk, group=group)
        if self.world > 1 and shard.transport != "p2p":
            raise ValueError("PlaneBFS on more than one rank needs the peer-store transport")
        if self.chunk > (1 << 31) // 12 - 1:
            raise ValueError("chunk: the new nodes of a round are counted in 32 bits")
        self.overlap = bool(overlap)
        self.count_group = _count_group(group) if (self.world > 1 and self.overlap) else group

    def _dump_trace(self, trace, path):
        import json

        torch.cuda.synchronize(self.dev)
        t0 = trace[0][2]
        rows = [(name, k, round(1e3 * t0.elapsed_time(ev), 1)) for name, k, ev in trace[1:]]  # microseconds since the start
        with open("%s_rank%d.json" % (path, self.rank), "w") as f:
            json.dump(rows, f)

    def run(self):
        try:
            return self._run()
        finally:
            if getattr(self, "_trace", None):
                self._dump_trace(*self._trace)

    def _run(self):
        s, G, dev, L = self.shard, self.world, self.dev, self.shard.L
        rank, C_max = self.rank, self.chunk
        if s.owner_of_row(self.root, G) == rank:
            s.seed(self.root)
        main = torch.cuda.current_stream(dev)
        pipe = self.overlap and self.timers is None
        st_stage = torch.cuda.Stream(dev, priority=-1) if pipe else main
        st_commit = torch.cuda.Stream(dev) if pipe else main
        ctl = torch.zeros(16, dtype=torch.int64, device=dev)
        ctl[1] = 1
        c2 = (C_max + 1) // 2 * 2
        mask16 = [torch.zeros(c2, dtype=torch.int16, device=dev) for _ in range(2)]
        mask32 = [m.view(torch.int32) for m in mask16]
        mins = [torch.empty(2, dtype=torch.int64, device=dev) for _ in range(2)]
        starts = [torch.zeros(1, dtype=torch.int64, device=dev) for _ in range(2)]  # n_local before / behind the round's probe
        ends = [torch.zeros(1, dtype=torch.int64, device=dev) for _ in range(2)]
        excl = torch.empty(C_max, dtype=torch.int32, device=dev)
        cursors = [torch.zeros(G + 1, dtype=torch.int64, device=dev) for _ in range(3)]
        allcs = [torch.empty(G * (G + 1), dtype=torch.int64, device=dev) for _ in range(3)]
        need = int(L.acx_pbfs_decide(0, 0, C_max, 0, 0, 0, 0, 0, 1, 0, 0, 0, main.cuda_stream))
        if need < 0:
            _lib.check(need, "acx_pbfs_decide")
        temp = torch.empty(max(need, 1), dtype=torch.uint8, device=dev)
        recv = s.recv_bufs if s.transport == "p2p" else s.sends
        dsts = [(ctypes.c_void_p * G)(*s.dst[b]) for b in range(3)]
        # the host runs at most `ahead` rounds in front of the device: it waits for the event of round k - ahead, looks
        # at that round's copy of the stop code (pinned memory) and stops enqueueing when the search has ended
        ahead = 4
        LOGCAP = 4096  # rounds of one level whose place ranges are logged (a longer level is grouped from all its nodes)
        round_log = torch.zeros((LOGCAP, 4), dtype=torch.int64, device=dev)
        prev_log, level_k0 = None, 0  # the log of the previous level (numpy [rounds, 4]), first round of this level
        mirror = torch.zeros((ahead + 1, 16), dtype=torch.int64).pin_memory()
        events = [None] * (ahead + 1)
        commit_ev, probed_ev = {}, {}  # round -> event behind its commit / its probe (the last few are kept)
        lo, hi, first = 0, 1, 0
        n_local = s.n_local()
        # ACX_BFS_TRACE=<path prefix>: a device-side time line of the pipeline (one event per step and stream), for tuning
        trace_to = os.environ.get("ACX_BFS_TRACE") if pipe else None
        trace = []

        def mk(stream, name, k):
            if trace_to:
                ev = torch.cuda.Event(enable_timing=True)
                ev.record(stream)
                trace.append((name, k, ev))

        mk(main, "t0", -1)
        self._trace = (trace, trace_to) if trace_to else None

        def stage(k, off, a, b):
            """stage stream: expand the local parents of round k and route the records to their owners"""
            cur = cursors[k % 3]
            with torch.cuda.stream(st_stage):
                if pipe and (k - 3) in commit_ev:
                    st_stage.wait_event(commit_ev[k - 3])  # on one rank the decision of round k-3 read these counts
                cur.zero_()
                mk(st_stage, "expand_begin", k)
                if b > a:
                    _lib.check(L.acx_pbfs_expand2(_ptr(s.nodes), _ptr(s.action), loc[a:b].data_ptr(), gids[a:b].data_ptr(), off, b - a,
                                                  s.m, int(s.cyc), G, s.rcap, _ptr(cur), dsts[k % 3], _ptr(ctl), st_stage.cuda_stream), "acx_pbfs_expand2")
                mk(st_stage, "expand_end", k)
            return cur

        def gather(k, cur):
            """stage stream: the count all-gather of round k, once this rank has probed round k-2 -- when it completes, every
            rank has written its records of round k and is done with the receive buffer (one of three, read by the probe
            of its round only) that round k+1 will land in.  It waits neither for the probe of round k-1 nor for the
            commit of round k-2, and it runs on a communicator of its own (_count_group): its latency hides behind probe k-1."""
            allc = allcs[k % 3]
            with torch.cuda.stream(st_stage):
                if pipe:
                    if (k - 2) in probed_ev:
                        st_stage.wait_event(probed_ev[k - 2])
                    if (k - 3) in commit_ev:
                        st_stage.wait_event(commit_ev[k - 3])  # the decision of round k-3 read this ring slot of the counts
                mk(st_stage, "allgather_begin", k)
                if G > 1:
                    dist.all_gather_into_tensor(allc, cur, group=self.count_group)
                mk(st_stage, "allgather_end", k)
                ev = torch.cuda.Event()
                ev.record(st_stage)
            return (allc if G > 1 else cur), ev

        while lo < hi:
            self.levels.append(hi - lo)
            self._tick(None)
            gl = s.gid[first:n_local]
            nwin = (hi - lo + C_max - 1) // C_max
            stopped = False
            # The level's windows (rounds) are grouped WB at a time: a level that the budget cuts after a few rounds -- the last
            # one, the largest -- is not sorted beyond them, and the grouping buffers hold one block's nodes at a time.
            for wb in range(0, nwin, WB):
                nb = min(WB, nwin - wb)
                blo = lo + wb * C_max
                bhi = min(hi, blo + nb * C_max)
                if wb and pipe:
                    main.wait_stream(st_stage)  # the staged rounds of the previous block read the grouping buffers
                if nwin == 1:  # one round: the local frontier as it is (no grouping pass, no read-back)
                    loc = torch.arange(first, n_local, dtype=torch.int64, device=dev, out=s._part[0][: n_local - first])
                    gids, cuts = gl, [0, n_local - first]
                else:
                    a, b = block_places(prev_log, blo, bhi, first, n_local)
                    loc, gids, cuts = s.window_partition(s.gid[a:b], a, blo, C_max, nb)
                self._tick("level_sort")
                offs = list(range(blo, bhi, C_max))
                if pipe:
                    st_stage.wait_stream(main)  # loc / gids of this block; every commit of the previous level (joined below)
                    st_commit.wait_stream(main)
                cur = stage(self.rounds, offs[0], cuts[0], cuts[1])
                self._tick("expand_exchange")
                pending = gather(self.rounds, cur)
                self._tick("count_allgather")
                for i, off in enumerate(offs):
                    C = min(C_max, hi - off)
                    k = self.rounds
                    ring = k % (ahead + 1)
                    if events[ring] is not None:
                        events[ring].synchronize()
                        if int(mirror[ring, 0]) != 0:
                            stopped = True
                            break
                    r2 = k % 2
                    buf = recv[k % 3]
                    counts, ready = pending
                    nxt = stage(k + 1, offs[i + 1], cuts[i + 1], cuts[i + 2]) if i + 1 < len(offs) else None
                    self._tick("expand_exchange")
                    rc_ptr = counts.data_ptr() + 8 * rank  # records rank s staged for this rank: counts[s * (G + 1) + rank]
                    # ---- probe stream (main)
                    if pipe:
                        main.wait_event(ready)
                        if (k - 2) in commit_ev:
                            main.wait_event(commit_ev[k - 2])  # slot / mask / mins of this ring position are free again
                    mask16[r2][: (C + 1) // 2 * 2].zero_()
                    mins[r2].fill_(SENT)
                    starts[r2].copy_(s.n_local_dev)
                    mk(main, "probe_begin", k)
                    _lib.check(L.acx_pbfs_probe3(_ptr(buf), rc_ptr, G + 1, G, s.rcap, _ptr(s.table), s.tcap, _ptr(s.nodes), s.cap, _ptr(s.gid),
                                                 _ptr(s.n_local_dev), _ptr(starts[r2]), _ptr(mins[r2]), _ptr(mask32[r2]), _ptr(ctl),
                                                 main.cuda_stream), "acx_pbfs_probe3")
                    ends[r2].copy_(s.n_local_dev)
                    mk(main, "probe_end", k)
                    self._tick("probe")
                    probed = torch.cuda.Event()
                    probed.record(main)
                    probed_ev[k] = probed
                    probed_ev.pop(k - 4, None)
                    # ---- commit stream
                    with torch.cuda.stream(st_commit):
                        if pipe:
                            st_commit.wait_event(probed)
                        mk(st_commit, "reduce_begin", k)
                        if G > 1:
                            if self.timers is not None:
                                dist.barrier(group=self.group)
                                self._tick("wait_slowest_rank")
                            dist.all_reduce(mask32[r2][: (C + 1) // 2], group=self.group)
                            dist.all_reduce(mins[r2], op=dist.ReduceOp.MIN, group=self.group)
                        mk(st_commit, "reduce_end", k)
                        cs = st_commit.cuda_stream
                        rcd = int(L.acx_pbfs_decide(_ptr(mask16[r2]), _ptr(excl), C, _ptr(mins[r2]), off, self.budget, _ptr(ctl), _ptr(counts), G,
                                                    rank, _ptr(temp), need, cs))
                        if rcd:
                            _lib.check(rcd, "acx_pbfs_decide")
                        self._tick("reduce_scan_decide")
                        _lib.check(L.acx_pbfs_commit3(_ptr(starts[r2]), _ptr(ends[r2]), G * s.rcap, s.cap, _ptr(excl), _ptr(mask32[r2]), off,
                                                      _ptr(ctl), _ptr(s.gid), _ptr(s.parent), _ptr(s.action),
                                                  round_log[k - level_k0].data_ptr() if k - level_k0 < LOGCAP else 0, cs), "acx_pbfs_commit3")
                        self._tick("commit")
                        mk(st_commit, "commit_end", k)
                        commit_ev[k] = torch.cuda.Event()
                        commit_ev[k].record(st_commit)
                        commit_ev.pop(k - 5, None)
                        mirror[ring].copy_(ctl, non_blocking=True)
                        events[ring] = torch.cuda.Event()
                        events[ring].record(st_commit)
                    self.rounds += 1
                    if nxt is not None:
                        pending = gather(k + 1, nxt)
                        self._tick("count_allgather")
                if stopped:
                    break
            first = n_local
            if pipe:  # join: the level's commits, and a staged round that may still be in flight after an early exit
                main.wait_stream(st_commit)
                main.wait_stream(st_stage)
            mk(main, "level_end", len(self.levels))
            n_logged = self.rounds - level_k0
            logged = n_logged <= LOGCAP
            c = torch.cat([ctl, s.n_local_dev, round_log[:n_logged].flatten() if logged else round_log[:0].flatten()]).tolist()  # the one read-back of the level
            n_local = c[16]
            prev_log = np.asarray(c[17:], dtype=np.int64).reshape(-1, 4) if logged else None
            level_k0 = self.rounds
            self.n_nodes, self.n_expanded = c[1], c[2]
            if c[6]:
                raise RuntimeError("plane bfs: a send region or the node store overflowed (hash imbalance above the head room)")
            self.n_candidates, self.comm_bytes = c[8], 32 * (c[8] - c[9])
            if c[0]:
                if c[0] == STOP_SOLVED:
                    return True, self._path(c[5] + c[4] // 12, c[4] % 12)
                if c[0] == STOP_ERROR:
                    raise AssertionError("a relator became empty during the search: not a valid presentation")
                return False, None
            lo, hi = hi, self.n_nodes
        return False, None
