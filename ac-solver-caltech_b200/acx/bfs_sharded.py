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
    return ((rowb + 7) // 8 * 8 + 16 + 15) // 16 * 16


_POP_LUT = {}


def popcount_lut(mask):
    """popcount of 12-bit masks through a 4096-entry table (one gather)"""
    key = str(mask.device)
    if key not in _POP_LUT:
        _POP_LUT[key] = popcount12(torch.arange(4096, dtype=torch.int64, device=mask.device))
    return _POP_LUT[key][mask.to(torch.int64)]


class _RawCuda:
    """device memory owned by libacx (acx_peer_alloc) exposed to torch through the CUDA array interface"""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (int(nbytes),), "typestr": "|u1", "data": (int(ptr), False), "version": 2}


class GpuShard:
    """One rank's shard on its GPU: rows, records and table of the states it owns (libacx kernels).

    transport "p2p" (default for max_relator_length 36 on more than one GPU): every rank owns two peer-visible
    receive buffers (CUDA IPC); the fused expand kernel of rank s stores each record straight into region s of
    its owner's buffer over NVLink -- the kernel IS the all-to-all, there is no staging copy and no collective on
    the data path (only the count all-gather, which doubles as the barrier).  transport "nccl": records are
    bucketed into a local send buffer and exchanged with an all-to-all-v."""

    n_bufs = 2  # exchange buffers: round k + 1 may be staged while round k is still being probed / committed

    def __init__(self, m, cyclical, cap_local, max_parents, device, transport="auto"):
        self.m, self.rowb, self.cyc = m, 2 * m, bool(cyclical)
        self.rec = record_bytes(self.rowb)
        self.dev = torch.device(device)
        self.stream = _prep(self.dev)
        d = self.dev
        self.cap = int(cap_local)
        self.fmax = int(max_parents)
        self.tcap = _next_pow2(2 * (self.cap + 12 * self.fmax))
        self.nodes = torch.zeros((self.cap, self.rowb), dtype=torch.int8, device=d)
        self.gid = torch.full((self.cap,), -1, dtype=torch.int64, device=d)
        self.parent = torch.full((self.cap,), -1, dtype=torch.int64, device=d)
        self.action = torch.zeros((self.cap,), dtype=torch.uint8, device=d)
        self.total = torch.zeros((self.cap,), dtype=torch.int16, device=d)
        self.table = torch.zeros((self.tcap,), dtype=torch.int64, device=d)
        self.n_local_dev = torch.zeros((1,), dtype=torch.int64, device=d)
        self.L = _lib.lib()
        assert self.L.acx_bfs_record_bytes(m) == self.rec
        self.transport = transport
        self._own, self._opened = [], []
        self.world = None

    # -- wiring -----------------------------------------------------------------------------------------------
    def connect(self, group=None):
        """allocate the exchange buffers for this process group (collective: every rank calls it)"""
        if self.world is not None:
            return
        G, rank = dist.get_world_size(group), dist.get_rank(group)
        self.world, self.rank = G, rank
        self.fused = self._fusable(G)
        f = self.fmax
        self.rcap = self._region_cap(G, f)
        if self.transport == "auto":
            self.transport = "p2p" if (self.fused and G > 1) else "nccl"
        if self.transport == "p2p" and not (self.fused and G > 1):
            raise ValueError("transport 'p2p' needs the fused expand kernel (max_relator_length 36 or plane nodes) and 2..32 ranks")
        nbytes = G * self.rcap * self.rec
        if self.transport == "p2p":
            import ctypes

            self.recv_bufs, self.dst = [], []
            for _ in range(self.n_bufs):
                ptr = ctypes.c_void_p()
                _lib.check(self.L.acx_peer_alloc(nbytes, ctypes.byref(ptr)), "acx_peer_alloc")
                self._own.append(ptr.value)
                h = (ctypes.c_uint8 * 64)()
                _lib.check(self.L.acx_peer_export(ptr, h), "acx_peer_export")
                mine = torch.tensor(list(h), dtype=torch.uint8, device=self.dev)
                allh = torch.empty((G * 64,), dtype=torch.uint8, device=self.dev)
                dist.all_gather_into_tensor(allh, mine, group=group)
                allh = allh.view(G, 64).cpu().numpy()
                bases = []
                for g in range(G):
                    if g == rank:
                        bases.append(ptr.value)
                        continue
                    hg = (ctypes.c_uint8 * 64)(*[int(v) for v in allh[g]])
                    pg = ctypes.c_void_p()
                    _lib.check(self.L.acx_peer_open(hg, ctypes.byref(pg)), "acx_peer_open")
                    self._opened.append(pg.value)
                    bases.append(pg.value)
                self.recv_bufs.append(torch.as_tensor(_RawCuda(ptr.value, nbytes), device=self.dev).view(G * self.rcap, self.rec))
                # my region inside every owner's receive buffer
                self.dst.append([bases[g] + rank * self.rcap * self.rec for g in range(G)])
        elif self.fused:
            # two send buffers: round k + 1 may be staged while round k is still being exchanged / probed
            self.sends = [torch.empty((G * self.rcap, self.rec), dtype=torch.uint8, device=self.dev) for _ in range(self.n_bufs)]
            self.dst = [[b.data_ptr() + g * self.rcap * self.rec for g in range(G)] for b in self.sends]
        self._group = group
        self._part = self._alloc_part()

    def _alloc_part(self):
        """frontier index buffers of a level (local index, global id), allocated once: a level can be most of the shard"""
        return (torch.empty((self.cap,), dtype=torch.int64, device=self.dev), torch.empty((self.cap,), dtype=torch.int64, device=self.dev))

    def _fusable(self, world):
        return self.m == 36 and world <= 32

    def _region_cap(self, G, f):
        """records one sender may put into one owner's region per round: all 12 f for one rank, else the mean plus 15 %"""
        return 12 * f if G == 1 else min(12 * f, int(12 * f / G * 1.15) + 8192)

    def node_total(self, k):
        return self.total[k].to(torch.int64)

    def close(self):
        """collective: nobody may still be writing into a peer buffer when it is unmapped / freed"""
        if self._own or self._opened:
            torch.cuda.synchronize(self.dev)
            dist.barrier(group=self._group)
            for p in self._opened:
                self.L.acx_peer_close(p)
            self.recv_bufs = []
            dist.barrier(group=self._group)
            for p in self._own:
                self.L.acx_peer_free(p)
            self._own, self._opened = [], []

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

    # -- one round --------------------------------------------------------------------------------------------
    def stage(self, loc, gids, ord_base, k):
        """expand the parents nodes[loc] (global ids `gids`) and route the candidate records towards their owners.
        Returns counts int64[world + 1] on the device: records for each owner, then a region-overflow flag."""
        import ctypes

        G, f, d = self.world, loc.shape[0], self.dev
        if f > self.fmax:
            raise RuntimeError("sharded bfs: %d local parents in one round, buffers sized for %d" % (f, self.fmax))
        n = 12 * f
        counts = torch.zeros((G + 1,), dtype=torch.int64, device=d)
        self._packed = None
        if f == 0:
            return counts
        loc, gids = loc.contiguous(), gids.contiguous()
        if self.fused:
            dst = (ctypes.c_void_p * G)(*self.dst[k % 2])
            _lib.check(self.L.acx_bfs_expand_bucket(_ptr(self.nodes), _ptr(loc), _ptr(gids), int(ord_base), f, self.m,
                                                    int(self.cyc), G, self.rcap, _ptr(counts), dst, self.stream),
                       "acx_bfs_expand_bucket")
            return counts
        rows = self.nodes[loc]
        send = torch.empty((n, self.rec), dtype=torch.uint8, device=d)
        ch = torch.empty((n, self.rowb), dtype=torch.int8, device=d)
        ln = torch.empty((n, 2), dtype=torch.uint8, device=d)
        st = torch.empty((n,), dtype=torch.uint8, device=d)
        _lib.check(self.L.acx_expand12(_ptr(rows), _ptr(ch), _ptr(ln), _ptr(st), f, self.m, int(self.cyc), self.stream), "acx_expand12")
        own = torch.empty((n,), dtype=torch.uint8, device=d)
        po = (gids - ord_base).contiguous()
        _lib.check(self.L.acx_bfs_bucket(_ptr(ch), 0, 0, 0, n, self.m, G, _ptr(own), _ptr(counts), 0, 0, 0, 0, self.stream),
                   "acx_bfs_bucket")
        offsets = torch.cumsum(counts[:G], 0) - counts[:G]
        cursor = torch.zeros((G,), dtype=torch.int64, device=d)
        _lib.check(self.L.acx_bfs_bucket(_ptr(ch), _ptr(ln), _ptr(st), _ptr(po), n, self.m, G, _ptr(own), _ptr(counts),
                                         _ptr(offsets), _ptr(cursor), _ptr(send), 1, self.stream), "acx_bfs_bucket")
        self._packed = send
        return counts

    def deliver(self, allc, k):
        """allc[s][o] = records rank s staged for owner o.  After this call the records this rank owns are in
        self._recv as len(self._regions) regions of self._rcap_now records."""
        G, rank = self.world, self.rank
        scl = allc[rank][:G]
        rcl = [allc[g][rank] for g in range(G)]
        if self.transport == "p2p":  # already here: the senders' kernels wrote them
            self._recv, self._regions, self._rcap_now = self.recv_bufs[k % 2], rcl, self.rcap
            return
        if self._packed is not None or not self.fused:
            send = self._packed if self._packed is not None else torch.empty((0, self.rec), dtype=torch.uint8, device=self.dev)
            starts = [sum(scl[:g]) for g in range(G)]
        else:
            send, starts = self.sends[k % 2], [g * self.rcap for g in range(G)]
        if G == 1:
            self._recv, self._regions, self._rcap_now = send[starts[0]:starts[0] + scl[0]], [scl[0]], max(scl[0], 1)
            return
        recv = torch.empty((sum(rcl), self.rec), dtype=torch.uint8, device=self.dev)
        dist.all_to_all(list(recv.split(rcl)), [send[starts[g]:starts[g] + scl[g]] for g in range(G)], group=self._group)
        self._recv, self._regions, self._rcap_now = recv, [sum(rcl)], max(sum(rcl), 1)

    def window_partition(self, gl, first, lo, chunk, nwin):
        """local frontier (global ids gl, stored from local index `first`) grouped by round: (loc, gids, cuts)"""
        n = gl.shape[0]
        d = self.dev
        counts = torch.zeros((nwin,), dtype=torch.int64, device=d)
        loc, gids = self._part[0][:n], self._part[1][:n]
        if n:
            gl = gl.contiguous()
            _lib.check(self.L.acx_bfs_window_partition(_ptr(gl), n, lo, chunk, nwin, _ptr(counts), 0, 0, first, 0, 0, 0, self.stream),
                       "acx_bfs_window_partition")
        ends = torch.cumsum(counts, 0)
        if n:
            offsets = (ends - counts).contiguous()
            cursor = torch.zeros((nwin,), dtype=torch.int64, device=d)
            _lib.check(self.L.acx_bfs_window_partition(_ptr(gl), n, lo, chunk, nwin, _ptr(counts), _ptr(offsets), _ptr(cursor), first,
                                                       _ptr(loc), _ptr(gids), 1, self.stream), "acx_bfs_window_partition")
        return loc, gids, [0] + ends.tolist()

    def mask_scan(self, mask):
        """(exclusive scan of popcount(mask), total) on the device, one cub pass"""
        n = mask.shape[0]
        excl = torch.empty((n,), dtype=torch.int64, device=self.dev)
        total = torch.zeros((1,), dtype=torch.int64, device=self.dev)
        need = int(self.L.acx_bfs_mask_scan(0, n, 0, 0, 0, 0, self.stream))
        if need < 0:
            _lib.check(need, "acx_bfs_mask_scan")
        temp = torch.empty((max(need, 1),), dtype=torch.uint8, device=self.dev)
        _lib.check(int(self.L.acx_bfs_mask_scan(_ptr(mask), n, _ptr(excl), _ptr(total), _ptr(temp), need, self.stream)),
                   "acx_bfs_mask_scan")
        return excl, total

    def _region_args(self):
        import ctypes

        return (ctypes.c_int64 * len(self._regions))(*self._regions), len(self._regions), self._rcap_now

    def probe(self, n_parents):
        rc, nreg, rcap = self._region_args()
        self._slot = torch.empty((nreg * rcap,), dtype=torch.int64, device=self.dev)
        mask = torch.zeros((n_parents,), dtype=torch.int32, device=self.dev)
        mins = torch.full((2,), -1, dtype=torch.int64, device=self.dev)
        if sum(self._regions):
            _lib.check(self.L.acx_bfs_probe_packed(_ptr(self._recv), rc, nreg, rcap, self.m, _ptr(self.table), self.tcap,
                                                   _ptr(self.nodes), _ptr(self._slot), _ptr(mins), _ptr(mask), self.stream),
                       "acx_bfs_probe_packed")
        mins = torch.where(mins < 0, torch.full_like(mins, SENT), mins)
        return mask, mins

    def commit(self, excl_par, gmask, parent_base, gid_base):
        rc, nreg, rcap = self._region_args()
        if sum(self._regions):
            _lib.check(self.L.acx_bfs_commit_packed(_ptr(self._recv), rc, nreg, rcap, self.m, _ptr(self._slot), _ptr(excl_par),
                                                    _ptr(gmask), parent_base, gid_base, _ptr(self.n_local_dev), _ptr(self.table),
                                                    _ptr(self.nodes), _ptr(self.gid), _ptr(self.parent), _ptr(self.action),
                                                    self.total.data_ptr(), self.stream), "acx_bfs_commit_packed")


class PlaneShard(GpuShard):
    """The shard on bit-plane nodes (csrc/acx_pbfs.cu): a node / a candidate record is 32 bytes instead of 2m / 2m+24,
    the moves run on the planes in registers and nothing is packed or unpacked on the way.  For max_relator_length
    <= 36 and roots inside the plane domain (PlaneShard.accepts); same protocol and transports as GpuShard."""

    def __init__(self, m, cyclical, cap_local, max_parents, device, transport="auto"):
        if m > 36:
            raise ValueError("plane nodes hold max_relator_length <= 36")
        self.m, self.rowb, self.cyc = m, 2 * m, bool(cyclical)
        self.rec = 32
        self.dev = torch.device(device)
        self.stream = _prep(self.dev)
        d = self.dev
        self.cap = int(cap_local)
        self.fmax = int(max_parents)
        self.tcap = _next_pow2(2 * (self.cap + 12 * self.fmax))
        self.nodes = torch.zeros((self.cap, 4), dtype=torch.int64, device=d)
        self.gid = torch.full((self.cap,), -1, dtype=torch.int64, device=d)
        self.parent = torch.full((self.cap,), -1, dtype=torch.int64, device=d)
        self.action = torch.zeros((self.cap,), dtype=torch.uint8, device=d)
        self.table = torch.zeros((self.tcap,), dtype=torch.int64, device=d)
        self.n_local_dev = torch.zeros((1,), dtype=torch.int64, device=d)
        self.L = _lib.lib()
        self.transport = transport
        self._own, self._opened = [], []
        self.world = None

    def _fusable(self, world):
        return world <= 32

    def pack(self, rows):
        """byte rows [n, 2m] (numpy) -> (plane nodes int64 [n, 4] on the device, ok uint8 [n])"""
        r = torch.from_numpy(np.ascontiguousarray(rows, dtype=np.int8)).to(self.dev)
        out = torch.empty((r.shape[0], 4), dtype=torch.int64, device=self.dev)
        ok = torch.empty((r.shape[0],), dtype=torch.uint8, device=self.dev)
        _lib.check(self.L.acx_pbfs_pack(_ptr(r), r.shape[0], self.m, int(self.cyc), _ptr(out), _ptr(ok), self.stream), "acx_pbfs_pack")
        return out, ok

    def unpack(self, nodes):
        rows = torch.empty((nodes.shape[0], self.rowb), dtype=torch.int8, device=self.dev)
        _lib.check(self.L.acx_pbfs_unpack(_ptr(nodes.contiguous()), nodes.shape[0], self.m, _ptr(rows), self.stream), "acx_pbfs_unpack")
        return rows

    def accepts(self, root):
        """is the root inside the plane domain (letters +-1, +-2, freely reduced, cyclically reduced when cyclical)?"""
        node, ok = self.pack(root[None, :])
        self._root_node = node
        return bool(ok[0])

    def owner_of_row(self, row, world):
        if not self.accepts(row):
            raise ValueError("root outside the plane domain: use GpuShard")
        h = torch.empty((1,), dtype=torch.int64, device=self.dev)
        _lib.check(self.L.acx_pbfs_hash(_ptr(self._root_node), 1, _ptr(h), self.stream), "acx_pbfs_hash")
        return int(owner_of(h, world)[0])

    def seed(self, root):
        self.nodes[0].copy_(self._root_node[0])
        self.gid[0] = 0
        _lib.check(self.L.acx_pbfs_seed(_ptr(self.nodes), _ptr(self.table), self.tcap, 0, self.stream), "acx_pbfs_seed")
        self.n_local_dev.fill_(1)

    def node_total(self, k):
        w = self.nodes[k]
        return ((w[0] >> 40) & 63) + ((w[1] >> 40) & 63)

    overlap = True  # stage(k + 1) may run on a side stream while round k is probed and committed

    def stage(self, loc, gids, ord_base, k, stream=None, ctas=0):
        import ctypes

        G, f = self.world, loc.shape[0]
        if f > self.fmax:
            raise RuntimeError("sharded bfs: %d local parents in one round, buffers sized for %d" % (f, self.fmax))
        counts = torch.zeros((G + 1,), dtype=torch.int64, device=self.dev)
        self._packed = None
        if f:
            loc, gids = loc.contiguous(), gids.contiguous()
            dst = (ctypes.c_void_p * G)(*self.dst[k % 2])
            _lib.check(self.L.acx_pbfs_expand(_ptr(self.nodes), _ptr(loc), _ptr(gids), int(ord_base), f, self.m, int(self.cyc), G,
                                              self.rcap, _ptr(counts), dst, ctas, self.stream if stream is None else stream),
                       "acx_pbfs_expand")
        return counts

    def probe(self, n_parents):
        rc, nreg, rcap = self._region_args()
        self._slot = torch.empty((nreg * rcap,), dtype=torch.int64, device=self.dev)
        mask = torch.zeros((n_parents,), dtype=torch.int32, device=self.dev)
        mins = torch.full((2,), -1, dtype=torch.int64, device=self.dev)
        if sum(self._regions):
            _lib.check(self.L.acx_pbfs_probe(_ptr(self._recv), rc, nreg, rcap, _ptr(self.table), self.tcap, _ptr(self.nodes),
                                             _ptr(self._slot), _ptr(mins), _ptr(mask), self.stream), "acx_pbfs_probe")
        mins = torch.where(mins < 0, torch.full_like(mins, SENT), mins)
        return mask, mins

    def commit(self, excl_par, gmask, parent_base, gid_base):
        rc, nreg, rcap = self._region_args()
        if sum(self._regions):
            _lib.check(self.L.acx_pbfs_commit(_ptr(self._recv), rc, nreg, rcap, _ptr(self._slot), _ptr(excl_par), _ptr(gmask),
                                              parent_base, gid_base, _ptr(self.n_local_dev), _ptr(self.table), _ptr(self.nodes),
                                              _ptr(self.gid), _ptr(self.parent), _ptr(self.action), self.stream), "acx_pbfs_commit")


def make_shard(root, cyclical, cap_local, max_parents, device, transport="auto"):
    """PlaneShard when the root is inside the plane domain (every Miller-Schupp / AK(n) presentation), else GpuShard"""
    root = np.ascontiguousarray(np.asarray(root), dtype=np.int8)
    m = len(root) // 2
    if m <= 36:
        sh = PlaneShard(m, cyclical, cap_local, max_parents, device, transport)
        if sh.accepts(root):
            return sh
        del sh
    return GpuShard(m, cyclical, cap_local, max_parents, device, "nccl" if (transport == "auto" and m != 36) else transport)


class ShardedBFS:
    def __init__(self, root, budget, shard, chunk=1 << 20, group=None, pipeline=False, stage_ctas=2):
        self.root = np.ascontiguousarray(np.asarray(root), dtype=np.int8)
        self.budget = max(1, int(budget))
        self.shard = shard
        self.group = group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.chunk = int(chunk)
        self.dev = shard.dev
        self.n_nodes, self.n_expanded, self.levels = 1, 0, []
        self.comm_bytes, self.rounds, self.n_candidates = 0, 0, 0
        # pipeline=True stages round k + 1 on a side stream while round k is probed and committed.  Measured on 8 B200:
        # no gain (the NVLink-bound staging kernel and the latency-bound probe slow each other down), so it is off.
        self.pipeline, self.stage_ctas = bool(pipeline), int(stage_ctas)
        shard.connect(group)
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

    def _drain(self):
        side, main = self._side
        if side is not None:
            main.wait_stream(side)  # a staged round may still be in flight

    def _exchange(self, counts, k):
        """every rank learns how many records each rank staged for each owner (this all-gather is also the barrier
        behind which the peer stores of the round are complete), then the shard takes delivery."""
        G = self.world
        if G == 1:
            allc = [counts.tolist()]
        else:
            allc = torch.empty((G * (G + 1),), dtype=torch.int64, device=self.dev)
            dist.all_gather_into_tensor(allc, counts, group=self.group)
            allc = allc.view(G, G + 1).tolist()
        if any(row[G] for row in allc):
            raise RuntimeError("sharded bfs: a send region overflowed (hash imbalance above the 15% head room)")
        self.shard.deliver(allc, k)
        self.n_candidates += sum(allc[self.rank][:G])  # records this rank staged (children that differ from their parent)
        return self.shard.rec * (sum(allc[self.rank][:G]) - allc[self.rank][self.rank])

    def _lookup(self, gid):
        """(parent gid, action, total) of node `gid`, from whichever rank owns it"""
        s = self.shard
        hit = torch.nonzero(s.gid[: s.n_local()] == gid).flatten()
        rec = torch.zeros((3,), dtype=torch.int64, device=self.dev)
        if hit.numel():
            k = hit[0]
            rec = torch.stack([s.parent[k] + 1, s.action[k].to(torch.int64), s.node_total(k)])
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
        side = main = None
        if self.pipeline and G > 1 and getattr(s, "overlap", False) and dev.type == "cuda":
            # high priority: the staging kernel is NVLink-bound and needs few SM slots, but it needs them early
            side, main = torch.cuda.Stream(dev, priority=-1), torch.cuda.current_stream(dev)
        self._side = (side, main)
        lo, hi = 0, 1
        first = 0  # local index of the first node this rank discovered in the current frontier level
        while lo < hi:
            self.levels.append(hi - lo)
            self._tick(None)
            n_local = s.n_local()
            # group the local frontier by round (window of `chunk` consecutive global ids); inside a window the
            # order is free, so a narrow-key sort on the window index is enough (2-4 radix passes, not 8)
            gl = s.gid[first:n_local]
            nwin = (hi - lo + self.chunk - 1) // self.chunk
            if hasattr(s, "window_partition"):
                loc, gids, cuts = s.window_partition(gl, first, lo, self.chunk, nwin)
            else:
                widx = (gl - lo) // self.chunk
                widx, perm = torch.sort(widx.to(torch.int16 if nwin < 32768 else torch.int32))
                loc = perm + first
                gids = gl[perm]
                cuts = torch.searchsorted(widx, torch.arange(nwin + 1, device=dev, dtype=widx.dtype)).tolist()
            first = n_local
            self._tick("level_sort")
            rounds = [(off, min(self.chunk, hi - off), cuts[k], cuts[k + 1]) for k, off in enumerate(range(lo, hi, self.chunk))]
            k0 = self.rounds
            if side is not None:
                side.wait_stream(main)  # loc / gids of this level are ready

            def launch(i):
                off, C, a, b = rounds[i]
                if side is None:
                    return s.stage(loc[a:b], gids[a:b], off, k0 + i), None
                with torch.cuda.stream(side):
                    c = s.stage(loc[a:b], gids[a:b], off, k0 + i, stream=side.cuda_stream, ctas=self.stage_ctas)
                    c.record_stream(main)
                    return c, side.record_event()

            pending = launch(0)
            for i, (off, C, a, b) in enumerate(rounds):
                counts, ev = pending
                if ev is not None:
                    main.wait_event(ev)
                self._tick("expand_bucket")
                self.comm_bytes += self._exchange(counts, k0 + i)
                self.rounds += 1
                # every rank has now finished round i - 1 (its all-gather contribution comes after its commit), so
                # the buffers of round i + 1 are free: stage it next to the probe / commit of round i
                pending = launch(i + 1) if i + 1 < len(rounds) else None
                self._tick("all_to_all")
                mask, mins = s.probe(C)
                self._tick("probe")
                if G > 1:
                    if self.timers is not None:
                        dist.barrier(group=self.group)
                        self._tick("wait_slowest_rank")
                    dist.all_reduce(mask, group=self.group)
                    dist.all_reduce(mins, op=dist.ReduceOp.MIN, group=self.group)
                if hasattr(s, "mask_scan"):
                    excl, total = s.mask_scan(mask)
                else:
                    cnt = popcount_lut(mask)
                    excl = torch.cumsum(cnt, dim=0) - cnt
                    total = excl[-1:] + cnt[-1:]
                c_s, c_e, n_new = torch.cat([mins, total]).tolist()
                self._tick("reduce_scan")
                events = []
                if c_s < SENT:
                    events.append((c_s, 0, "solved"))
                if c_e < SENT:
                    events.append((c_e, 0, "error"))
                p_star = None
                if self.n_nodes + n_new >= self.budget:
                    # first parent after which len(tree_nodes) >= max_nodes_to_explore (breadth_first.py:91-95):
                    # inclusive count of parent p = excl[p + 1] (or the round total for the last parent)
                    idx = int(torch.searchsorted(excl, torch.tensor([self.budget - self.n_nodes], device=dev))[0])
                    p_star = max(idx - 1, 0)
                    events.append((12 * p_star + 11, 1, "budget"))
                if events:
                    pos, _, what = min(events)
                    if what == "solved":
                        self.n_expanded += pos // 12 + 1
                        p, aa = pos // 12, pos % 12
                        before = int(excl[p]) + int(popcount12(mask[p].to(torch.int64) & ((1 << aa) - 1)))
                        self.n_nodes += before
                        self._drain()
                        return True, self._path(off + p, aa)
                    if what == "error":
                        self._drain()
                        raise AssertionError("a relator became empty during the search: not a valid presentation")
                    self.n_expanded += p_star + 1
                    self.n_nodes += int(excl[p_star]) + int(popcount12(mask[p_star].to(torch.int64)))
                    self._drain()
                    return False, None
                s.commit(excl, mask, off, self.n_nodes)
                self._tick("commit")
                self.n_nodes += n_new
                self.n_expanded += C
            lo, hi = hi, self.n_nodes
        return False, None
