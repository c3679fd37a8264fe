"""TEST INFRASTRUCTURE: a CPU stand-in for acx.bfs_sharded.GpuShard built on the oracle, so that the
host-side logic of the sharded bfs (partitioning, all-to-all-v, mask all-reduce, global ids, budget cut)
can run under gloo with world_size 2 on a machine without GPUs.  Never used by the product."""
import numpy as np
import torch

import ac_oracle

SENT = 1 << 62


class OracleShard:
    def __init__(self, m, cyclical, cap_local):
        self.m, self.rowb, self.cyc = m, 2 * m, bool(cyclical)
        self.dev = torch.device("cpu")
        self.nodes = torch.zeros((cap_local, self.rowb), dtype=torch.int8)
        self.gid = torch.full((cap_local,), -1, dtype=torch.int64)
        self.parent = torch.full((cap_local,), -1, dtype=torch.int64)
        self.action = torch.zeros((cap_local,), dtype=torch.uint8)
        self.total = torch.zeros((cap_local,), dtype=torch.int16)
        self._n = 0
        self.rec = ((self.rowb + 7) // 8 * 8 + 16 + 15) // 16 * 16
        self.table = {}  # row bytes -> ("node", idx) | ("cand", j)

    def seed(self, root):
        self.nodes[0] = torch.from_numpy(root)
        self.total[0] = int(np.count_nonzero(root))
        self.gid[0] = 0
        self.table[root.tobytes()] = ("node", 0)
        self._n = 1

    def n_local(self):
        return self._n

    def _hash(self, a):
        a = a.view(np.uint8).astype(np.uint64)
        h = np.full(a.shape[0], 1469598103934665603, dtype=np.uint64)
        with np.errstate(over="ignore"):
            for k in range(a.shape[1]):
                h = (h ^ a[:, k]) * np.uint64(1099511628211)
            h ^= h >> np.uint64(29)
        return h

    def owner_of_row(self, row, world):
        return int(((self._hash(np.ascontiguousarray(row[None, :])) >> np.uint64(20)) & np.uint64(0x7FFFFFFF))[0] % world)

    def connect(self, group=None):
        import torch.distributed as dist

        self._group, self.world, self.rank = group, dist.get_world_size(group), dist.get_rank(group)

    def stage(self, loc, gids, ord_base, k):
        world = self.world
        rows, parent_ord = self.nodes[loc], gids - ord_base
        f = rows.shape[0]
        n = 12 * f
        off = (self.rowb + 7) // 8 * 8
        send = np.zeros((n, self.rec), dtype=np.uint8)
        counts = np.zeros(world + 1, dtype=np.int64)
        if f:
            ch, ln, st = ac_oracle.expand12_batch(rows.numpy(), cyclical=self.cyc)
            ch = np.ascontiguousarray(ch.reshape(n, self.rowb))
            own = ((self._hash(ch) >> np.uint64(20)) & np.uint64(0x7FFFFFFF)) % np.uint64(world)
            order = (parent_ord.numpy()[:, None] * 12 + np.arange(12)[None, :]).reshape(-1).astype(np.int64)
            rec = np.zeros((n, self.rec), dtype=np.uint8)
            rec[:, :self.rowb] = ch.view(np.uint8)
            rec[:, off:off + 8] = order.view(np.uint8).reshape(n, 8)
            rec[:, off + 8:off + 10] = ln.reshape(n, 2).astype(np.uint8)
            rec[:, off + 10] = st.reshape(n)
            # any order inside an owner group is legal: scramble it to prove the receiver does not depend on it
            perm = np.lexsort((np.random.default_rng(n).permutation(n), own))
            send = rec[perm]
            counts[:world] = np.bincount(own.astype(np.int64), minlength=world)
        self._send = torch.from_numpy(send)
        return torch.from_numpy(counts)

    def deliver(self, allc, k):
        import torch.distributed as dist

        G = self.world
        scl, rcl = allc[self.rank][:G], [allc[g][self.rank] for g in range(G)]
        recv = torch.empty((sum(rcl), self.rec), dtype=torch.uint8)
        dist.all_to_all_single(recv, self._send, output_split_sizes=rcl, input_split_sizes=scl, group=self._group)
        self._recv = recv

    def _fields(self, recv):
        r = recv.numpy()
        off = (self.rowb + 7) // 8 * 8
        rows = np.ascontiguousarray(r[:, :self.rowb]).view(np.int8)
        order = np.ascontiguousarray(r[:, off:off + 8]).view(np.int64).reshape(-1)
        return rows, order, r[:, off + 8:off + 10], r[:, off + 10]

    def probe(self, n_parents):
        recv = self._recv
        n = recv.shape[0]
        c, o, l, s = self._fields(recv)
        slot = torch.full((n,), -1, dtype=torch.int64)
        mins = [SENT, SENT]
        keys = [None] * n
        for j in range(n):
            if s[j] != 0:
                mins[1] = min(mins[1], int(o[j]))
            elif int(l[j, 0]) + int(l[j, 1]) == 2:
                mins[0] = min(mins[0], int(o[j]))
            else:
                key = c[j].tobytes()
                keys[j] = key
                e = self.table.get(key)
                if e is None or (e[0] == "cand" and o[e[1]] > o[j]):
                    self.table[key] = ("cand", j)
        mask = torch.zeros((n_parents,), dtype=torch.int32)
        for j in range(n):
            if keys[j] is not None and self.table[keys[j]] == ("cand", j):
                slot[j] = 0
                mask[int(o[j]) // 12] |= 1 << (int(o[j]) % 12)
        self._keys, self._slot = keys, slot
        return mask, torch.tensor(mins, dtype=torch.int64)

    def commit(self, excl_par, gmask, parent_base, gid_base):
        recv, slot = self._recv, self._slot
        c, o, l, s = self._fields(recv)
        for j in range(recv.shape[0]):
            if slot[j] >= 0:
                k = self._n
                self._n += 1
                par, act = int(o[j]) // 12, int(o[j]) % 12
                self.nodes[k] = torch.from_numpy(c[j].copy())
                self.gid[k] = gid_base + int(excl_par[par]) + bin(int(gmask[par]) & ((1 << act) - 1)).count("1")
                self.parent[k] = parent_base + par
                self.action[k] = act
                self.total[k] = int(l[j, 0]) + int(l[j, 1])
                self.table[self._keys[j]] = ("node", k)
