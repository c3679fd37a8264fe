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
        self.n_local = 0
        self.table = {}  # row bytes -> ("node", idx) | ("cand", j)

    def seed(self, root):
        self.nodes[0] = torch.from_numpy(root)
        self.total[0] = int(np.count_nonzero(root))
        self.gid[0] = 0
        self.table[root.tobytes()] = ("node", 0)
        self.n_local = 1

    def hashes(self, rows):
        a = rows.numpy().view(np.uint8).astype(np.uint64)
        h = np.full(a.shape[0], 1469598103934665603, dtype=np.uint64)
        with np.errstate(over="ignore"):
            for k in range(a.shape[1]):
                h = (h ^ a[:, k]) * np.uint64(1099511628211)
            h ^= h >> np.uint64(29)
        return torch.from_numpy(h.view(np.int64).copy())

    def expand(self, rows):
        f = rows.shape[0]
        if f == 0:
            return (torch.empty((0, self.rowb), dtype=torch.int8), torch.empty((0, 2), dtype=torch.uint8),
                    torch.empty((0,), dtype=torch.uint8))
        ch, ln, st = ac_oracle.expand12_batch(rows.numpy(), cyclical=self.cyc)
        return (torch.from_numpy(ch.reshape(f * 12, self.rowb)), torch.from_numpy(ln.reshape(f * 12, 2).astype(np.uint8)),
                torch.from_numpy(st.reshape(f * 12)))

    def probe(self, cand, lens, status, order):
        n = cand.shape[0]
        slot = torch.full((n,), -1, dtype=torch.int64)
        mins = [SENT, SENT]
        c, l, s, o = cand.numpy(), lens.numpy(), status.numpy(), order.numpy()
        keys = [None] * n
        for j in range(n):
            if s[j] != 0:
                mins[1] = min(mins[1], int(o[j]))
                slot[j] = -2
            elif int(l[j, 0]) + int(l[j, 1]) == 2:
                mins[0] = min(mins[0], int(o[j]))
                slot[j] = -2
            else:
                key = c[j].tobytes()
                keys[j] = key
                e = self.table.get(key)
                if e is None or (e[0] == "cand" and o[e[1]] > o[j]):
                    self.table[key] = ("cand", j)
                    slot[j] = 0
                elif e[0] == "cand":
                    slot[j] = 0
        is_new = torch.zeros((n,), dtype=torch.uint8)
        for j in range(n):
            if keys[j] is not None and self.table[keys[j]] == ("cand", j):
                is_new[j] = 1
        self._keys = keys
        return slot, is_new, torch.tensor(mins, dtype=torch.int64)

    def commit(self, cand, lens, slot, is_new, excl, order, parent_base):
        for j in range(cand.shape[0]):
            if is_new[j]:
                k = self.n_local + int(excl[j])
                self.nodes[k] = cand[j]
                self.parent[k] = parent_base + int(order[j]) // 12
                self.action[k] = int(order[j]) % 12
                self.total[k] = int(lens[j, 0]) + int(lens[j, 1])
                self.table[self._keys[j]] = ("node", k)
