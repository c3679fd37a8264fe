"""CPU: host-side logic of the device-driven plane bfs (acx.pbfs) that needs no kernel."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "ac-solver-caltech_b200"))


def test_block_places_covers_every_node_of_the_block():
    """random levels: rounds with random numbers of new nodes (some with none on this rank, some with holes between the
    nodes); the place span chosen for a block of windows holds every local node whose id falls into the block, and is
    not the whole level when the block is a small part of it"""
    from acx.pbfs import block_places

    rng = np.random.default_rng(11)
    for trial in range(200):
        n_rounds = int(rng.integers(1, 40))
        first = int(rng.integers(0, 1000))
        gid0 = int(rng.integers(0, 10 ** 6))
        place, gid, rows, node_gid, node_place = first, gid0, [], [], []
        for r in range(n_rounds):
            n_new = int(rng.integers(0, 500))                   # new nodes of the round over all ranks
            local = np.sort(rng.choice(n_new, size=int(rng.integers(0, n_new + 1)), replace=False)) if n_new else np.zeros(0, int)
            holes = int(rng.integers(0, 20)) if len(local) else 0
            span = len(local) + holes                           # places the round reserved on this rank
            where = np.sort(rng.choice(span, size=len(local), replace=False)) if span else np.zeros(0, int)
            rows.append((place, place + span, gid, n_new))
            node_gid += (gid + local).tolist()
            node_place += (place + where).tolist()
            place += span
            gid += n_new
        n_local = place
        log = np.array(rows, dtype=np.int64)
        node_gid, node_place = np.array(node_gid, dtype=np.int64), np.array(node_place, dtype=np.int64)
        chunk = int(rng.integers(1, 300))
        for _ in range(10):
            blo = gid0 + chunk * int(rng.integers(0, max(1, (gid - gid0) // chunk + 1)))
            bhi = blo + chunk * int(rng.integers(1, 5))
            a, b = block_places(log, blo, bhi, first, n_local)
            assert first <= a <= b <= n_local
            inside = (node_gid >= blo) & (node_gid < bhi)
            assert ((node_place[inside] >= a) & (node_place[inside] < b)).all()
            # tight: the span is made of whole rounds that reach into the block
            touched = [(p0, p1) for p0, p1, g0, nn in rows if g0 < bhi and g0 + nn > blo and p1 > p0]
            assert (a, b) == ((min(p for p, _ in touched), max(q for _, q in touched)) if touched else (first, first))
        assert block_places(None, 0, 10, first, n_local) == (first, n_local)
