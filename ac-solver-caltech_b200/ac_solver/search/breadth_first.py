"""bfs with the reference's signature and return value (ac_solver/search/breadth_first.py:15-97)."""
import numpy as np

from ac_solver import _device
from acx import bfs as _level
from acx import search as _search

# below this budget one warp replays the reference loop (acx_search_batch, latency-bound but launch-free);
# above it the level-synchronous engine expands whole frontiers per launch (identical results, see tests)
LEVEL_SYNCHRONOUS_FROM = 1 << 16


def bfs(presentation, max_nodes_to_explore=10000, verbose=False, cyclically_reduce_after_moves=False):
    """FIFO search.  Returns (True, path) or (False, None) (breadth_first.py:84-97)."""
    row = np.array(presentation, dtype=np.int8)
    if max_nodes_to_explore >= LEVEL_SYNCHRONOUS_FROM:
        stats = {}
        ok, path = _level.bfs_level(row, max_nodes_to_explore, cyclically_reduce_after_moves, str(_device.device()), stats=stats)
        n_seen = stats["n_seen"]
    else:
        ok, path, n_seen, _ = _search.search_one(row, max_nodes_to_explore, cyclically_reduce_after_moves, False,
                                                 str(_device.device()))
    if ok and verbose:
        print(f"Found trivial state after exploring {n_seen} nodes; path length {len(path) - 1}")
    if not ok and n_seen >= max_nodes_to_explore:
        print(f"Exiting search as number of explored nodes = {n_seen} has exceeded the limit {max_nodes_to_explore}")
    return ok, (path if ok else None)
