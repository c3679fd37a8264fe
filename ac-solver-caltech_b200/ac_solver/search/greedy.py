"""greedy_search with the reference's signature and return value (ac_solver/search/greedy.py:15-121)."""
import numpy as np

from ac_solver import _device
from acx import search as _search


def greedy_search(presentation, max_nodes_to_explore=10000, verbose=False, cyclically_reduce_after_moves=False):
    """Best-first search by total length.  Returns (is_search_successful, path) where path is the
    list [(-1, len(root)), (action, total_length), ...] -- for a failed search the reference's last
    expansion (greedy.py:121)."""
    row = np.array(presentation, dtype=np.int8)
    ok, path, n_seen, _ = _search.search_one(row, max_nodes_to_explore, cyclically_reduce_after_moves, True,
                                             str(_device.device()))
    if ok and verbose:
        print(f"Found trivial state after exploring {n_seen} nodes; path length {len(path) - 1}")
    if not ok and n_seen >= max_nodes_to_explore:
        print(f"Exiting search as number of explored nodes = {n_seen} has exceeded the limit {max_nodes_to_explore}")
    return ok, path
