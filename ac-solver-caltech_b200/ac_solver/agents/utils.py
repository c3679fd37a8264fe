"""Dataset loader of the env-vector layer (reference: ac_solver/agents/utils.py:10-34).

The reference reads `all_presentations.txt` / `greedy_solved_presentations.txt` from its package
data; here the same 1190 rows (and the index of the 533 greedy-solved ones) ship as
ac_solver/data/miller_schupp.npz, written from those files by tests/golden/make_golden.py."""
import os

import numpy as np

_DATA = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "data", "miller_schupp.npz")


def load_initial_states_from_text_file(states_type):
    assert states_type in ["solved", "all"], "states_type must be 'solved' or 'all'"
    z = np.load(_DATA)
    widths, flat = z["widths"], z["flat"]
    off = np.concatenate([[0], np.cumsum(widths)])
    rows = [flat[off[k]:off[k + 1]].tolist() for k in range(len(widths))]
    if states_type == "solved":
        rows = [rows[k] for k in z["greedy_solved_idx"]]
    print(f"Loaded {len(rows)} presentations from "
          f"{'greedy_solved' if states_type == 'solved' else 'all'}_presentations.txt.")
    return rows
