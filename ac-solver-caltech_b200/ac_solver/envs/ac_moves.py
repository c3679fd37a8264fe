"""ac_solver.envs.ac_moves of the reference (ac_solver/envs/ac_moves.py) on the GPU engine."""
import numpy as np

from ac_solver import _device

# (kind, i, sign or generator) -> id of the ACMove that applies that primitive (ac_moves.py:192-206)
_CONCAT_ID = {(1, 1): 0, (0, -1): 1, (1, -1): 2, (0, 1): 3}
_CONJ_ID = {(1, -1): 4, (0, -2): 5, (1, -2): 6, (0, 1): 7, (1, 1): 8, (0, 2): 9, (1, 2): 10, (0, -1): 11}


def _raw(presentation, max_relator_length, move_id, lengths):
    p = np.asarray(presentation)
    out, lens, st = _device.rows_primitive(p, max_relator_length, move_id, _device.MODE_RAW_MOVE)
    assert st == 0, "expect a valid presentation"
    if isinstance(lengths, list) and len(lengths) == 2:  # the reference mutates the caller's list (ac_moves.py:65)
        lengths[0], lengths[1] = lens[0], lens[1]
    return out.astype(p.dtype), lens


def concatenate_relators(presentation, max_relator_length, i, j, sign, lengths):
    """ac_moves.py:4-76: r_i <- r_i r_j^{sign}, junction cancellation only; unchanged if too long."""
    assert all(arg in [0, 1] for arg in [i, j]) and i == 1 - j, f"expect i and j to be 0 or 1 and i != j; got i = {i}, j = {j}"
    assert sign in [1, -1], f"expect sign to be +1 or -1, received {sign}"
    return _raw(presentation, max_relator_length, _CONCAT_ID[(i, sign)], lengths)


def conjugate(presentation, max_relator_length, i, j, sign, lengths):
    """ac_moves.py:79-156: r_i <- g r_i g^-1 with g = sign * j; one letter tested at each end."""
    assert i in [0, 1], f"expect i to be 0 or 1. i = {i}"
    assert j in [1, 2], f"expect j to be 1 or 2. j = {j}"
    assert sign in [1, -1], f"expect sign to be +1 or -1. sign = {sign}"
    return _raw(presentation, max_relator_length, _CONJ_ID[(i, sign * j)], lengths)


def ACMove(move_id, presentation, max_relator_length, lengths, cyclical=True):
    """ac_moves.py:159-231: one of the 12 AC moves followed by simplify_presentation.

    The input array is not modified; the result has the input's dtype; an AssertionError is raised
    where the reference raises one (invalid presentation / a relator cancelling to nothing)."""
    assert move_id in range(0, 12), f"Expect n to be in range 0-11 (both inclusive). Got {move_id}"
    p = np.asarray(presentation)
    mode = _device.MODE_CYCLICAL if cyclical else 0
    out, lens, st = _device.rows_primitive(p, max_relator_length, move_id, mode)
    assert st == 0, f"{presentation} is not a valid presentation. Expect all zeros to be padded to the right."
    return out.astype(p.dtype), lens
