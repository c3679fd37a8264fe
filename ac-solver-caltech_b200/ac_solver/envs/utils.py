"""ac_solver.envs.utils of the reference (ac_solver/envs/utils.py), arithmetic on the GPU engine.

Predicates and padding helpers are host-side array checks, as in the reference; the reductions
(simplify_relator, simplify_presentation) run in libacx.so (acx_rows_primitive).
"""
import numpy as np

from ac_solver import _device


def is_array_valid_presentation(array):
    """utils.py:13-54: even length, both halves non-empty, zeros only on the right of each half."""
    assert isinstance(array, (list, np.ndarray)), f"array must be a list or a numpy array, got {type(array)}"
    a = np.asarray(array)
    if a.ndim != 1 or len(a) % 2 != 0 or len(a) == 0:
        return False
    m = len(a) // 2
    for r in (a[:m], a[m:]):
        n = int(np.count_nonzero(r))
        if n == 0 or np.any(r[:n] == 0):
            return False
    return True


def is_presentation_trivial(presentation):
    """utils.py:57-87: valid, both relators of length one, on distinct generators."""
    a = np.asarray(presentation)
    if not is_array_valid_presentation(a):
        return False
    m = len(a) // 2
    if np.count_nonzero(a[:m]) != 1 or np.count_nonzero(a[m:]) != 1:
        return False
    return sorted((abs(int(a[0])), abs(int(a[m])))) == [1, 2]


def generate_trivial_states(max_relator_length):
    """utils.py:91-114: the eight trivial states, in the reference's order."""
    states = []
    for i in [1, 2]:
        for sign1 in [-1, 1]:
            for sign2 in [-1, 1]:
                states += [[sign1 * i] + [0] * (max_relator_length - 1) + [sign2 * (3 - i)] + [0] * (max_relator_length - 1)]
    return np.array(states)


def convert_relators_to_presentation(relator1, relator2, max_relator_length):
    """utils.py:117-148"""
    assert 0 not in relator1 and 0 not in relator2, "relator1 and relator2 must not be padded with zeros."
    assert max_relator_length >= max(len(relator1), len(relator2)), \
        "max_relator_length must be greater than or equal to the lengths of relator1 and rel2."
    assert isinstance(relator1, (list, np.ndarray)) and isinstance(relator2, (list, np.ndarray)), \
        f"got types {type(relator1)} for relator1 and {type(relator2)} for relator2"
    r1, r2 = list(relator1), list(relator2)
    return np.array(r1 + [0] * (max_relator_length - len(r1)) + r2 + [0] * (max_relator_length - len(r2)), dtype=np.int8)


def change_max_relator_length_of_presentation(presentation, new_max_length):
    """utils.py:151-175"""
    p = np.asarray(presentation)
    old = len(p) // 2
    n1, n2 = int(np.count_nonzero(p[:old])), int(np.count_nonzero(p[old:]))
    return convert_relators_to_presentation(p[:n1], p[old:old + n2], new_max_length)


def simplify_relator(relator, max_relator_length, cyclical=False, padded=True):
    """utils.py:178-243: free (and optionally cyclic) reduction of one relator."""
    assert isinstance(relator, np.ndarray), "expect relator to be a numpy array"
    r = np.asarray(relator)
    n_in = int(np.count_nonzero(r))
    if n_in == 0:
        out = np.zeros(max_relator_length if padded else len(r), dtype=r.dtype)
        return (out if padded else r.copy()), 0
    width = max(len(r), max_relator_length, 1)
    row = np.zeros(2 * width, dtype=np.int8)
    row[: len(r)] = r
    row[width] = 1  # a one-letter companion relator keeps the row a valid presentation
    mode = _device.MODE_NO_MOVE | (_device.MODE_CYCLICAL if cyclical else 0)
    out, lens, st = _device.rows_primitive(row, width, 0, mode)
    assert st == 0, "expect all zeros to be at the right end"
    n = lens[0]
    assert max_relator_length >= n, "Increase max length! Word length is bigger than maximum allowed length."
    if padded:
        res = out[:max_relator_length].astype(r.dtype)
    else:  # np.delete keeps the input's own padding, minus the removed letters (utils.py:214, :230)
        res = out[: len(r) - (n_in - n)].astype(r.dtype)
    return res, n


def simplify_presentation(presentation, max_relator_length, lengths_of_words, cyclical=True):
    """utils.py:246-283"""
    p = np.array(presentation)
    assert is_array_valid_presentation(p), f"{presentation} is not a valid presentation. Expect all zeros to be padded to the right."
    mode = _device.MODE_NO_MOVE | (_device.MODE_CYCLICAL if cyclical else 0)
    out, lens, st = _device.rows_primitive(p, max_relator_length, 0, mode)
    assert st == 0, f"{presentation} is not a valid presentation."
    return out.astype(p.dtype), lens
