"""Shared generators for the parity tests (seeded; no reference import at run time)."""
import numpy as np

AK2 = np.array([1, 1, -2, -2, -2, 0, 0, 1, 2, 1, -2, -1, -2, 0], dtype=np.int8)
AK3 = [1, 1, 1, -2, -2, -2, -2], [1, 2, 1, -2, -1, -2]


def pad_row(r0, r1, m):
    row = np.zeros(2 * m, dtype=np.int8)
    row[: len(r0)] = r0
    row[m : m + len(r1)] = r1
    return row


def repad(row, m_new):
    """change_max_relator_length_of_presentation (utils.py:151-175) for one row."""
    m = len(row) // 2
    r0, r1 = row[:m][row[:m] != 0], row[m:][row[m:] != 0]
    return pad_row(r0, r1, m_new)


def reduced_word(rng, n):
    w = []
    while len(w) < n:
        c = int(rng.choice([1, -1, 2, -2]))
        if w and w[-1] == -c:
            continue
        w.append(c)
    return w


def dataset_rows(golden_dir):
    """The 1190 Miller-Schupp rows of all_presentations.txt (ragged widths)."""
    import os

    z = np.load(os.path.join(golden_dir, "dataset.npz"))
    widths, flat = z["widths"], z["flat"]
    off = np.concatenate([[0], np.cumsum(widths)])
    return [flat[off[k] : off[k + 1]].copy() for k in range(len(widths))]


def random_rows(rng, n, m, kinds=(0, 1, 2, 3, 4)):
    """Mixed test rows: 0 random letters (non-reduced), 1 reduced, 2 reduced near the cap,
    3 second relator related to the inverse of the first (deep cancellation), 4 weird alphabet /
    invalid padding."""
    rows = np.zeros((n, 2 * m), dtype=np.int8)
    for k in range(n):
        kind = kinds[k % len(kinds)]
        for h in range(2):
            if kind == 0:
                ln = int(rng.integers(1, m + 1))
                w = rng.choice([1, -1, 2, -2], size=ln)
            elif kind == 2:
                ln = int(rng.integers(max(1, m - 3), m + 1))
                w = reduced_word(rng, ln)
            elif kind == 4:
                ln = int(rng.integers(0, m + 1))
                w = rng.choice([1, -1, 2, -2, 3, -3, 4, -128, 127], size=ln)
            else:
                ln = int(rng.integers(1, m + 1))
                w = reduced_word(rng, ln)
            rows[k, h * m : h * m + ln] = w
            if kind == 4 and ln > 1 and rng.random() < 0.3:
                rows[k, h * m + int(rng.integers(0, ln))] = 0
        if kind == 3:
            l0 = int(np.count_nonzero(rows[k, :m]))
            inv = -rows[k, :l0][::-1]
            cut = int(rng.integers(0, l0 + 1))
            tail = reduced_word(rng, int(rng.integers(0, max(1, m - cut))))
            cand = (list(inv[:cut]) + tail)[:m] if rng.random() < 0.5 else (tail + list(inv[l0 - cut :]))[:m]
            if not cand:
                cand = [1]
            rows[k, m:] = 0
            rows[k, m : m + len(cand)] = cand
    return rows


def walk_states(oracle, roots, m, n, max_walk, seed, cyclical=True):
    """n states = random AC walks (<= max_walk uniform moves) from the given roots, via the oracle."""
    rng = np.random.default_rng(seed)
    states = np.stack([repad(roots[k % len(roots)], m) for k in range(n)])
    walk = rng.integers(0, max_walk + 1, size=n)
    for t in range(max_walk):
        live = walk > t
        if not live.any():
            break
        mv = rng.integers(0, 12, size=n).astype(np.uint8)
        nxt, _, st = oracle.acmove_batch(states, mv, cyclical=cyclical, threads=8)
        assert (st == 0).all()
        states[live] = nxt[live]
    return states
