"""The product's row arithmetic (bit planes + byte slow path), compiled for the host, against the
oracle and the reference-generated golden vectors.  CPU only; the same code runs on the GPU."""
import os

import numpy as np
import pytest

from helpers import random_rows
from host_harness import harness


def _variants(m):
    """0/1: generic multi-word planes (dynamic / static); 2/3: 64-bit planes (dynamic / static 36)"""
    v = [0]
    if m in (36, 128):
        v.append(1)
    if m <= 64:
        v.append(2)
    if m == 36:
        v.append(3)
    return v


@pytest.mark.parametrize("m", [4, 7, 18, 36, 128])
def test_host_rows_golden(golden_dir, m):
    g = np.load(os.path.join(golden_dir, "moves_golden.npz"))
    rows, mv, cyc, out, lens, raised = (g[f"{k}_{m}"] for k in ("rows", "moves", "cyc", "out", "lens", "raised"))
    for c in (0, 1):
        sel = cyc == c
        for variant in _variants(m):
            for inplace in (0, 1):
                o, l, st = harness.move_batch(rows[sel], mv[sel], c, variant, inplace)
                assert np.array_equal(o, out[sel])
                ok = raised[sel] == 0
                assert np.array_equal(l[ok].astype(np.int32), lens[sel][ok])
                assert np.array_equal(st != 0, raised[sel] != 0)


@pytest.mark.parametrize("m", [1, 2, 3, 5, 8, 12, 31, 32, 33, 36, 40, 63, 64, 65, 100, 127, 128])
def test_host_rows_fuzz(oracle, m):
    rng = np.random.default_rng(100 + m)
    rows = random_rows(rng, 4000, m)
    mv = rng.integers(0, 12, len(rows)).astype(np.uint8)
    for c in (0, 1):
        eo, el, es = oracle.acmove_batch(rows, mv, cyclical=c, threads=4)
        for variant in _variants(m):
            o, l, st = harness.move_batch(rows, mv, c, variant, 1)
            assert np.array_equal(o, eo)
            assert np.array_equal(l.astype(np.int32), el)
            assert np.array_equal(st, es)


def test_host_certified_rows(oracle, golden_dir):
    """the certified-row fast path of the env kernel (rotation conjugation, 2-plane pack) on AC-walk
    states and on the reduced golden rows; uncertified rows fall back to the checked path"""
    from helpers import dataset_rows, walk_states

    roots = dataset_rows(golden_dir)
    states = walk_states(oracle, roots, 36, 6000, 40, seed=5, cyclical=True)
    rng = np.random.default_rng(9)
    # short relators (length 1..3) exercise the rotation edge cases
    short = random_rows(rng, 3000, 36, kinds=(1,))
    short[:, 3:36] = 0
    short[:, 36 + 2:] = 0
    mixed = random_rows(rng, 3000, 36)
    rows = np.concatenate([states, short, mixed])
    for rep in range(3):
        mv = rng.integers(0, 12, len(rows)).astype(np.uint8)
        eo, el, es = oracle.acmove_batch(rows, mv, cyclical=1, threads=4)
        for variant in (4, 5, 6):  # plane path; class-split byte/plane path with 0x00 / 0xFF in the stage padding
            o, l, st = harness.move_batch(rows, mv, 1, variant, 1)
            assert np.array_equal(o, eo)
            assert np.array_equal(l.astype(np.int32), el)
            assert np.array_equal(st, es)


@pytest.mark.parametrize("cyc", [0, 1])
def test_host_register_conjugations(oracle, golden_dir, cyc):
    """the register-resident conjugations of the frontier expansion (conj_row_words of acx_envrow.cuh, both forms):
    relators of every length up to the guard (n + 2 > m leaves the word alone), both cancellations"""
    from helpers import dataset_rows, walk_states

    roots = dataset_rows(golden_dir)
    states = walk_states(oracle, roots, 36, 4000, 40, seed=11, cyclical=bool(cyc))
    rng = np.random.default_rng(13)
    full = random_rows(rng, 4000, 36, kinds=(1,))  # freely reduced, every length 1..36
    short = random_rows(rng, 2000, 36, kinds=(1,))
    short[:, 3:36] = 0
    short[:, 36 + 2:] = 0
    mixed = random_rows(rng, 2000, 36)
    rows = np.concatenate([states, full, short, mixed])
    for rep in range(3):
        mv = rng.integers(4, 12, len(rows)).astype(np.uint8)
        if rep == 2:
            mv = rng.integers(0, 12, len(rows)).astype(np.uint8)
        eo, el, es = oracle.acmove_batch(rows, mv, cyclical=cyc, threads=4)
        o, l, st = harness.move_batch(rows, mv, cyc, 7, 1)
        assert np.array_equal(o, eo)
        assert np.array_equal(l.astype(np.int32), el)
        assert np.array_equal(st, es)
