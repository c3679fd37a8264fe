"""GPU parity: the CUDA path through the C ABI (acx.batch -> libacx.so) against the oracle and the
reference-generated golden vectors.  Bit-exact is the bar (integer/byte work)."""
import os

import numpy as np
import pytest

from helpers import dataset_rows, random_rows, repad, walk_states

pytestmark = pytest.mark.gpu


def _t(a, dev):
    import torch

    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


def _move(states, moves, cyc, dev, inplace=False):
    from acx import batch

    s = _t(states, dev)
    out, lens, st = batch.move_batch(s, _t(moves, dev), cyclical=cyc, out=s if inplace else None)
    return out.cpu().numpy(), lens.cpu().numpy(), st.cpu().numpy()


@pytest.mark.parametrize("m", [4, 7, 18, 36, 128])
def test_move_batch_golden(cuda, golden_dir, m):
    g = np.load(os.path.join(golden_dir, "moves_golden.npz"))
    rows, mv, cyc, out, lens, raised = (g[f"{k}_{m}"] for k in ("rows", "moves", "cyc", "out", "lens", "raised"))
    for c in (0, 1):
        sel = cyc == c
        o, l, st = _move(rows[sel], mv[sel], c, cuda)
        assert np.array_equal(o, out[sel])
        ok = raised[sel] == 0
        assert np.array_equal(l[ok].astype(np.int32), lens[sel][ok])
        assert np.array_equal(st != 0, raised[sel] != 0)


@pytest.mark.parametrize("m", [1, 2, 3, 5, 8, 12, 31, 32, 33, 36, 40, 63, 64, 65, 100, 127, 128])
@pytest.mark.parametrize("n", [1, 127, 128, 129, 5000])
def test_move_batch_fuzz(cuda, oracle, m, n):
    """ragged sizes (tile tails, single row), weird alphabets, invalid rows, deep cancellations"""
    rng = np.random.default_rng(1000 * m + n)
    rows = random_rows(rng, n, m)
    mv = rng.integers(0, 12, n).astype(np.uint8)
    for c in (0, 1):
        eo, el, es = oracle.acmove_batch(rows, mv, cyclical=c, threads=4)
        for inplace in (False, True):
            o, l, st = _move(rows, mv, c, cuda, inplace)
            assert np.array_equal(o, eo)
            assert np.array_equal(l.astype(np.int32), el)
            assert np.array_equal(st, es)


def test_move_batch_empty_and_unaligned(cuda, oracle):
    import torch
    from acx import batch

    e = torch.empty((0, 72), dtype=torch.int8, device=cuda)
    out, lens, st = batch.move_batch(e, torch.empty((0,), dtype=torch.uint8, device=cuda))
    assert out.shape == (0, 72)
    # a view that starts 72 bytes into an allocation is only 8-byte aligned: no bulk-copy path
    rng = np.random.default_rng(5)
    rows = random_rows(rng, 1001, 36, kinds=(1, 2, 3))
    mv = rng.integers(0, 12, 1000).astype(np.uint8)
    big = _t(rows, cuda)
    view = big[1:]
    assert view.data_ptr() % 16 != 0
    out, lens, st = batch.move_batch(view, _t(mv, cuda), cyclical=True)
    eo, el, es = oracle.acmove_batch(rows[1:], mv, cyclical=1)
    assert np.array_equal(out.cpu().numpy(), eo) and np.array_equal(st.cpu().numpy(), es)


@pytest.mark.parametrize("m,n", [(36, 1 << 20), (128, 1 << 18)])
def test_move_batch_walk_states_large(cuda, oracle, golden_dir, m, n):
    """BASELINE config #2 shape (1M rows, m=36, cyclical) and the m=128 width of config #5: random AC
    walks from the Miller-Schupp rows; full-size compare against the multi-threaded oracle."""
    roots = dataset_rows(golden_dir)
    base = walk_states(oracle, roots, m, 8192, 48, seed=0, cyclical=True)
    rng = np.random.default_rng(1)
    states = base[rng.integers(0, len(base), n)]
    mv = rng.integers(0, 12, n).astype(np.uint8)
    for c in (1, 0):
        eo, el, es = oracle.acmove_batch(states, mv, cyclical=c, threads=os.cpu_count() or 4)
        o, l, st = _move(states, mv, c, cuda)
        assert np.array_equal(o, eo)
        assert np.array_equal(l.astype(np.int32), el)
        assert np.array_equal(st, es)
        assert (es == 0).all()


def test_env_step_golden(cuda, golden_dir):
    """ACEnv.step trajectories recorded from the reference (obs, reward, done, truncated, lengths)."""
    import torch
    from acx import batch

    g = np.load(os.path.join(golden_dir, "env_golden.npz"))
    init, horizon, actions = g["init"], g["horizon"], g["actions"]
    for hz in np.unique(horizon):
        sel = np.where(horizon == hz)[0]
        buf = batch.EnvBuffers(_t(init[sel], cuda), horizon=int(hz), pool=_t(init[sel], cuda))
        idx = torch.arange(len(sel), dtype=torch.int32, device=cuda)
        prev_reset_state = None
        for t in range(actions.shape[0]):
            # without reset first: compare the post-move observation, then redo with reset
            snap = buf.state.clone(), buf.steps.clone()
            batch.env_step(buf, _t(actions[t, sel], cuda))
            assert np.array_equal(buf.state.cpu().numpy(), g["obs"][t, sel])
            assert np.array_equal(buf.reward.cpu().numpy(), g["reward"][t, sel])
            assert np.array_equal(buf.done.cpu().numpy(), g["done"][t, sel])
            assert np.array_equal(buf.truncated.cpu().numpy(), g["truncated"][t, sel])
            assert np.array_equal(buf.lens.cpu().numpy(), g["lens"][t, sel])
            assert (buf.status.cpu().numpy() == 0).all()
            buf.set_state(snap[0], snap[1])
            batch.env_step(buf, _t(actions[t, sel], cuda), next_idx=idx)  # fused reset of finished envs
            fin = (g["done"][t, sel] | g["truncated"][t, sel]).astype(bool)
            exp = g["obs"][t, sel].copy()
            exp[fin] = init[sel][fin]
            assert np.array_equal(buf.state.cpu().numpy(), exp)
            assert (buf.steps.cpu().numpy()[fin] == 0).all()
        del prev_reset_state


@pytest.mark.parametrize("n", [1, 130, 70001])
def test_env_step_vs_oracle(cuda, oracle, golden_dir, n):
    """random trajectories with auto-reset from a pool, short horizon so that truncation fires"""
    import torch
    from acx import batch

    m, horizon = 36, 9
    roots = dataset_rows(golden_dir)
    pool = np.stack([repad(r, m) for r in roots[:64]])
    rng = np.random.default_rng(n)
    start = rng.integers(0, len(pool), n)
    state = pool[start].copy()
    steps = np.zeros(n, dtype=np.int32)
    buf = batch.EnvBuffers(_t(state, cuda), horizon=horizon, pool=_t(pool, cuda))
    for t in range(40):
        act = rng.integers(0, 12, n).astype(np.uint8)
        nxt = rng.integers(0, len(pool), n).astype(np.int32)
        rew, done, trunc, lens, st = oracle.env_step_batch(state, steps, act, horizon, pool=pool, next_idx=nxt, threads=4)
        batch.env_step(buf, _t(act, cuda), next_idx=_t(nxt, cuda))
        assert np.array_equal(buf.state.cpu().numpy(), state)
        assert np.array_equal(buf.steps.cpu().numpy(), steps)
        assert np.array_equal(buf.reward.cpu().numpy(), rew)
        assert np.array_equal(buf.reward_clipped.cpu().numpy(), np.clip(rew, -10, 1000).astype(np.float32))
        assert np.array_equal(buf.done.cpu().numpy(), done)
        assert np.array_equal(buf.truncated.cpu().numpy(), trunc)
        assert np.array_equal(buf.lens.cpu().numpy(), lens)
        assert np.array_equal(buf.status.cpu().numpy(), st)


@pytest.mark.parametrize("m", [7, 18, 36, 128])
@pytest.mark.parametrize("f", [1, 7, 8, 9, 3000])
def test_expand12_vs_oracle(cuda, oracle, m, f):
    from acx import batch

    rng = np.random.default_rng(m * 7 + f)
    rows = random_rows(rng, f, m)
    for c in (0, 1):
        ec, el, es = oracle.expand12_batch(rows, cyclical=c, threads=4)
        ch, l, st = batch.expand12(_t(rows, cuda), cyclical=c)
        assert np.array_equal(ch.cpu().numpy(), ec)
        assert np.array_equal(l.cpu().numpy().astype(np.int32), el)
        assert np.array_equal(st.cpu().numpy(), es)


def test_env_step_certified_vs_checked(cuda, oracle, golden_dir):
    """the certified-row kernel and the checked kernel give identical trajectories (and match the oracle);
    an uncertifiable pool row switches the buffers to the checked kernel"""
    import torch
    from acx import batch

    m, horizon, n = 36, 25, 40000
    roots = dataset_rows(golden_dir)
    pool = np.stack([repad(r, m) for r in roots[:300]])
    states = walk_states(oracle, roots, m, 4096, 30, seed=11, cyclical=True)
    rng = np.random.default_rng(2)
    state = states[rng.integers(0, len(states), n)].copy()
    steps = np.zeros(n, dtype=np.int32)
    a = batch.EnvBuffers(_t(state, cuda), horizon=horizon, pool=_t(pool, cuda))
    b = batch.EnvBuffers(_t(state, cuda), horizon=horizon, pool=_t(pool, cuda))
    assert a.certified and b.certified
    b.certified = False  # force the checked kernel
    for t in range(60):
        act = rng.integers(0, 12, n).astype(np.uint8)
        nxt = rng.integers(0, len(pool), n).astype(np.int32)
        rew, done, trunc, lens, st = oracle.env_step_batch(state, steps, act, horizon, pool=pool, next_idx=nxt, threads=8)
        for buf in (a, b):
            batch.env_step(buf, _t(act, cuda), next_idx=_t(nxt, cuda))
            assert np.array_equal(buf.state.cpu().numpy(), state)
            assert np.array_equal(buf.step_out.cpu().numpy()[:, :4].copy().view(np.int32)[:, 0], rew)
            assert np.array_equal(buf.done.cpu().numpy(), done)
            assert np.array_equal(buf.truncated.cpu().numpy(), trunc)
            assert np.array_equal(buf.lens.cpu().numpy(), lens)
            assert np.array_equal(buf.steps.cpu().numpy(), steps)
    bad_pool = pool.copy()
    bad_pool[5, :4] = [1, 2, -2, 1]  # not freely reduced
    c = batch.EnvBuffers(_t(state, cuda), horizon=horizon, pool=_t(bad_pool, cuda))
    assert not c.certified
