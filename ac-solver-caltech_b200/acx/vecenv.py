"""VecACEnv -- N lockstep ACEnv instances resident on one GPU (the env-vector layer).

Replaces `gym.vector.SyncVectorEnv([make_env(...)] * num_envs)` of the reference
(ac_solver/agents/environment.py:96-101), whose `step` is a sequential Python loop over envs, and
the per-env reset the PPO rollout applies to finished envs (agents/training.py:351-352):

    envs = VecACEnv(initial_states, horizon_length=2000)         # initial_states: [P, 2m] pool
    obs, _ = envs.reset()
    obs, reward, done, truncated, infos = envs.step(actions)      # numpy in, numpy out (host API)
    obs_t, rew_t, done_t, trunc_t = envs.step_device(actions_t)   # torch CUDA tensors, no host hop

State, lengths and step counters live in HBM; one `acx_env_step` launch advances every env,
computes reward/done/truncated and re-seats finished envs on `pool[next_state_index[e]]`.
Rewards are clipped to [min_rew, max_rew] as `TransformReward(np.clip)` does (environment.py:48-52)
when clip_rewards is set.

Action log (`ACEnv.actions`, ac_env.py:92,110): with track_actions the step kernel keeps the last
`horizon_length` action vectors in a device ring and drops (env, episode length, final state) of every
finished env into a side buffer; `step` turns those into the `infos` gymnasium 0.28.1's SyncVectorEnv
hands to the rollout loop (agents/training.py:273-280):
    infos["final_info"][i] == {"actions": [...]}   for a solved env (ACEnv returns {} when only truncated)
    infos["final_observation"][i]                   the state BEFORE the reset
    infos["_final_info"][i], infos["_final_observation"][i]   masks
and the gymnasium >= 1.0 spelling infos["actions"][i] / infos["_actions"][i].
"""
import numpy as np
import torch

from . import batch


class _EnvHandle:
    """`envs.envs[i]`: what the PPO loop pokes at for a single env (agents/training.py:352,362-363)."""

    def __init__(self, vec, index):
        self._vec, self._i = vec, index

    def reset(self, *, seed=None, options=None):
        v, i = self._vec, self._i
        if options and "starting_state" in options:
            start = options["starting_state"]
            if isinstance(start, torch.Tensor):
                row = start.detach().to(device=v.device, dtype=torch.int8)
            else:
                row = torch.from_numpy(np.ascontiguousarray(np.asarray(start), dtype=np.int8)).to(v.device)
        else:
            row = v.pool[int(v.start_index[i])]
        v.buf.set_row(i, row, steps=0)
        return v.buf.state[i].cpu().numpy(), {}

    @property
    def unwrapped(self):
        return self

    @property
    def max_reward(self):
        return self._vec.max_reward

    @property
    def max_relator_length(self):
        return self._vec.max_relator_length

    @property
    def horizon_length(self):
        return self._vec.horizon_length

    @property
    def state(self):
        return self._vec.buf.state[self._i].cpu().numpy()

    @property
    def lengths(self):
        w = int(self._vec.buf.meta[self._i]) & 0xFFFFFFFF
        return [(w >> 16) & 0xFF, w >> 24]

    @property
    def count_steps(self):
        return int(self._vec.buf.meta[self._i]) & 0xFFFF

    @property
    def actions(self):
        """actions of the running episode (ACEnv.actions, ac_env.py:88,92)"""
        v = self._vec
        if v.buf.ring is None:
            raise AttributeError("VecACEnv was built with track_actions=False: no action log is kept")
        n = self.count_steps
        t = v.buf.t
        rows = (torch.arange(t - n, t, device=v.device) % v.buf.horizon).long()
        return [int(a) for a in v.buf.ring[rows, self._i].cpu().numpy()]


class _LazyHandles:
    """envs.envs[i] for very large batches without materialising N Python objects."""

    def __init__(self, vec):
        self._vec = vec

    def __getitem__(self, i):
        return _EnvHandle(self._vec, i)

    def __len__(self):
        return self._vec.num_envs


class _Space:
    def __init__(self, shape=None, n=None, dtype=None):
        self.shape, self.n, self.dtype = shape, n, dtype


class _Sparse:
    """infos[key] of a vector env, kept sparse: index i -> value, None for every env not in the step's event set
    (gymnasium materialises an object array of length num_envs per key)."""

    def __init__(self, n, items, default=None):
        self._n, self._items, self._default = n, items, default

    def __getitem__(self, i):
        i = int(i)
        if not -self._n <= i < self._n:
            raise IndexError(i)
        return self._items.get(i % self._n, self._default)

    def __len__(self):
        return self._n

    def __iter__(self):
        return (self[i] for i in range(self._n))

    def keys(self):
        return self._items.keys()


class VecACEnv:
    def __init__(self, initial_states, horizon_length=1000, num_envs=None, start_index=None, device="cuda:0",
                 clip_rewards=True, min_rew=-10.0, max_rew=1000.0, track_actions=None):
        pool = np.ascontiguousarray(np.asarray(initial_states), dtype=np.int8)
        if pool.ndim != 2 or pool.shape[1] % 2:
            raise ValueError("initial_states must be [P, 2*max_relator_length]")
        self.device = torch.device(device)
        self.pool = torch.from_numpy(pool).to(self.device)
        self.num_envs = int(num_envs if num_envs is not None else pool.shape[0])
        self.max_relator_length = pool.shape[1] // 2
        self.horizon_length = int(horizon_length)
        self.max_reward = self.horizon_length * self.max_relator_length * 2  # ac_env.py:76
        self.clip = (float(min_rew), float(max_rew)) if clip_rewards else (-3.0e38, 3.0e38)
        n = self.num_envs
        # the action ring costs horizon_length bytes per env: on by default up to 2^17 envs (262 MB at horizon 2000)
        self.track_actions = bool(track_actions) if track_actions is not None else n <= (1 << 17)
        if start_index is None:
            start_index = np.arange(n) % pool.shape[0]
        self.start_index = torch.as_tensor(np.asarray(start_index), dtype=torch.int64, device=self.device)
        # where each env restarts when it finishes (device resident; update with set_next_states)
        self.next_state_index = self.start_index.to(torch.int32).contiguous()
        self.buf = None
        self.envs = [_EnvHandle(self, i) for i in range(n)] if n <= 4096 else _LazyHandles(self)
        self.single_observation_space = _Space(shape=(pool.shape[1],), dtype=np.int8)
        self.single_action_space = _Space(shape=(), n=12, dtype=np.int64)
        # pinned host staging for the numpy API
        row = pool.shape[1]
        self._h_act = torch.empty(n, dtype=torch.uint8).pin_memory()
        self._h_obs = torch.empty((n, row), dtype=torch.int8).pin_memory()
        self._h_rew = torch.empty(n, dtype=torch.float32).pin_memory()
        self._h_out = torch.empty((n, 8), dtype=torch.uint8).pin_memory()  # reward | done | truncated | lens
        self._h_cnt = torch.zeros(2, dtype=torch.int32).pin_memory()       # finished envs, rows with a status
        self._d_act = torch.empty(n, dtype=torch.uint8, device=self.device)
        self._d_cnt = torch.zeros(2, dtype=torch.int32, device=self.device)
        self.h2d_bytes_per_step = n
        self.d2h_bytes_per_step = n * row + 4 * n + 8 * n + 8

    # -- gymnasium-vector-like surface --------------------------------------------------------
    def reset(self, seed=None, options=None):
        state = self.pool[self.start_index].contiguous()
        self.buf = batch.EnvBuffers(state, self.horizon_length, pool=self.pool, track_actions=self.track_actions)
        return self.buf.state.cpu().numpy(), {}

    def set_next_states(self, index):
        """Per-env pool index used at the next done/truncated (the curriculum of training.py:319-350)."""
        self.next_state_index.copy_(torch.as_tensor(index, dtype=torch.int32, device=self.device))

    def step_device(self, actions, chained=False):
        """actions: uint8 CUDA tensor [N].  Returns device tensors (views of the env buffers).  Row anomalies
        (the reference raises AssertionError, utils.py:264-266) are counted in `self.buf.err_count` and flagged per
        row in `self.buf.status`; `check_errors()` raises for them (deferred: no host sync on this path)."""
        b = batch.env_step(self.buf, actions, next_idx=self.next_state_index, clip=self.clip, chained=chained)
        return b.state, b.reward_clipped, b.done, b.truncated

    def check_errors(self):
        n = int(self.buf.err_count.item())
        if n:
            self.buf.err_count.zero_()
            raise AssertionError("%d env row(s) hit an invalid presentation / an emptied relator (utils.py:264-266)" % n)

    def _finished(self, n_fin, rec):
        """infos of this step from the finished-episode side buffer (device) and the step records (host)"""
        b, n = self.buf, self.num_envs
        k = min(n_fin, b.fin_cap)
        env = b.fin_env[:k].long()
        length = b.fin_steps[:k].long()
        lmax = int(length.max().item())
        t = b.t - 1  # the step that just ran
        rows = (t - length[:, None] + 1 + torch.arange(lmax, device=self.device)[None, :]) % b.horizon
        acts = b.ring[rows, env[:, None]].cpu().numpy()
        env_h, len_h, obs_h = env.cpu().numpy(), length.cpu().numpy(), b.fin_state[:k].cpu().numpy()
        b.fin_count.zero_()
        final_obs, final_info, actions = {}, {}, {}
        for j in range(k):
            e = int(env_h[j])
            final_obs[e] = obs_h[j]
            if rec[e, 4]:  # done: ACEnv.step returns {"actions": ...} (ac_env.py:110); a truncated env returns {}
                lst = [int(a) for a in acts[j, :len_h[j]]]
                final_info[e] = {"actions": lst}
                actions[e] = lst
            else:
                final_info[e] = {}
        infos = {"final_observation": _Sparse(n, final_obs), "_final_observation": _Sparse(n, {e: True for e in final_obs}, False),
                 "final_info": _Sparse(n, final_info), "_final_info": _Sparse(n, {e: True for e in final_info}, False)}
        if actions:
            infos["actions"] = _Sparse(n, actions)
            infos["_actions"] = _Sparse(n, {e: True for e in actions}, False)
        if n_fin > k:
            infos["fin_overflow"] = n_fin - k
        return infos

    def step(self, actions):
        """numpy actions [N] (any int dtype) -> (obs int8 [N,2m], reward f32, done bool, truncated bool, infos).
        The returned arrays are views of pinned staging buffers that the next step overwrites.  Raises
        AssertionError where the reference's ACMove would (invalid row / emptied relator, utils.py:264-266)."""
        a = np.asarray(actions)
        try:  # dtype conversion into pinned memory: torch's vectorised cast (0.1 ms per 2^20 int64 actions, numpy's takes 0.45)
            self._h_act.copy_(torch.from_numpy(a))
        except (TypeError, ValueError, RuntimeError):  # a dtype or layout torch does not take
            self._h_act.numpy()[:] = a
        self._d_act.copy_(self._h_act, non_blocking=True)
        b = batch.env_step(self.buf, self._d_act, next_idx=self.next_state_index, clip=self.clip)
        self._h_obs.copy_(b.state, non_blocking=True)
        self._h_rew.copy_(b.reward_clipped, non_blocking=True)
        self._h_out.copy_(b.step_out, non_blocking=True)
        if b.fin_count is not None:
            self._d_cnt[0:1].copy_(b.fin_count, non_blocking=True)
        self._d_cnt[1:2].copy_(b.err_count, non_blocking=True)
        self._h_cnt.copy_(self._d_cnt, non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        rec = self._h_out.numpy()
        n_fin, n_err = int(self._h_cnt[0]), int(self._h_cnt[1])
        if n_err:
            b.err_count.zero_()
            bad = np.flatnonzero(b.status.cpu().numpy()) if b.status is not None else []
            raise AssertionError("env rows %s: not a valid presentation / a move emptied a relator (utils.py:264-266)" % list(bad[:8]))
        infos = self._finished(n_fin, rec) if (n_fin and b.fin_count is not None) else {}
        return (self._h_obs.numpy(), self._h_rew.numpy(), rec[:, 4].view(np.bool_), rec[:, 5].view(np.bool_), infos)

    @property
    def lengths(self):
        return self.buf.lens

    @property
    def count_steps(self):
        return self.buf.steps
