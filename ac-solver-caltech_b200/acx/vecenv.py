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
"""
import numpy as np
import torch

from . import batch


class _EnvHandle:
    """`envs.envs[i]`: what the PPO loop pokes at for a single env (agents/training.py:352)."""

    def __init__(self, vec, index):
        self._vec, self._i = vec, index

    def reset(self, *, seed=None, options=None):
        v, i = self._vec, self._i
        if options and "starting_state" in options:
            row = torch.from_numpy(np.ascontiguousarray(np.asarray(options["starting_state"]), dtype=np.int8)).to(v.device)
        else:
            row = v.pool[int(v.start_index[i])]
        v.buf.state[i].copy_(row)
        v.buf.steps[i] = 0
        m = v.max_relator_length
        v.buf.lens[i, 0] = int((row[:m] != 0).sum())
        v.buf.lens[i, 1] = int((row[m:] != 0).sum())
        if v.buf.certified and not batch.certify(row[None, :].contiguous()):
            v.buf.certified = False  # an arbitrary starting state: back to the checked kernel
        return v.buf.state[i].cpu().numpy(), {}

    @property
    def state(self):
        return self._vec.buf.state[self._i].cpu().numpy()

    @property
    def lengths(self):
        return [int(x) for x in self._vec.buf.lens[self._i].cpu().numpy()]

    @property
    def count_steps(self):
        return int(self._vec.buf.steps[self._i])


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


class VecACEnv:
    def __init__(self, initial_states, horizon_length=1000, num_envs=None, start_index=None, device="cuda:0",
                 clip_rewards=True, min_rew=-10.0, max_rew=1000.0):
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
        if start_index is None:
            start_index = np.arange(n) % pool.shape[0]
        self.start_index = torch.as_tensor(np.asarray(start_index), dtype=torch.int64, device=self.device)
        # where each env restarts when it finishes (device resident; update with set_next_states)
        self.next_state_index = self.start_index.to(torch.int32).contiguous()
        self.buf = None
        self.envs = [_EnvHandle(self, i) for i in range(n)] if n <= 4096 else _LazyHandles(self)
        self.single_observation_space = _Space(shape=(pool.shape[1],), dtype=np.int8)
        self.single_action_space = _Space(n=12)
        # pinned host staging for the numpy API
        row = pool.shape[1]
        self._h_act = torch.empty(n, dtype=torch.uint8).pin_memory()
        self._h_obs = torch.empty((n, row), dtype=torch.int8).pin_memory()
        self._h_rew = torch.empty(n, dtype=torch.float32).pin_memory()
        self._h_out = torch.empty((n, 8), dtype=torch.uint8).pin_memory()  # reward | done | truncated | lens
        self._d_act = torch.empty(n, dtype=torch.uint8, device=self.device)
        self.h2d_bytes_per_step = n
        self.d2h_bytes_per_step = n * row + 4 * n + 8 * n

    # -- gymnasium-vector-like surface --------------------------------------------------------
    def reset(self, seed=None, options=None):
        state = self.pool[self.start_index].contiguous()
        self.buf = batch.EnvBuffers(state, self.horizon_length, pool=self.pool)
        return self.buf.state.cpu().numpy(), {}

    def set_next_states(self, index):
        """Per-env pool index used at the next done/truncated (the curriculum of training.py:319-350)."""
        self.next_state_index.copy_(torch.as_tensor(index, dtype=torch.int32, device=self.device))

    def step_device(self, actions):
        """actions: uint8 CUDA tensor [N].  Returns device tensors (views of the env buffers)."""
        b = batch.env_step(self.buf, actions, next_idx=self.next_state_index, clip=self.clip)
        return b.state, b.reward_clipped, b.done, b.truncated

    def step(self, actions):
        """numpy actions [N] (any int dtype) -> (obs int8 [N,2m], reward f32, done bool, truncated bool, {}).
        The returned arrays are views of pinned staging buffers that the next step overwrites."""
        a = np.asarray(actions)
        self._h_act.numpy()[:] = a  # dtype conversion into pinned memory
        self._d_act.copy_(self._h_act, non_blocking=True)
        b = batch.env_step(self.buf, self._d_act, next_idx=self.next_state_index, clip=self.clip)
        self._h_obs.copy_(b.state, non_blocking=True)
        self._h_rew.copy_(b.reward_clipped, non_blocking=True)
        self._h_out.copy_(b.step_out, non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        rec = self._h_out.numpy()
        return (self._h_obs.numpy(), self._h_rew.numpy(), rec[:, 4].view(np.bool_), rec[:, 5].view(np.bool_), {})

    @property
    def lengths(self):
        return self.buf.lens

    @property
    def count_steps(self):
        return self.buf.steps
