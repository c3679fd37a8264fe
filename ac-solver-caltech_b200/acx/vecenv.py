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
        # pinned host staging for the numpy API
        row = pool.shape[1]
        self._h_act = torch.empty(n, dtype=torch.uint8).pin_memory()
        self._h_obs = torch.empty((n, row), dtype=torch.int8).pin_memory()
        self._h_rew = torch.empty(n, dtype=torch.float32).pin_memory()
        self._h_done = torch.empty(n, dtype=torch.uint8).pin_memory()
        self._h_trunc = torch.empty(n, dtype=torch.uint8).pin_memory()
        self._d_act = torch.empty(n, dtype=torch.uint8, device=self.device)
        self.h2d_bytes_per_step = n
        self.d2h_bytes_per_step = n * row + 4 * n + n + n

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
        """numpy actions [N] (any int dtype) -> (obs int8 [N,2m], reward f32, done bool, truncated bool, {})."""
        a = np.asarray(actions)
        self._h_act.numpy()[:] = a  # dtype conversion into pinned memory
        self._d_act.copy_(self._h_act, non_blocking=True)
        b = batch.env_step(self.buf, self._d_act, next_idx=self.next_state_index, clip=self.clip)
        self._h_obs.copy_(b.state, non_blocking=True)
        self._h_rew.copy_(b.reward_clipped, non_blocking=True)
        self._h_done.copy_(b.done, non_blocking=True)
        self._h_trunc.copy_(b.truncated, non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        return (self._h_obs.numpy(), self._h_rew.numpy(), self._h_done.numpy().view(np.bool_),
                self._h_trunc.numpy().view(np.bool_), {})

    @property
    def lengths(self):
        return self.buf.lens

    @property
    def count_steps(self):
        return self.buf.steps
