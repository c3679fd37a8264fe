"""ACEnvConfig / ACEnv with the reference's surface (ac_solver/envs/ac_env.py:14-132).

A single ACEnv is a VecACEnv-of-one under the hood: its state lives on the GPU and `step` is one
`acx_env_step` launch.  For throughput use acx.vecenv.VecACEnv (ac_solver.agents.environment.get_env)."""
from dataclasses import dataclass, field
from typing import Union

import numpy as np
import torch

from ac_solver import _device
from ac_solver.envs.utils import is_array_valid_presentation
from acx import batch


class _Discrete:
    """stand-in for gymnasium.spaces.Discrete when gymnasium is not installed"""

    def __init__(self, n):
        self.n, self.shape, self.dtype = n, (), np.int64

    def sample(self):
        return int(np.random.randint(self.n))

    def contains(self, x):
        return isinstance(x, (int, np.integer)) and 0 <= int(x) < self.n


class _Box:
    """stand-in for gymnasium.spaces.Box when gymnasium is not installed"""

    def __init__(self, low, high, dtype):
        self.low, self.high, self.dtype, self.shape = low, high, dtype, low.shape

    def sample(self):
        return np.random.randint(self.low, self.high + 1).astype(self.dtype)

    def contains(self, x):
        x = np.asarray(x)
        return x.shape == self.shape and bool(np.all(x >= self.low)) and bool(np.all(x <= self.high))


# The reference's ACEnv is a gymnasium.Env (ac_env.py:8-9,53) with gymnasium spaces, which gymnasium's wrappers
# (NormalizeReward, TransformReward) and SyncVectorEnv check for.  With gymnasium installed the drop-in is one too;
# without it (this image) the stand-ins above provide shape / n / dtype / sample / contains.
try:  # pragma: no cover - depends on the environment
    import gymnasium as _gym
    from gymnasium.spaces import Box as _GymBox, Discrete as _GymDiscrete

    _EnvBase = _gym.Env

    def _make_spaces(m):
        return (_GymBox(np.ones(m * 2, dtype=np.int8) * -2, np.ones(m * 2, dtype=np.int8) * 2, dtype=np.int8), _GymDiscrete(12))
except ImportError:
    _EnvBase = object

    def _make_spaces(m):
        return _Box(np.ones(m * 2, dtype=np.int8) * -2, np.ones(m * 2, dtype=np.int8) * 2, np.int8), _Discrete(12)


@dataclass
class ACEnvConfig:
    initial_state: Union[np.ndarray, list] = field(default_factory=lambda: np.array([1, 0, 2, 0]))
    horizon_length: any = 1000
    use_supermoves: any = False

    def __post_init__(self):  # ac_env.py:22-35
        if isinstance(self.initial_state, list):
            self.initial_state = np.array(self.initial_state)
        if not isinstance(self.initial_state, np.ndarray):
            raise TypeError("initial_state must be a numpy array")
        if self.initial_state.ndim != 1:
            raise ValueError("initial_state must be a 1-dimensional array")
        if len(self.initial_state) % 2 != 0:
            raise ValueError("initial state must have even length")
        if not is_array_valid_presentation(self.initial_state):
            raise ValueError("initial state must be a valid presentation")

    @property
    def max_relator_length(self):
        return len(self.initial_state) // 2

    @classmethod
    def from_dict(cls, config_dict):
        return cls(initial_state=np.array(config_dict.get("initial_state", cls().initial_state)),
                   horizon_length=config_dict.get("horizon_length", cls().horizon_length),
                   use_supermoves=config_dict.get("use_supermoves", cls().use_supermoves))


class ACEnv(_EnvBase):
    def __init__(self, config: ACEnvConfig = None):
        config = config if config is not None else ACEnvConfig()
        self.n_gen = 2
        self.max_relator_length = config.max_relator_length
        self.initial_state = config.initial_state
        self.horizon_length = config.horizon_length
        if config.use_supermoves:
            raise NotImplementedError("ACEnv with supermoves is not yet implemented in this library.")
        m = self.max_relator_length
        self.observation_space, self.action_space = _make_spaces(m)
        self.max_reward = self.horizon_length * m * self.n_gen
        self._dev = _device.device()
        self._load(self.initial_state)

    def _load(self, state):
        s = np.asarray(state)
        self._dtype = s.dtype
        t = torch.from_numpy(np.ascontiguousarray(s, dtype=np.int8)[None, :].copy()).to(self._dev)
        self._buf = batch.EnvBuffers(t, int(self.horizon_length))
        self.state = np.copy(s)
        m = self.max_relator_length
        self.lengths = [int(np.count_nonzero(s[i * m:(i + 1) * m])) for i in range(2)]
        self.count_steps = 0
        self.actions = []

    def step(self, action):  # ac_env.py:91-111
        assert action in range(0, 12), f"Expect n to be in range 0-11 (both inclusive). Got {action}"
        self.actions += [action]
        act = torch.tensor([int(action)], dtype=torch.uint8, device=self._dev)
        b = batch.env_step(self._buf, act)
        rec = torch.cat([b.state.view(torch.uint8).flatten(), b.step_out.flatten(), b.status]).cpu().numpy()
        row = 2 * self.max_relator_length
        assert rec[row + 8] == 0, f"{self.state} is not a valid presentation. Expect all zeros to be padded to the right."
        self.state = rec[:row].view(np.int8).astype(self._dtype)
        reward = int(rec[row:row + 4].view(np.int32)[0])
        done, truncated = bool(rec[row + 4]), bool(rec[row + 5])
        self.lengths = [int(rec[row + 6]), int(rec[row + 7])]
        self.count_steps += 1
        return self.state, reward, done, truncated, ({"actions": self.actions.copy()} if done else {})

    def reset(self, *, seed=None, options=None):  # ac_env.py:113-129
        start = options["starting_state"] if options and "starting_state" in options else self.initial_state
        self._load(np.copy(start))
        return self.state, {}

    def render(self):
        pass
