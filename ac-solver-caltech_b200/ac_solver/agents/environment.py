"""Env-vector layer (reference: ac_solver/agents/environment.py:19-127), GPU resident.

`get_env(args)` keeps the reference's return tuple but `envs` is an acx.vecenv.VecACEnv: all
`num_envs` environments advance in ONE kernel launch per step instead of gymnasium's sequential
SyncVectorEnv loop, rewards are clipped in the same launch (TransformReward(np.clip), :48-52), and
finished envs are re-seated on the device (agents/training.py:351-352).  NormalizeReward
(:45-46, off by default) is a learner-side statistic and is not part of this layer."""
import numpy as np

from acx.vecenv import VecACEnv
from ac_solver.agents.utils import load_initial_states_from_text_file
from ac_solver.envs.utils import change_max_relator_length_of_presentation, convert_relators_to_presentation


class _ClippedACEnv:
    """TransformReward(env, np.clip) of the reference's make_env (environment.py:48-52) around the scalar ACEnv"""

    def __init__(self, env, lo, hi):
        self.env, self._lo, self._hi = env, lo, hi

    def step(self, action):
        obs, reward, done, truncated, info = self.env.step(action)
        return obs, np.clip(reward, self._lo, self._hi), done, truncated, info

    def reset(self, **kw):
        return self.env.reset(**kw)

    @property
    def unwrapped(self):
        return self.env

    def __getattr__(self, name):
        return getattr(self.env, name)


def make_env(presentation, args):
    """A thunk building the scalar ACEnv with the reference's per-env configuration (environment.py:19-57)."""

    def thunk():
        from ac_solver.envs.ac_env import ACEnv, ACEnvConfig

        if getattr(args, "norm_rewards", False):
            raise NotImplementedError("NormalizeReward is a learner-side running statistic; apply it to the reward tensor")
        env = ACEnv(ACEnvConfig.from_dict({"initial_state": presentation, "horizon_length": args.horizon_length,
                                           "use_supermoves": getattr(args, "use_supermoves", False)}))
        if getattr(args, "clip_rewards", False):
            assert args.min_rew < args.max_rew, "min_rew must be less than max_rew"
            env = _ClippedACEnv(env, args.min_rew, args.max_rew)
        return env

    return thunk


def get_env(args):
    if getattr(args, "use_supermoves", False):
        raise NotImplementedError("ACEnv with supermoves is not yet implemented in this library.")
    if getattr(args, "norm_rewards", False):
        raise NotImplementedError("NormalizeReward is a learner-side running statistic; apply it to the reward tensor")
    if args.fixed_init_state:
        presentation = convert_relators_to_presentation(args.relator1, args.relator2, args.max_relator_length)
        initial_states = [presentation]
        start_index = np.zeros(args.num_envs, dtype=np.int64)
    else:
        initial_states = load_initial_states_from_text_file(states_type=args.states_type)
        # the reference asserts this (environment.py:80-83); `allow_env_wraparound` lifts it for batches larger than
        # the dataset (BASELINE configs[2]: 65536 envs over the 1190 presentations, env e starts on state e % P)
        assert args.num_envs <= len(initial_states) or getattr(args, "allow_env_wraparound", False), \
            "Expect number of environments to be less than number of distinct initial states for now"
        args.max_relator_length = 36  # environment.py:87
        initial_states = [change_max_relator_length_of_presentation(np.array(s), args.max_relator_length)
                          for s in initial_states]
        start_index = np.arange(args.num_envs, dtype=np.int64) % len(initial_states)
    if getattr(args, "clip_rewards", False):
        assert args.min_rew < args.max_rew, "min_rew must be less than max_rew"
    envs = VecACEnv(np.stack(initial_states), horizon_length=args.horizon_length, num_envs=args.num_envs,
                    start_index=start_index, clip_rewards=bool(getattr(args, "clip_rewards", False)),
                    min_rew=getattr(args, "min_rew", -10.0), max_rew=getattr(args, "max_rew", 1000.0),
                    device=getattr(args, "device", "cuda:0"))
    curr_states = [int(k) for k in start_index]
    states_processed = set(curr_states)
    success_record = {"solved": set(), "unsolved": set(range(len(initial_states)))}
    ACMoves_hist = {}
    return envs, initial_states, curr_states, success_record, ACMoves_hist, states_processed
