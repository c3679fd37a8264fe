"""Import the UNMODIFIED reference from /root/reference (build container only).

TEST INFRASTRUCTURE ONLY.  /root/reference does not exist on the GPU box, so
nothing that runs there (pytest -m gpu, smoke(), bench.py) may import this
module; it is used by oracle/validate_vs_reference.py and
tests/golden/make_golden.py to pin the oracle and to generate fixtures.

`gymnasium` (pinned 0.28.1 by the reference, pyproject.toml:40) is not installed
here.  ac_solver/__init__.py imports ac_env eagerly, which needs only
`gymnasium.Env` as a base class and the `Discrete`/`Box` constructors
(ac_env.py:8-9, :66-73); a stand-in providing exactly those names and no
arithmetic is registered before the import.
"""
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("ACX_REFERENCE_ROOT", "/root/reference")


def available():
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "ac_solver"))


def _install_gymnasium_stand_in():
    if "gymnasium" in sys.modules:
        return
    gym = types.ModuleType("gymnasium")
    spaces = types.ModuleType("gymnasium.spaces")

    class Env:  # base class only
        pass

    class Discrete:
        def __init__(self, n):
            self.n = n

    class Box:
        def __init__(self, low, high, dtype=None):
            self.low, self.high, self.dtype = low, high, dtype
            self.shape = getattr(low, "shape", None)

    gym.Env = Env
    spaces.Discrete, spaces.Box = Discrete, Box
    gym.spaces = spaces
    sys.modules["gymnasium"] = gym
    sys.modules["gymnasium.spaces"] = spaces


def load():
    """Returns a namespace with the reference's hot-path callables."""
    if not available():
        raise RuntimeError(f"reference not found at {REFERENCE_ROOT}")
    _install_gymnasium_stand_in()
    # a product package of the same name may already be imported: evict it
    saved = {k: sys.modules.pop(k) for k in list(sys.modules) if k == "ac_solver" or k.startswith("ac_solver.")}
    sys.path.insert(0, REFERENCE_ROOT)
    try:
        import importlib

        ns = types.SimpleNamespace()
        moves = importlib.import_module("ac_solver.envs.ac_moves")
        utils = importlib.import_module("ac_solver.envs.utils")
        env = importlib.import_module("ac_solver.envs.ac_env")
        greedy = importlib.import_module("ac_solver.search.greedy")
        bfs = importlib.import_module("ac_solver.search.breadth_first")
        ns.ACMove, ns.concatenate_relators, ns.conjugate = moves.ACMove, moves.concatenate_relators, moves.conjugate
        ns.simplify_relator, ns.simplify_presentation = utils.simplify_relator, utils.simplify_presentation
        ns.is_array_valid_presentation = utils.is_array_valid_presentation
        ns.is_presentation_trivial = utils.is_presentation_trivial
        ns.generate_trivial_states = utils.generate_trivial_states
        ns.convert_relators_to_presentation = utils.convert_relators_to_presentation
        ns.change_max_relator_length_of_presentation = utils.change_max_relator_length_of_presentation
        ns.ACEnv, ns.ACEnvConfig = env.ACEnv, env.ACEnvConfig
        ns.greedy_search, ns.bfs = greedy.greedy_search, bfs.bfs
        ns.root = REFERENCE_ROOT
        return ns
    finally:
        sys.path.remove(REFERENCE_ROOT)
        for k in [k for k in sys.modules if k == "ac_solver" or k.startswith("ac_solver.")]:
            sys.modules.pop(k)
        sys.modules.update(saved)
