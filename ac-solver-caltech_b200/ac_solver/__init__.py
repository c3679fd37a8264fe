"""Drop-in for the hot path of AC-Solver (Avi161/AC-Solver-Caltech) on top of the acx CUDA engine.

Same import paths and call signatures as the reference for: ACMove, concatenate_relators, conjugate,
simplify_relator, simplify_presentation, the predicates and padding helpers, ACEnvConfig/ACEnv,
greedy_search and bfs; plus the GPU-resident env vector (`ac_solver.agents.environment.get_env`).
All word arithmetic runs in libacx.so on the GPU; there is no CPU implementation behind it.
"""
from ac_solver.envs.ac_env import ACEnvConfig, ACEnv  # noqa: F401
from ac_solver.envs.ac_moves import ACMove  # noqa: F401
from ac_solver.search.breadth_first import bfs  # noqa: F401
from ac_solver.search.greedy import greedy_search  # noqa: F401
