"""one single-GPU PlaneBFS run from AK(3) (m = 36) at a small budget: the command the bfs kernels are profiled on"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "ac-solver-caltech_b200"))
import bench_bfs  # noqa: E402

budget = int(float(sys.argv[1])) if len(sys.argv) > 1 else 6 * 10 ** 7
torch.cuda.set_device(0)
dev = torch.device("cuda", 0)
bench_bfs.init_single_process_group(dev)
os.environ["ACX_BFS_OVERLAP"] = "0"
r = bench_bfs._timed_run(bench_bfs.ak3(), budget, False, 1, dev, 1 << 20)
print({k: r[k] for k in ("n_seen", "nodes_expanded", "wall_s", "rounds", "levels")})
