import sys, os, time, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "ac-solver-caltech_b200"))
import numpy as np
import bench_bfs
from acx.pbfs import make_plane_shard
torch.cuda.set_device(0); dev = torch.device("cuda", 0)
bench_bfs.init_single_process_group(dev)
n = 600_000_000
sh = make_plane_shard(bench_bfs.ak3(), False, n + 1000, 1 << 20, str(dev))
sh.connect()
g = torch.randperm(n, device=dev)  # worst case: random windows
g2 = torch.arange(n, device=dev) + torch.randint(0, 5_000_000, (n,), device=dev) - 2_500_000  # realistic: near-sorted
g2.clamp_(0, n - 1)
def t(f, reps=3):
    torch.cuda.synchronize(); best = 1e9
    for _ in range(reps):
        a = time.perf_counter(); f(); torch.cuda.synchronize(); best = min(best, time.perf_counter() - a)
    return best * 1e3
for name, gg in (("random", g), ("near-sorted", g2)):
    sh.gid[:n].copy_(gg)
    gl = sh.gid[:n]
    print(name, "all 573 windows: %.2f ms" % t(lambda: sh.window_partition(gl, 0, 0, 1 << 20, (n + (1 << 20) - 1) >> 20)))
    print(name, "first 64 windows: %.2f ms" % t(lambda: sh.window_partition(gl, 0, 0, 1 << 20, 64)))
print("reference: gid.sum() %.2f ms" % t(lambda: sh.gid[:n].sum()))
