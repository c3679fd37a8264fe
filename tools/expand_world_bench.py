"""acx_pbfs_expand2 on ONE GPU with the records routed to `world` local regions: what sorting the records into `world`
runs costs without any NVLink in the way (the 8-GPU search spends 2.6 x the one-GPU time per parent in this kernel).

    python tools/expand_world_bench.py [budget of the search that provides the parents]"""
import ctypes
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "ac-solver-caltech_b200"))
import bench_bfs  # noqa: E402
from acx import _lib  # noqa: E402
from acx.batch import _ptr  # noqa: E402
from acx.pbfs import PlaneBFS, make_plane_shard  # noqa: E402

budget = int(float(sys.argv[1])) if len(sys.argv) > 1 else 3 * 10 ** 7
torch.cuda.set_device(0)
dev = torch.device("cuda", 0)
bench_bfs.init_single_process_group(dev)
shard = make_plane_shard(bench_bfs.ak3(), False, budget + 4096, 1 << 20, str(dev))
eng = PlaneBFS(bench_bfs.ak3(), budget, shard, chunk=1 << 20)
eng.run()
n = int(shard.n_local_dev.item())
gid = shard.gid[:n]
ok = torch.nonzero((gid >= 0) & (gid < (1 << 62))).flatten()
f = 1 << 20
loc = ok[-f:].contiguous()  # the last 2^20 committed nodes: deep parents, as in the big rounds
gids = gid[loc].contiguous()
L = _lib.lib()
ctl = torch.zeros(16, dtype=torch.int64, device=dev)
stream = torch.cuda.current_stream(dev)
for world in (1, 2, 4, 8):
    rcap = (12 * f // world * 2 + 8192) // 4 * 4
    bufs = torch.empty((world, rcap, 32), dtype=torch.uint8, device=dev)
    dst = (ctypes.c_void_p * world)(*[bufs[o].data_ptr() for o in range(world)])
    cur = torch.zeros(world + 1, dtype=torch.int64, device=dev)
    ts = []
    for rep in range(5):
        cur.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        _lib.check(L.acx_pbfs_expand2(_ptr(shard.nodes), _ptr(shard.action), _ptr(loc), _ptr(gids), int(gids.min()), f, 36, 0, world, rcap,
                                      _ptr(cur), dst, _ptr(ctl), stream.cuda_stream), "acx_pbfs_expand2")
        b.record(stream)
        b.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    print({"world": world, "us_per_2^20_parents": round(min(ts), 1), "records": int(cur[:world].sum()), "overflow": int(cur[world])})
