"""One-shot spherical Delaunay throughput sweep (BASELINE.json configs[4]): N = 1M..10M points,
uniform random (U) and jittered Fibonacci (FIB) sets, points already resident on the device,
fm_triangulate timed with CUDA events on the handle's stream (median of 3).  Prints one JSON
line per case.  Usage: python tools/dt_sweep.py [N ...]
Under torchrun (one process per GPU) the meshes are sharded with attached peers, i.e. the point
location of every insertion round is split over the ranks; times are the max over ranks.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 tools/dt_sweep.py"""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import fishman_b200 as fb  # noqa: E402
from fishman_b200 import points  # noqa: E402

rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
dist = None
if world > 1:
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))

sizes = [int(a) for a in sys.argv[1:]] or [1000000, 2000000, 4000000, 6000000, 8000000, 10000000]
for n in sizes:
    for name, p in (("U", points.uniform(n, 0)), ("FIB", points.fibonacci(n, 0))):
        m = fb.Mesh.from_xyz(p, rank=rank, size=world)
        if dist:
            m.peer_connect(dist)
        m.delaunay_triangulation()  # warm-up (allocations, first-touch)
        ts = []
        for _ in range(3):
            m.stats_reset()
            m.timer_begin()
            m.delaunay_triangulation()
            ts.append(m.timer_end())
        st = m.stats()
        # cheap validity evidence at full size: counts + a warm repair must find nothing to flip
        m.stats_reset()
        m.retriangulate(warm=True)
        idem = m.stats()["flips"]
        ms = float(np.median(ts))
        if dist:
            t = torch.tensor([ms], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        if rank == 0:
            print(json.dumps({"set": name, "n": n, "n_gpus": world, "ms": round(ms, 3), "mpts_per_s": round(n / ms / 1e3, 2),
                              "rounds": st["rounds"], "subrounds": st["subrounds"], "flips": st["flips"],
                              "launches": st["launches"], "flips_on_recheck": idem}), flush=True)
        m.close()
if dist:
    dist.destroy_process_group()
