"""Two ranks on ONE GPU must be refused by fm_peer_attach (their barrier kernels would wait on each
other on one device).  Run under torchrun with 2 processes; exit code 0 = both ranks were refused
with the expected message."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import fishman_b200 as fb  # noqa: E402
from fishman_b200 import points  # noqa: E402

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(0)
dist.init_process_group("gloo")
m = fb.Mesh.from_xyz(points.ico(3), rank=rank, size=world)
try:
    m.peer_connect(dist)
    ok = False
except fb.FishmanError as e:
    ok = "same GPU" in str(e)
    print(f"rank {rank}: refused: {e}", flush=True)
m.close()
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
