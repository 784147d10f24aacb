"""Where an end-to-end step (block upload -> one Lloyd iteration -> block download) spends its time
on every rank.  torchrun ... tools/e2e_breakdown.py [n] [steps]"""
import os, sys, time
import numpy as np
import torch
import torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import fishman_b200 as fb
from fishman_b200 import points

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2621442
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 50
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
m = fb.Mesh.from_xyz(points.fibonacci(n, 0, 0.2), rank=rank, size=world)
if world > 1:
    m.peer_connect(dist)
rho = fb.density_uniform()
m.lloyd(rho, max_iter=20, warm=True)
cnt = m.block_count()
hx = [fb.pinned_empty(max(cnt, 1)) for _ in range(3)]
m.get_generators_block_soa(*hx)
t = np.zeros(4)
if world > 1:
    dist.barrier()
T0 = time.perf_counter()
for _ in range(steps):
    a = time.perf_counter(); m.set_generators_block_soa(*hx)
    b = time.perf_counter(); m.lloyd(rho, max_iter=1, warm=True)
    c = time.perf_counter(); m.get_generators_block_soa(*hx)
    d = time.perf_counter()
    t += (b - a, c - b, d - c, d - a)
tot = (time.perf_counter() - T0) / steps * 1e3
print(f"rank {rank}/{world}: ms per step {tot:.3f}: upload {t[0]/steps*1e3:.3f} lloyd {t[1]/steps*1e3:.3f} download {t[2]/steps*1e3:.3f}  (block {cnt} cells, {24*cnt/1e6:.1f} MB each way)", flush=True)
m.close()
if world > 1:
    dist.destroy_process_group()
