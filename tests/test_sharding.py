"""N > 1 host logic on CPU: world_size 2 and 3, gloo backend.  Each rank runs the centroid
stage for its own block only (here through the CPU oracle, standing in for the rank's GPU),
the blocks are exchanged with the same all-gather the GPU path uses, and every rank must end
up with exactly the generators of a single-process Lloyd step."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from fishman_b200 import points, sharding
from oracle import oracle as O


def test_block_bounds_cover_exactly():
    for n in (4, 7, 10242, 2621442):
        for size in (1, 2, 3, 8):
            seen = np.zeros(n, dtype=int)
            for r in range(size):
                npad, b, per, valid = sharding.block_bounds(n, r, size)
                assert npad == per * size and npad >= n and npad - n < size
                seen[b:b + valid] += 1
            assert (seen == 1).all()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, size, port, n, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=size)
    p = points.fibonacci(n, 3)
    npad, begin, per, valid = sharding.block_bounds(n, rank, size)
    full = torch.zeros(npad, 4, dtype=torch.float64)
    full[:n, :3] = torch.from_numpy(p)
    full[:n, 3] = torch.arange(n, dtype=torch.float64)
    md_all = []
    for it in range(3):
        cur = full[:n, :3].numpy().copy()
        o = O.Oracle(cur)  # replicated triangulation
        o.delaunay_triangulation()
        o.voronoi_diagram()
        new = o.centroids(O.density_uniform())
        if valid:  # only the rank's block is updated locally
            full[begin:begin + valid, :3] = torch.from_numpy(new[begin:begin + valid])
            md = float(np.linalg.norm(new[begin:begin + valid] - cur[begin:begin + valid], axis=1).max())
        else:
            md = 0.0
        sharding.allgather_blocks(dist, full, begin, per, in_place=False)
        md_all.append(sharding.reduce_max(dist, md, "cpu"))
    np.save(os.path.join(out, f"r{rank}.npy"), full[:n].numpy())
    np.save(os.path.join(out, f"md{rank}.npy"), np.array(md_all))
    dist.destroy_process_group()


@pytest.mark.parametrize("size,n", [(2, 1000), (3, 1001)])
def test_sharded_lloyd_equals_single_process(tmp_path, size, n):
    mp.spawn(_worker, args=(size, _free_port(), n, str(tmp_path)), nprocs=size, join=True)
    o = O.Oracle(points.fibonacci(n, 3))
    md = [o.lloyd_step(O.density_uniform()) for _ in range(3)]
    ref = o.generators()
    for r in range(size):
        got = np.load(tmp_path / f"r{r}.npy")
        assert np.array_equal(got[:, :3], ref), f"rank {r}"
        assert np.array_equal(got[:, 3], np.arange(n))
        assert np.allclose(np.load(tmp_path / f"md{r}.npy"), md, rtol=0, atol=1e-15)
