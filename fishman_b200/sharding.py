"""Multi-GPU plumbing of the Lloyd loop: one process per GPU, the
triangulation replicated, the centroid / quadrature / update stage sharded over contiguous
blocks of the internal (space-filling-curve) order, one all-gather of generator records per
iteration.  The block arithmetic here is the same as fm_create_dist's (csrc/fishman_b200.cu,
create_impl) and is what the world_size-2 gloo tests exercise on CPU."""
import torch


def block_bounds(n, rank, size):
    """(n_padded, begin, count, valid): records per rank = ceil(n / size); rank r owns
    [r * per, (r + 1) * per) of the padded array, of which `valid` are real generators."""
    per = (n + size - 1) // size
    begin = rank * per
    valid = 0 if begin >= n else min(per, n - begin)
    return per * size, begin, per, valid


class DeviceRecords:
    """torch view of the handle's generator records: float64 [n_padded, 4] = (x, y, z, id bits)."""

    def __init__(self, ptr, n_padded, device):
        self.__cuda_array_interface__ = {"shape": (int(n_padded), 4), "typestr": "<f8",
                                         "data": (int(ptr), False), "version": 2}
        self.tensor = torch.as_tensor(self, device=device)


def allgather_blocks(dist, full, begin, count, in_place=True):
    """Every rank contributes full[begin:begin+count]; afterwards full is identical on all
    ranks.  in_place: NCCL accepts the send buffer inside the receive buffer (no staging
    copy); gloo does not, so the CPU tests pass in_place=False."""
    src = full[begin:begin + count]
    if not in_place:
        src = src.clone()
    dist.all_gather_into_tensor(full.view(-1), src.reshape(-1))
    return full


def reduce_max(dist, value, device):
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
