"""Multi-GPU plumbing: one process per GPU, mesh replicated, query batch sharded, no per-query traffic.

The path has no exchange step: the only collective is the one-time broadcast of the
flattened mesh buffers (NCCL over NVLink through torch.distributed) and the max-over-ranks reduction
of the timing.  The pure host logic here (sharding, header broadcast, timing reduction) also runs on
the gloo backend and is covered by world_size-2 CPU tests.
"""
import numpy as np


def shard_bounds(n, world_size, rank):
    """contiguous 1/world_size slice [lo, hi) of n queries for `rank` (sizes differ by at most one)."""
    lo = n * rank // world_size
    hi = n * (rank + 1) // world_size
    return lo, hi


def broadcast_header(header, grid_params, src=0):
    """broadcast the mesh header (int64[16]) and grid parameters (float64[18]) from `src`."""
    import torch.distributed as dist
    obj = [None]
    if dist.get_rank() == src:
        obj = [(np.asarray(header, dtype=np.int64).tolist(), np.asarray(grid_params, dtype=np.float64).tolist())]
    dist.broadcast_object_list(obj, src=src)
    h, g = obj[0]
    return np.array(h, dtype=np.int64), np.array(g, dtype=np.float64)


def max_over_ranks(value, device=None):
    """max of a python float over all ranks (timing: the slowest rank defines the step time)."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device if device is not None else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(value, device=None):
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device if device is not None else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


def replicate_mesh(mesh, src=0, device=None, **options):
    """Rank `src` passes its Mesh; every other rank passes None and receives a replica whose device buffers
    are filled by an NCCL broadcast.  Returns the local Mesh."""
    import torch
    import torch.distributed as dist
    from .mesh import Mesh
    rank = dist.get_rank()
    if rank == src:
        header, gp = mesh.header()
    else:
        header, gp = None, None
    header, gp = broadcast_header(header, gp, src=src)
    if rank != src:
        mesh = Mesh.from_header(header, gp, **options)
    dev = device if device is not None else torch.device("cuda", torch.cuda.current_device())
    for i, (_, nbytes) in enumerate(mesh.buffers()):
        if nbytes == 0:
            continue
        buf = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        if rank == src:
            mesh.buffer_read(i, buf.data_ptr())
        dist.broadcast(buf, src=src)
        if rank != src:
            mesh.buffer_write(i, buf.data_ptr())
        del buf
    torch.cuda.synchronize(dev)
    return mesh
