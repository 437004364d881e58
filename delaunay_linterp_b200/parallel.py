"""Multi-GPU plumbing: one process per GPU, mesh replicated, query batch sharded, no per-query traffic.

The path has no exchange step: the only collective is the one-time broadcast of the
flattened mesh buffers (NCCL over NVLink through torch.distributed) and the max-over-ranks reduction
of the timing.  The pure host logic here (sharding, header broadcast, timing reduction) also runs on
the gloo backend and is covered by world_size-2 CPU tests.
"""
import ctypes as C
import os

import numpy as np


def shard_bounds(n, world_size, rank):
    """contiguous 1/world_size slice [lo, hi) of n queries for `rank` (sizes differ by at most one)."""
    lo = n * rank // world_size
    hi = n * (rank + 1) // world_size
    return lo, hi


def broadcast_header(header, grid_params, src=0):
    """broadcast the mesh header (int64[16]) and grid parameters (float64[3 * DLI_MAX_DIM]) from `src`."""
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


_NCCL = None


class _NcclUniqueId(C.Structure):
    _fields_ = [("internal", C.c_char * 128)]


def _nccl_lib():
    """the NCCL this process uses (torch's bundled copy first), loaded into the global symbol scope so that libdli.so's
    dli_mesh_replicate resolves the same library (it looks NCCL up at run time)."""
    global _NCCL
    if _NCCL is not None:
        return _NCCL
    cands = []
    try:
        import nvidia.nccl
        cands.append(os.path.join(list(nvidia.nccl.__path__)[0], "lib", "libnccl.so.2"))
    except Exception:
        pass
    cands += ["libnccl.so.2", "libnccl.so"]
    for c in cands:
        try:
            L = C.CDLL(c, mode=C.RTLD_GLOBAL)
        except OSError:
            continue
        L.ncclGetUniqueId.argtypes = [C.POINTER(_NcclUniqueId)]
        L.ncclCommInitRank.argtypes = [C.POINTER(C.c_void_p), C.c_int, _NcclUniqueId, C.c_int]
        L.ncclCommDestroy.argtypes = [C.c_void_p]
        L.ncclBroadcast.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        _NCCL = L
        return L
    raise OSError("libnccl.so.2 not found")


def nccl_communicator(src=0):
    """a raw ncclComm_t over all ranks of the torch.distributed job (the unique id travels through torch.distributed).
    The current CUDA device must be this rank's device.  Returns (comm, destroy)."""
    import torch.distributed as dist
    L = _nccl_lib()
    uid = _NcclUniqueId()
    obj = [None]
    if dist.get_rank() == src:
        rc = L.ncclGetUniqueId(C.byref(uid))
        if rc != 0:
            raise RuntimeError("ncclGetUniqueId failed: %d" % rc)
        obj = [C.string_at(C.byref(uid), 128)]
    dist.broadcast_object_list(obj, src=src)
    C.memmove(C.byref(uid), obj[0], 128)
    comm = C.c_void_p()
    rc = L.ncclCommInitRank(C.byref(comm), dist.get_world_size(), uid, dist.get_rank())
    if rc != 0:
        raise RuntimeError("ncclCommInitRank failed: %d" % rc)
    return comm, (lambda: L.ncclCommDestroy(comm))


_COMM = None        # (ncclComm_t, world_size): created once per process, kept until shutdown_nccl()


def cached_nccl_communicator():
    """the raw communicator of this process, created on first use and kept: ncclCommInitRank costs 0.2 - 0.6 s per call and
    grows with the number of ranks (it was paid on every replicate_mesh call).  Collective on first use."""
    global _COMM
    import torch.distributed as dist
    if _COMM is None or _COMM[1] != dist.get_world_size():
        comm, _destroy = nccl_communicator(src=0)
        # NCCL connects its channels lazily, inside the first collective (0.3 s at 2 GPUs): pay that here, with an
        # 8-byte broadcast, so that the first dli_mesh_replicate costs what the mesh bytes cost
        import torch
        warm = torch.zeros(8, dtype=torch.uint8, device="cuda")
        rc = _nccl_lib().ncclBroadcast(warm.data_ptr(), warm.data_ptr(), 8, 0, 0, comm, torch.cuda.current_stream().cuda_stream)
        torch.cuda.current_stream().synchronize()
        if rc != 0:
            raise RuntimeError("ncclBroadcast (warm-up) failed: %d" % rc)
        _COMM = (comm, dist.get_world_size())
    return _COMM[0]


def shutdown_nccl():
    """destroy the cached communicator (optional; call on every rank before destroy_process_group)."""
    global _COMM
    if _COMM is not None:
        _nccl_lib().ncclCommDestroy(_COMM[0])
        _COMM = None


def replicate_mesh_nccl(mesh, src=0):
    """dli_mesh_replicate (include/dli.h) with a raw NCCL communicator: what a C / C++ caller does.  Rank `src` passes
    its Mesh, the others None; returns the local Mesh."""
    import torch.distributed as dist
    from ._capi import check, lib
    from .mesh import Mesh
    comm = cached_nccl_communicator()
    h = C.c_void_p(mesh._h.value if (mesh is not None and dist.get_rank() == src) else None)
    check(lib().dli_mesh_replicate(C.byref(h), comm, int(src)))
    if dist.get_rank() == src:
        return mesh
    return Mesh._from_handle(h, lib().dli_mesh_dim(h))


def replicate_mesh(mesh, src=0, device=None, use_c_abi=True, **options):
    """Rank `src` passes its Mesh; every other rank passes None and receives a replica whose device buffers
    are filled by an NCCL broadcast.  Returns the local Mesh.  On the NCCL backend the work is done by
    dli_mesh_replicate (one ncclBroadcast per buffer, straight from / into the mesh buffers); otherwise (gloo, or no
    NCCL library on ANY rank -- the ranks agree on the path with an all-reduce, a mixed choice would deadlock) by
    torch.distributed broadcasts of the raw buffers.  Either way the replica answers like the source: the options
    that change results travel in the header; `options` only chooses the launch shape of the replica."""
    import torch
    import torch.distributed as dist
    from .mesh import Mesh
    if use_c_abi and dist.get_backend() == "nccl":
        try:
            _nccl_lib()
            have = 1
        except OSError:
            have = 0
        flag = torch.tensor([have], dtype=torch.int32, device=device if device is not None else torch.device("cuda", torch.cuda.current_device()))
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if int(flag.item()) == 1:
            return replicate_mesh_nccl(mesh, src=src)
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
