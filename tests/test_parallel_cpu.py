"""CPU, world_size 2 over gloo: the host logic of the multi-GPU path (sharding, header broadcast, timing reduce)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from delaunay_linterp_b200 import parallel


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        header = np.arange(16, dtype=np.int64) * 3 + 1 if rank == 0 else None
        gp = np.linspace(0.0, 1.7, 12) if rank == 0 else None
        h, g = parallel.broadcast_header(header, gp, src=0)
        ok = np.array_equal(h, np.arange(16, dtype=np.int64) * 3 + 1) and np.array_equal(g, np.linspace(0.0, 1.7, 12))
        mx = parallel.max_over_ranks(10.0 + rank)
        sm = parallel.sum_over_ranks(1.0 + rank)
        # every rank evaluates its shard; gathering the shards reproduces the full batch order
        n = 1001
        lo, hi = parallel.shard_bounds(n, world, rank)
        xs = np.arange(n, dtype=np.float64)
        part = torch.zeros(n, dtype=torch.float64)
        part[lo:hi] = torch.from_numpy(xs[lo:hi] * 2.0)
        dist.all_reduce(part)
        ok = ok and np.array_equal(part.numpy(), xs * 2.0)
        out.put((rank, bool(ok), mx, sm, lo, hi))
    finally:
        dist.destroy_process_group()


def test_world_size_2_gloo():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(r[1] for r in res)
    assert all(r[2] == 11.0 for r in res) and all(r[3] == 3.0 for r in res)
    assert res[0][4] == 0 and res[0][5] == res[1][4] and res[1][5] == 1001


def test_shard_bounds_partition():
    for n in (0, 1, 7, 100, 1_000_000_007):
        for w in (1, 2, 4, 8):
            b = [parallel.shard_bounds(n, w, r) for r in range(w)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(w - 1))
            sizes = [hi - lo for lo, hi in b]
            assert max(sizes) - min(sizes) <= 1
