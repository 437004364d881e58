"""BASELINE config 5 under torchrun: 16M-vertex 2D mesh built on rank 0, replicated by one NCCL broadcast, 1B queries
split across the ranks (125M per GPU at 8), Philox seed 2006 + rank, generated on the device.  Prints one JSON line.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 scripts/cfg5_multi.py"""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch, torch.distributed as dist
import helpers
import delaunay_linterp_b200 as dl
from delaunay_linterp_b200 import parallel

rank, world, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
NV = int(os.environ.get("CFG5_NV", 16000000)); NQ_TOTAL = int(os.environ.get("CFG5_NQ", 1000000000)); STEPS = 5
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
if world > 1: dist.init_process_group("nccl", device_id=dev)
t0 = time.perf_counter(); mesh = None
if rank == 0:
    pts = np.random.default_rng(1005).random((NV, 2))
    t = dl.DelaunayInterp2(); t.insert_many(pts, helpers.cobb_douglas(pts), use_gpu=False)
    c, f, tri, nb = t.export(); del t
    t_tri = time.perf_counter() - t0
    mesh = dl.Mesh(c, f, tri, nb, device=lr)
t_build = time.perf_counter() - t0
t0 = time.perf_counter()
if world > 1: mesh = parallel.replicate_mesh(mesh, src=0, device=dev)
t_repl = time.perf_counter() - t0
lo, hi = parallel.shard_bounds(NQ_TOTAL, world, rank); nq = hi - lo
g = torch.Generator(device=dev); g.manual_seed(2006 + rank)
xq = torch.rand(nq, 2, dtype=torch.float64, device=dev, generator=g) * 0.96 + 0.02
ys = torch.empty(nq, dtype=torch.float64, device=dev)
for _ in range(3): mesh.interp(xq, out=ys)
torch.cuda.synchronize(dev)
if world > 1: dist.barrier()
torch.cuda.synchronize(dev)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(STEPS): mesh.interp(xq, out=ys)
e1.record(); torch.cuda.synchronize(dev)
if world > 1: dist.barrier()
ms = parallel.max_over_ranks(e0.elapsed_time(e1), device=dev) / STEPS
nan = parallel.sum_over_ranks(float(torch.isnan(ys).sum().item()), device=dev)
if rank == 0:
    print(json.dumps({"config": "cfg5: 16M-vertex 2D mesh replicated, 1B queries sharded", "n_gpus": world, "mesh_vertices": NV,
                      "mesh_faces": int(mesh.n_simplices), "mesh_device_MB": round(mesh.device_bytes / 1e6), "queries_total": NQ_TOTAL,
                      "queries_per_gpu": nq, "ms_per_step": round(ms, 3), "value_gqps": round(NQ_TOTAL / ms / 1e6, 2),
                      "per_gpu_gqps": round(nq / ms / 1e6, 2), "roofline_frac_per_gpu(108B)": round(nq / ms * 1e3 * 108 / 6462.4e9, 3),
                      "triangulate_s": round(t_tri, 1), "flatten_s": round(t_build - t_tri, 1), "replicate_s": round(t_repl, 2), "nan": nan}), flush=True)
if world > 1: dist.barrier(); dist.destroy_process_group()
