"""GPU, >= 2 devices: replicated mesh + sharded queries give bit-identical results to the 1-GPU run."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SCRIPT = r'''
import os, sys, numpy as np, torch, torch.distributed as dist
sys.path.insert(0, %(root)r); sys.path.insert(0, os.path.join(%(root)r, "tests"))
import helpers, delaunay_linterp_b200 as dl
from delaunay_linterp_b200 import parallel
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
mesh = None
if rank == 0:
    pts, f, sv, sn = helpers.random_mesh(2, 50000, 5)
    mesh = dl.Mesh(pts, f, sv, sn, device=lr, **%(opts)r)
mesh = parallel.replicate_mesh(mesh, src=0, device=dev, **%(opts)r)
q = np.random.default_rng(6).random((400000, 2)) * 1.04 - 0.02
lo, hi = parallel.shard_bounds(len(q), world, rank)
ys = mesh.interp(torch.from_numpy(q[lo:hi]).to(dev))
full = torch.zeros(len(q), dtype=torch.float64, device=dev)
full[lo:hi] = ys
dist.all_reduce(full)
if rank == 0:
    ref = mesh.interp(torch.from_numpy(q).to(dev))
    assert torch.equal(full, ref), "sharded result differs from the 1-GPU result"
    print("MULTI_OK", world)
dist.barrier(); dist.destroy_process_group()
'''


@pytest.mark.parametrize("opts", [{}, {"packed_faces": 3}])
def test_sharded_equals_single(gpu, tmp_path, opts):
    if gpu < 2:
        pytest.skip("needs >= 2 GPUs (run with gpurun --gpus 2)")
    script = tmp_path / "multi.py"
    script.write_text(SCRIPT % {"root": ROOT, "opts": opts})
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29533", str(script)],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "MULTI_OK 2" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]
