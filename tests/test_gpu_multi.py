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
    pts, f, sv, sn = helpers.random_mesh(%(dim)d, %(nv)d, 5)
    mesh = dl.Mesh(pts, f, sv, sn, device=lr, **%(opts)r)
# the receiving ranks pass no options: layout and result-affecting options travel in the header
mesh = parallel.replicate_mesh(mesh, src=0, device=dev, use_c_abi=%(c_abi)r)
mesh = parallel.replicate_mesh(mesh if rank == 0 else None, src=0, device=dev, use_c_abi=%(c_abi)r)   # again: cached communicator
q = np.random.default_rng(6).random((400000, %(dim)d)) * 1.04 - 0.02
lo, hi = parallel.shard_bounds(len(q), world, rank)
ys = mesh.interp(torch.from_numpy(q[lo:hi]).to(dev))
full = torch.zeros(len(q), dtype=torch.float64, device=dev)
full[lo:hi] = ys
dist.all_reduce(full)
if rank == 0:
    ref = mesh.interp(torch.from_numpy(q).to(dev))
    same = (full == ref) | (torch.isnan(full) & torch.isnan(ref))      # (dD: NaN outside the hull; all_reduce keeps NaN)
    assert bool(same.all()), "sharded result differs from the 1-GPU result"
    print("MULTI_OK", world)
dist.barrier(); parallel.shutdown_nccl(); dist.destroy_process_group()
'''


@pytest.mark.parametrize("c_abi", [True, False])   # dli_mesh_replicate (raw NCCL communicator) / torch.distributed broadcasts
@pytest.mark.parametrize("opts", [{}, {"packed_faces": 3}, {"packed_faces": 2}, {"outside_nan": 1}, {"epilogue": "closed_form"}])
def test_sharded_equals_single(gpu, tmp_path, opts, c_abi):
    if gpu < 2:
        pytest.skip("needs >= 2 GPUs (run with gpurun --gpus 2)")
    script = tmp_path / "multi.py"
    script.write_text(SCRIPT % {"root": ROOT, "opts": opts, "c_abi": c_abi, "dim": 2, "nv": 50000})
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29533", str(script)],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "MULTI_OK 2" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


@pytest.mark.parametrize("dim,nv,opts", [(3, 6000, {}), (4, 2000, {}), (4, 2000, {"start_records": -1}), (4, 2000, {"packed_faces": 5})])
def test_sharded_equals_single_dd(gpu, tmp_path, dim, nv, opts):
    """dD replicas: 3D records, 4D packed records with and without start copies, 4D split records"""
    if gpu < 2:
        pytest.skip("needs >= 2 GPUs (run with gpurun --gpus 2)")
    script = tmp_path / "multi.py"
    script.write_text(SCRIPT % {"root": ROOT, "opts": opts, "c_abi": True, "dim": dim, "nv": nv})
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29534", str(script)],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "MULTI_OK 2" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


def test_replicate_from_cpp_with_nccl(gpu, tmp_path):
    """tests/cpp/test_replicate.cpp: dli_mesh_replicate called from C++ with ncclCommInitAll communicators, one thread
    per GPU; replicas (2D and 3D) answer bit-identically."""
    if gpu < 2:
        pytest.skip("needs >= 2 GPUs (run with gpurun --gpus 2)")
    import delaunay_linterp_b200 as dl
    libdir = os.path.dirname(dl.LIB_PATH)
    exe = str(tmp_path / "test_replicate")
    env = dict(os.environ)
    env.pop("CXX", None)
    subprocess.check_call(["g++", "-std=c++11", "-O1", os.path.join(ROOT, "tests", "cpp", "test_replicate.cpp"),
                           "-I", os.path.join(ROOT, "include"), "-I", "/usr/local/cuda/include", "-L", libdir,
                           "-L", "/usr/local/cuda/lib64", "-ldli", "-lnccl", "-lcudart", "-lpthread",
                           "-Wl,-rpath," + libdir, "-o", exe], env=env)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "ok:" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


FAIL_SCRIPT = r'''
import os, sys, ctypes as C, torch, torch.distributed as dist
sys.path.insert(0, %(root)r)
import delaunay_linterp_b200 as dl
from delaunay_linterp_b200 import parallel
rank, lr = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
comm = parallel.cached_nccl_communicator()
h = C.c_void_p(None)                       # the root has no mesh to send
rc = dl.lib().dli_mesh_replicate(C.byref(h), comm, 0)
msg = dl.lib().dli_last_error().decode()
assert rc != 0 and not h.value, (rc, h.value)
assert ("must pass its mesh" in msg) if rank == 0 else ("no header received" in msg), msg
t = torch.tensor([1 if rc != 0 else 0], device=dev); dist.all_reduce(t)
if rank == 0:
    print("FAIL_OK", int(t.item()))
dist.barrier(); parallel.shutdown_nccl(); dist.destroy_process_group()
'''


def test_replicate_failure_returns_on_every_rank(gpu, tmp_path):
    """a root that cannot send (no mesh) must not leave the other ranks waiting in ncclBroadcast: it sends an
    empty header, every rank returns an error and nothing hangs."""
    if gpu < 2:
        pytest.skip("needs >= 2 GPUs (run with gpurun --gpus 2)")
    script = tmp_path / "fail.py"
    script.write_text(FAIL_SCRIPT % {"root": ROOT})
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29536", str(script)],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "FAIL_OK 2" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]
