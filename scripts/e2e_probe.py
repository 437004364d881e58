"""Developer probe: host-pointer path (pinned buffers) end to end vs the raw PCIe copies of the same bytes."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import helpers
import delaunay_linterp_b200 as dl
nq = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000
pts, f, sv, sn = helpers.random_mesh(2, 200000, 5)
m = dl.Mesh(pts, f, sv, sn)
hx = torch.empty((nq, 2), dtype=torch.float64, pin_memory=True); hy = torch.empty(nq, dtype=torch.float64, pin_memory=True)
hx.uniform_(0.02, 0.98)
dx = torch.empty((nq, 2), dtype=torch.float64, device="cuda"); dy = torch.empty(nq, dtype=torch.float64, device="cuda")
for rep in range(4):
    t = time.perf_counter(); m.interp(hx.numpy(), out=hy.numpy()); dt = time.perf_counter() - t
    print("host path rep %d: %.1f ms  %.2f Gq/s" % (rep, dt * 1e3, nq / dt / 1e9), flush=True)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
torch.cuda.synchronize(); t = time.perf_counter()
with torch.cuda.stream(s1): dx.copy_(hx, non_blocking=True)
with torch.cuda.stream(s2): hy.copy_(dy, non_blocking=True)
torch.cuda.synchronize(); dt = time.perf_counter() - t
print("raw copies of the same bytes, both directions at once: %.1f ms" % (dt * 1e3))
yd = m.interp(dx.copy_(hx)); torch.cuda.synchronize()
print("matches device path:", bool(torch.equal(yd.cpu(), hy)))
