"""Throughput of the other batched entry points on cfg2 (1M-vertex mesh, 100M device-resident queries):
interp, interp + simplex + status, locate, gradient."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import helpers
import delaunay_linterp_b200 as dl
pts = np.random.default_rng(1001).random((1000000, 2))
t = dl.DelaunayInterp2(); t.insert_many(pts, helpers.cobb_douglas(pts), use_gpu=False)
m = dl.Mesh(*t.export())
n = 100_000_000
g = torch.Generator(device="cuda"); g.manual_seed(2001)
xq = torch.rand(n, 2, dtype=torch.float64, device="cuda", generator=g) * 0.96 + 0.02
def timed(fn, reps=5):
    fn(); fn(); torch.cuda.synchronize(); ts = []
    for _ in range(reps):
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    return min(ts)
ys = torch.empty(n, dtype=torch.float64, device="cuda")
out = {}
for name, fn in (("interp", lambda: m.interp(xq, out=ys)),
                 ("interp+simplex+status", lambda: m.interp(xq, return_simplex=True, return_status=True)),
                 ("locate", lambda: m.locate(xq)),
                 ("gradient", lambda: m.gradient(xq))):
    ms = timed(fn)
    out[name] = {"ms": round(ms, 3), "gqps": round(n / ms / 1e6, 2)}
print(json.dumps(out))
