"""Throughput with a share of the batch outside the convex hull (2D: projected, src/delaunay_2_interp.h:530-545)."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import helpers
import delaunay_linterp_b200 as dl
pts = np.random.default_rng(1001).random((1000000, 2))
t = dl.DelaunayInterp2(); t.insert_many(pts, helpers.cobb_douglas(pts), use_gpu=False)
m = dl.Mesh(*t.export())
for lo, hi in ((0.02, 0.98), (-0.005, 1.005), (-0.05, 1.05)):
    xq = torch.rand(50000000, 2, dtype=torch.float64, device="cuda") * (hi - lo) + lo
    ys = torch.empty(xq.shape[0], dtype=torch.float64, device="cuda")
    st = torch.empty(xq.shape[0], dtype=torch.uint8, device="cuda")
    m.interp(xq, out=ys); m.interp(xq, out=ys); torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(); m.interp(xq, out=ys); e1.record(); torch.cuda.synchronize()
    _, s = m.interp(xq[:2000000], return_status=True)
    print(json.dumps({"window": [lo, hi], "outside_frac": round(float((s == 1).float().mean()), 4), "ms": round(e0.elapsed_time(e1), 3),
                      "gqps": round(xq.shape[0] / e0.elapsed_time(e1) / 1e6, 2)}))
