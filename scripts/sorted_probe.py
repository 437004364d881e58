"""cfg2 with spatially pre-sorted queries (report the pre-sorted throughput and the sort's own cost)."""
import json, os, sys, time
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
ys = torch.empty(n, dtype=torch.float64, device="cuda")
def timed(fn, reps=5):
    fn(); torch.cuda.synchronize(); ts = []
    for _ in range(reps):
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    return min(ts)
out = {"unsorted_ms": round(timed(lambda: m.interp(xq, out=ys)), 3)}
def sort_queries():
    B = 1024   # tiles per axis; row-major tile order (a Z-curve gives the same locality at this tile size)
    key = (xq[:, 1] * B).to(torch.int32) * B + (xq[:, 0] * B).to(torch.int32)
    order = torch.argsort(key)
    return xq[order], order
t_sort = timed(lambda: sort_queries(), reps=2)
xs, order = sort_queries()
out["sort_ms (torch: key + argsort + gather)"] = round(t_sort, 3)
out["sorted_ms"] = round(timed(lambda: m.interp(xs, out=ys)), 3)
ys2 = torch.empty_like(ys)
out["unpermute_ms (results back to caller order)"] = round(timed(lambda: ys2.index_copy_(0, order, ys), reps=2), 3)
out["gqps_unsorted"] = round(n / out["unsorted_ms"] / 1e6, 2); out["gqps_sorted_kernel_only"] = round(n / out["sorted_ms"] / 1e6, 2)
print(json.dumps(out))
