"""Developer probe for config 5 (16M-vertex 2D mesh): random vs spatially binned queries, hint-grid density.
usage: python scripts/cfg5_probe.py [--nv 16000000] [--nq 125000000] [--mode sweep|ncu]"""
import argparse, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import helpers
import delaunay_linterp_b200 as dl

ap = argparse.ArgumentParser()
ap.add_argument("--nv", type=int, default=16000000)
ap.add_argument("--nq", type=int, default=125000000)
ap.add_argument("--mode", default="sweep")
ap.add_argument("--cps", default="2,0.5,1,4")
ap.add_argument("--tiles", default="8,16,32,64,256")
ap.add_argument("--opts", default="{}")
a = ap.parse_args()
OPTS = json.loads(a.opts)

pts = np.random.default_rng(1005).random((a.nv, 2))
t0 = time.time(); t = dl.DelaunayInterp2(); t.insert_many(pts, helpers.cobb_douglas(pts), use_gpu=False)
c, f, tri, nb = t.export(); del t
print("triangulated %.1fs faces=%d" % (time.time() - t0, len(tri)), flush=True)
g = torch.Generator(device="cuda"); g.manual_seed(2006)
xq = torch.rand(a.nq, 2, dtype=torch.float64, device="cuda", generator=g) * 0.96 + 0.02
ys = torch.empty(a.nq, dtype=torch.float64, device="cuda")

def timeit(m, x, reps=3):
    for _ in range(2): m.interp(x, out=ys)
    torch.cuda.synchronize(); ts = []
    for _ in range(reps):
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(); m.interp(x, out=ys); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    return min(ts)

if a.mode == "ncu":
    m = dl.Mesh(c, f, tri, nb, **OPTS)
    for _ in range(5): m.interp(xq, out=ys)
    torch.cuda.synchronize()
    sys.exit(0)

for cps in [float(v) for v in a.cps.split(",")]:
    t0 = time.time(); m = dl.Mesh(c, f, tri, nb, hint_cells_per_simplex=cps, **OPTS); tf = time.time() - t0
    m.collect_stats(True); m.interp(xq[:4000000], out=ys[:4000000]); st = m.stats(); m.collect_stats(False)
    ms = timeit(m, xq)
    print(json.dumps(dict(opts=OPTS, stragglers=st["straggler_queries"], cps=cps, order="random", ms=round(ms, 3), gqps=round(a.nq / ms / 1e6, 2), steps=round(st["walk_steps"] / st["queries"], 3),
                          mesh_MB=round(m.device_bytes / 1e6), flatten_s=round(tf, 2))), flush=True)
    if cps == 2.0:
        for B in [int(v) for v in a.tiles.split(",") if v]:
            tile = (xq[:, 0] * B).long().clamp_(0, B - 1) + B * (xq[:, 1] * B).long().clamp_(0, B - 1)
            perm = torch.argsort(tile); del tile
            xs = xq[perm].contiguous(); del perm
            ms = timeit(m, xs); del xs
            print(json.dumps(dict(cps=cps, order="binned %dx%d tiles" % (B, B), ms=round(ms, 3), gqps=round(a.nq / ms / 1e6, 2))), flush=True)
    m.close(); del m
