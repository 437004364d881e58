"""Developer timing sweep (not the contract bench): config-2-like workload with option variants."""
import argparse, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import helpers
import delaunay_linterp_b200 as dl

ap = argparse.ArgumentParser()
ap.add_argument("--nv", type=int, default=1000000)
ap.add_argument("--nq", type=int, default=100000000)
ap.add_argument("--reps", type=int, default=5)
ap.add_argument("--variants", default="cps1,cps2,cps4,cps8,cf2,cf4")
a = ap.parse_args()
t = time.time()
pts, f, sv, sn = helpers.random_mesh(2, a.nv, 1001)
print("mesh built %.1fs ns=%d" % (time.time() - t, len(sv)), flush=True)
g = torch.Generator(device="cuda"); g.manual_seed(2001)
xq = torch.rand(a.nq, 2, dtype=torch.float64, device="cuda", generator=g) * 0.96 + 0.02
ys = torch.empty(a.nq, dtype=torch.float64, device="cuda")
VAR = {
 "noring": dict(prefetch_ring=-1), "ring_qpw2048": dict(queries_per_warp=2048), "ring_qpw4096": dict(queries_per_warp=4096), "ring_qpw512": dict(queries_per_warp=512),
 "ring_t64": dict(threads_per_block=64), "ring_t256": dict(threads_per_block=256),
 "auto": dict(), "ref": dict(epilogue="reference"), "auto_cps4": dict(hint_cells_per_simplex=4.0), "auto_cps1": dict(hint_cells_per_simplex=1.0),
 "auto_t128": dict(threads_per_block=128), "auto_t64": dict(threads_per_block=64), "auto_t256": dict(threads_per_block=256), "auto_qpw256": dict(queries_per_warp=256), "auto_qpw4096": dict(queries_per_warp=4096),
 "cps1": dict(hint_cells_per_simplex=1.0), "cps2": dict(hint_cells_per_simplex=2.0),
 "cps4": dict(hint_cells_per_simplex=4.0), "cps8": dict(hint_cells_per_simplex=8.0),
 "cf2": dict(hint_cells_per_simplex=2.0, epilogue="closed_form"), "cf4": dict(hint_cells_per_simplex=4.0, epilogue="closed_form"),
 "nopack": dict(hint_cells_per_simplex=2.0, packed_faces=-1), "cf2nopack": dict(hint_cells_per_simplex=2.0, packed_faces=-1, epilogue="closed_form"),
 "cps1cf": dict(hint_cells_per_simplex=1.0, epilogue="closed_form"),
 "t128": dict(hint_cells_per_simplex=2.0, threads_per_block=128),
 "qpw256": dict(hint_cells_per_simplex=2.0, queries_per_warp=256), "qpw512": dict(hint_cells_per_simplex=2.0, queries_per_warp=512),
 "qpw2048": dict(hint_cells_per_simplex=2.0, queries_per_warp=2048), "qpw4096": dict(hint_cells_per_simplex=2.0, queries_per_warp=4096),
}
for name in a.variants.split(","):
    t = time.time()
    m = dl.Mesh(pts, f, sv, sn, **VAR[name])
    tb = time.time() - t
    m.collect_stats(True)
    m.interp(xq[:10000000], out=ys[:10000000]); st = m.stats()
    m.collect_stats(False)
    for _ in range(3): m.interp(xq, out=ys)
    torch.cuda.synchronize()
    ts = []
    for _ in range(a.reps):
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(); m.interp(xq, out=ys); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    best = min(ts); med = sorted(ts)[len(ts)//2]
    print(json.dumps(dict(variant=name, build_s=round(tb,2), mesh_MB=round(m.device_bytes/1e6,1), ms_best=round(best,3), ms_med=round(med,3),
          gqps=round(a.nq/best/1e6,2), roofline_frac=round(a.nq/best*1e3*108/6562.6e9,4),
          steps_per_q=round(st["walk_steps"]/st["queries"],3), max_steps=st["max_walk_steps"], checksum=float(ys.sum()))), flush=True)
    m.close()
