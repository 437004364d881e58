"""Developer measurements of the other BASELINE configs (3: adaptive 2D, 4: 3D/4D, 5: 16M-vertex 2D)."""
import argparse, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import helpers
import delaunay_linterp_b200 as dl

ap = argparse.ArgumentParser()
ap.add_argument("--cfg", default="3,4a,4b,5")
ap.add_argument("--reps", type=int, default=3)
ap.add_argument("--opts", default="{}")
ap.add_argument("--opts-list", default="", help="JSON list of option dicts: the 4a/4b mesh is built once and measured with each")
a = ap.parse_args()
OPTS = json.loads(a.opts)
BALG = {2: 108, 3: 176, 4: 260}

def timeit(m, xq, ys, reps):
    for _ in range(3): m.interp(xq, out=ys)
    torch.cuda.synchronize(); ts = []
    for _ in range(reps):
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(); m.interp(xq, out=ys); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    return min(ts)

def report(name, m, dim, nq, lo, hi, seed, extra=None):
    g = torch.Generator(device="cuda"); g.manual_seed(seed)
    xq = torch.rand(nq, dim, dtype=torch.float64, device="cuda", generator=g) * (hi - lo) + lo
    ys = torch.empty(nq, dtype=torch.float64, device="cuda")
    m.collect_stats(True); n1 = min(nq, 4000000); m.interp(xq[:n1], out=ys[:n1]); st = m.stats(); m.collect_stats(False)
    ms = timeit(m, xq, ys, a.reps)
    out = dict(cfg=name, dim=dim, nq=nq, ns=int(m.n_simplices), mesh_MB=round(m.device_bytes / 1e6, 1), ms=round(ms, 3), gqps=round(nq / ms / 1e6, 3),
               roofline_frac=round(nq / ms * 1e3 * BALG[dim] / 6538.6e9, 4), steps_per_q=round(st["walk_steps"] / st["queries"], 3),
               max_steps=st["max_walk_steps"], stragglers=st["straggler_queries"], slow_simplices=st["slow_simplices"], nan_frac=float(torch.isnan(ys).float().mean()))
    if extra: out.update(extra)
    print(json.dumps(out), flush=True)

for cfg in a.cfg.split(","):
    if cfg == "3":
        fn = lambda x: x[0] ** 0.3 * x[1] ** 0.6
        t0 = time.time(); t = dl.DelaunayInterp2(fn)
        for p in ([0.01, 0.01], [0.01, 5.0], [5.0, 0.01], [5.0, 5.0]): t.insert(p, fn(p))
        for _ in range(99996): t.insert_largest_error_point()
        tb = time.time() - t0
        c, f, tri, nb = t.export()
        t0 = time.time(); m = dl.Mesh(c, f, tri, nb, **OPTS); tf = time.time() - t0
        # batched error evaluation over all faces + hull edges
        t0 = time.time(); cent, isx, iv = m.simplex_error_items(); phi = helpers.cobb_douglas(cent); err = m.simplex_error(phi); te = time.time() - t0
        report("3 adaptive 100k", m, 2, 50000000, 0.01, 5.0, 2003, dict(build_s=round(tb, 2), flatten_s=round(tf, 3), error_eval_s=round(te, 4), n_items=len(err)))
    elif cfg in ("4a", "4b"):
        dim = 3 if cfg == "4a" else 4
        t0 = time.time()
        pts = np.random.default_rng(1003 if dim == 3 else 1004).random((200000, dim))
        t = dl.Delaunay_incremental_interp_d(dim)
        t.insert_many(pts, helpers.cobb_douglas(pts), spatial_sort=True, use_gpu=False)   # native dD triangulation (host)
        pts, f, sv, sn = t.export(); del t
        tb = time.time() - t0
        for o in (json.loads(a.opts_list) if a.opts_list else [OPTS]):
            t0 = time.time(); m = dl.Mesh(pts, f, sv, sn, **o); tf = time.time() - t0
            report("4 %dD 200k" % dim, m, dim, 20000000, 0.1, 0.9, 2004 if dim == 3 else 2005, dict(build_s=round(tb, 2), flatten_s=round(tf, 3), opts=o))
            del m
    elif cfg == "5":
        pts = np.random.default_rng(1005).random((16000000, 2))
        t0 = time.time(); t = dl.DelaunayInterp2(); t.insert_many(pts, helpers.cobb_douglas(pts), use_gpu=False); tb = time.time() - t0
        c, f, tri, nb = t.export(); del t
        t0 = time.time(); m = dl.Mesh(c, f, tri, nb, **OPTS); tf = time.time() - t0
        report("5 16M-vertex", m, 2, 125000000, 0.02, 0.98, 2006, dict(build_s=round(tb, 2), flatten_s=round(tf, 3)))
