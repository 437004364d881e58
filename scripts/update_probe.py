"""Cost of keeping the device mesh in step with an adaptive loop that interleaves insert_largest_error_point and a
batched interp: in-place patch (dli_mesh_update) vs a full re-flatten per round."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import helpers
import delaunay_linterp_b200 as dl

NV = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
DIM = int(sys.argv[2]) if len(sys.argv) > 2 else 2          # 3 / 4: the dD class on the native dD triangulation
ROUNDS = 200 if DIM == 2 else 60
rng = np.random.default_rng(3)
if DIM == 2:
    fn = lambda x: x[0] ** 0.3 * x[1] ** 0.6
    t = dl.DelaunayInterp2(fn)
    pts = rng.random((NV, 2)) * 4.99 + 0.01
    pts[:4] = [[0.01, 0.01], [0.01, 5.0], [5.0, 0.01], [5.0, 5.0]]
    t.insert_many(pts, helpers.cobb_douglas(pts), use_gpu=True)
    nsimp = t.n_faces
else:
    al = np.array([0.3, 0.6, 0.2, 0.1])[:DIM]
    fn = lambda x: float(np.prod(np.asarray(x) ** al))
    t = dl.Delaunay_incremental_interp_d(DIM, fn)
    pts = rng.random((NV, DIM)) * 4.99 + 0.01
    t.insert_many(pts, helpers.cobb_douglas(pts), use_gpu=True)
    nsimp = t.number_of_simplices()
xq = torch.rand(1000000, DIM, dtype=torch.float64, device="cuda") * 4.9 + 0.05
t.interp(xq); torch.cuda.synchronize()
out = {"dim": DIM, "vertices": NV, "simplices": nsimp, "rounds": ROUNDS, "queries_per_round": xq.shape[0]}
for mode in ("patch", "full"):
    ts_sync, ts_total = [], []
    for r in range(ROUNDS):
        t0 = time.perf_counter()
        t.insert_largest_error_point()
        t1 = time.perf_counter()
        if mode == "full":
            t.set_options()          # forces a full re-flatten at the next interp
        t.interp(xq[:1]); torch.cuda.synchronize()      # brings the device mesh in step (1 query)
        t2 = time.perf_counter()
        t.interp(xq); torch.cuda.synchronize()
        t3 = time.perf_counter()
        ts_sync.append(t2 - t1); ts_total.append(t3 - t0)
    out[mode] = {"sync_ms_median": round(1e3 * float(np.median(ts_sync)), 4), "round_ms_median": round(1e3 * float(np.median(ts_total)), 4),
                 "sync_counts": t.sync_counts()}
print(json.dumps(out))
