"""Throughput on degenerate batches: lattice queries on a regular-grid mesh (the README example scaled up), where a
large share of the queries sits exactly on vertices / edges and the static filter cannot certify a face."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import helpers
import delaunay_linterp_b200 as dl

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
g = np.linspace(0.0, 5.0, n)
X, Y = np.meshgrid(g, g, indexing="ij")
coords = np.stack([X.ravel(), Y.ravel()], 1)
I, J = np.meshgrid(np.arange(n - 1), np.arange(n - 1), indexing="ij")
a = (I * n + J).ravel(); b = a + n; c = a + 1; d = a + n + 1
tri = np.concatenate([np.stack([a, b, c], 1), np.stack([b, d, c], 1)]).astype(np.int32)
f = helpers.cobb_douglas(coords)
m = dl.Mesh(coords, f, tri)
for name, nq1, off in (("lattice (all on vertices/edges)", 4 * n - 3, 0.0), ("generic", 4 * n - 3, 1e-3 / n)):
    q1 = np.linspace(0.0, 5.0, nq1) * (1 - off) + off
    Q = torch.tensor(np.stack(np.meshgrid(q1, q1, indexing="ij"), -1).reshape(-1, 2), device="cuda")
    Q = Q[torch.randperm(Q.shape[0], device="cuda")]
    ys = torch.empty(Q.shape[0], dtype=torch.float64, device="cuda")
    m.collect_stats(True); m.interp(Q[:2000000], out=ys[:2000000]); st = m.stats(); m.collect_stats(False)
    for _ in range(2): m.interp(Q, out=ys)
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(); m.interp(Q, out=ys); e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    err = float((ys - torch.tensor(helpers.cobb_douglas(Q.cpu().numpy()), device="cuda")).abs().max())
    print(json.dumps({"case": name, "mesh": "%dx%d grid" % (n, n), "queries": Q.shape[0], "ms": round(ms, 3),
                      "gqps": round(Q.shape[0] / ms / 1e6, 3), "stragglers_of_2M": st["straggler_queries"],
                      "exact_fallbacks_of_2M": st["exact_fallbacks"], "max_abs_err_vs_f": err}))
