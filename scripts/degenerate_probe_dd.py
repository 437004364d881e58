"""dD counterpart of degenerate_probe.py: lattice queries on a Kuhn-split regular grid (all on vertices / facets)."""
import itertools, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import helpers
import delaunay_linterp_b200 as dl

dim = int(sys.argv[1]) if len(sys.argv) > 1 else 3
n = int(sys.argv[2]) if len(sys.argv) > 2 else (48 if dim == 3 else 14)
g = np.arange(n, dtype=np.float64) / (n - 1)
grid = np.stack(np.meshgrid(*([g] * dim), indexing="ij"), -1).reshape(-1, dim)
cells = np.stack(np.meshgrid(*([np.arange(n - 1)] * dim), indexing="ij"), -1).reshape(-1, dim)
simp = []
for perm in itertools.permutations(range(dim)):
    cur = cells.copy()
    verts = [np.ravel_multi_index(cur.T, (n,) * dim)]
    for ax in perm:
        cur[:, ax] += 1
        verts.append(np.ravel_multi_index(cur.T, (n,) * dim))
    simp.append(np.stack(verts, 1))
simp = np.concatenate(simp).astype(np.int32)
vals = 1.0 + helpers.cobb_douglas(grid + 0.5)
m = dl.Mesh(grid, vals, simp, None)
nq = 4_000_000
rng = np.random.default_rng(1)
lat = rng.integers(0, 4 * (n - 1) + 1, (nq, dim)) / (4.0 * (n - 1))
gen = rng.random((nq, dim))
for name, q in (("lattice", lat), ("generic", gen)):
    Q = torch.tensor(q, device="cuda")
    ys = torch.empty(nq, dtype=torch.float64, device="cuda")
    m.collect_stats(True); m.interp(Q[:1000000], out=ys[:1000000]); st = m.stats(); m.collect_stats(False)
    m.interp(Q, out=ys); torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(); m.interp(Q, out=ys); e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    print(json.dumps({"dim": dim, "case": name, "simplices": int(simp.shape[0]), "queries": nq, "ms": round(ms, 3),
                      "gqps": round(nq / ms / 1e6, 3), "stragglers_of_1M": st["straggler_queries"],
                      "exact_fallbacks_of_1M": st["exact_fallbacks"], "nan": int(torch.isnan(ys).sum())}))
