"""CPU context numbers for cfg2, clearly NOT the reference and NOT part of the bench line: on the same
1M-vertex mesh and a 1M-query subsample of the cfg2 query stream, the host cores of this box give
  * the CPU oracle, reference-faithful (un-hinted walk)                 -- what bench.py's cpu_baseline times
  * the same oracle walking from the previous result ("hinted"), on unsorted and on Morton-sorted queries
  * scipy.interpolate.LinearNDInterpolator (Qhull find_simplex, epsilon-tolerant: not an exactness checker)
Developer script: uses oracle/ as a timing subject the way bench.py's cpu_baseline leg does; nothing under
delaunay_linterp_b200/ imports it.   usage: python scripts/cpu_context.py [--nq 1000000] [--out gpurun_out/cpu_context.json]"""
import argparse, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import bench
from oracle.oracle import OracleMesh

ap = argparse.ArgumentParser()
ap.add_argument("--nq", type=int, default=1_000_000)
ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "cpu_context.json"))
a = ap.parse_args()
ncores = os.cpu_count() or 1
res = {"workload": "cfg2 mesh (%d vertices), %d-query subsample, host cores of this box" % (bench.NV, a.nq), "cores": ncores,
       "rows": []}

t0 = time.perf_counter()
pts, f, sv, sn = bench.build_mesh_host_qhull(bench.NV)
res["mesh_build_s"] = round(time.perf_counter() - t0, 2)
om = OracleMesh(pts, f, sv, sn)
q = bench.host_queries(a.nq, bench.QUERY_SEED)


def row(name, fn, n):
    t = time.perf_counter(); out = fn(); dt = time.perf_counter() - t
    r = {"what": name, "queries": n, "seconds": round(dt, 3), "kq_per_s": round(n / dt / 1e3, 1)}
    res["rows"].append(r); print(json.dumps(r), flush=True)
    return out


n0 = min(a.nq, 20000)
row("oracle, un-hinted exact walk (reference-faithful), 1 thread", lambda: om.interp(q[:n0], mode=0, nthreads=1), n0)
n0t = min(a.nq, 20000 * ncores)
row("oracle, un-hinted exact walk, %d threads" % ncores, lambda: om.interp(q[:n0t], mode=0, nthreads=ncores), n0t)
n1 = min(a.nq, 100000)   # a random previous result is as far away as the fixed start face: bounded sample
row("oracle, walk from the previous result, unsorted queries, 1 thread", lambda: om.interp(q[:n1], mode=1, nthreads=1), n1)
row("oracle, walk from the previous result, unsorted queries, %d threads" % ncores,
    lambda: om.interp(q, mode=1, nthreads=ncores), a.nq)
# Morton order of the queries (16 bits per axis): consecutive queries are neighbours, the walk is a few faces long
ij = np.minimum((q * 65536.0).astype(np.uint64), 65535)


def spread(v):
    v = (v | (v << 8)) & 0x00FF00FF; v = (v | (v << 4)) & 0x0F0F0F0F
    v = (v | (v << 2)) & 0x33333333; v = (v | (v << 1)) & 0x55555555
    return v


t = time.perf_counter(); order = np.argsort(spread(ij[:, 0]) | (spread(ij[:, 1]) << 1), kind="stable"); qs = q[order]
res["morton_sort_s"] = round(time.perf_counter() - t, 3)
y1 = row("oracle, walk from the previous result, Morton-sorted queries, 1 thread", lambda: om.interp(qs, mode=1, nthreads=1), a.nq)[0]
row("oracle, walk from the previous result, Morton-sorted queries, %d threads" % ncores,
    lambda: om.interp(qs, mode=1, nthreads=ncores), a.nq)
try:
    from scipy.spatial import Delaunay
    from scipy.interpolate import LinearNDInterpolator
    t = time.perf_counter(); tri = Delaunay(pts); res["scipy_delaunay_build_s"] = round(time.perf_counter() - t, 2)
    li = LinearNDInterpolator(tri, f)
    li(q[:1000])
    row("scipy LinearNDInterpolator, unsorted queries, 1 thread", lambda: li(q[:n1]), n1)
    y2 = row("scipy LinearNDInterpolator, Morton-sorted queries, 1 thread", lambda: li(qs), a.nq)
    res["scipy_vs_oracle_max_rel_diff"] = float(np.nanmax(np.abs(y2 - y1) / np.maximum(np.abs(y1), 1e-300)))
except Exception as e:  # noqa: BLE001
    res["scipy"] = "unavailable (%s)" % type(e).__name__
os.makedirs(os.path.dirname(a.out), exist_ok=True)
with open(a.out, "w") as fh:
    json.dump(res, fh, indent=1)
print(json.dumps(res))
