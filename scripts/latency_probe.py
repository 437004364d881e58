"""Per-call latency of interp on the cfg2 mesh as a function of the batch size: host buffers (numpy in, numpy out:
the call a user of the reference's Python module makes) and device-resident tensors.  Wall clock around the call
(time.perf_counter, the call returns after its own stream synchronisation); median of `reps` calls.
usage: python scripts/latency_probe.py [--out gpurun_out/latency.json]"""
import argparse, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import helpers
import delaunay_linterp_b200 as dl

ap = argparse.ArgumentParser()
ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "latency.json"))
ap.add_argument("--sizes", default="1,32,1024,32768,1048576,16777216,100000000")
a = ap.parse_args()
pts = np.random.default_rng(1001).random((1000000, 2))
t = dl.DelaunayInterp2(); t.insert_many(pts, helpers.cobb_douglas(pts), use_gpu=False)
m = dl.Mesh(*t.export())
rng = np.random.default_rng(2001)
rows = []
for n in [int(v) for v in a.sizes.split(",")]:
    xh = rng.random((n, 2)) * 0.96 + 0.02
    xd = torch.from_numpy(xh).cuda(); yd = torch.empty(n, dtype=torch.float64, device="cuda")
    reps = 200 if n <= 32768 else (20 if n <= (1 << 24) else 5)

    def med(fn):
        for _ in range(5): fn()
        torch.cuda.synchronize(); ts = []
        for _ in range(reps):
            t0 = time.perf_counter(); fn(); torch.cuda.synchronize(); ts.append(time.perf_counter() - t0)
        return float(np.median(ts))
    th = med(lambda: m.interp(xh))
    yh = np.empty(n); yh[:] = 0.0
    tho = med(lambda: m.interp(xh, out=yh)) if n >= 1 << 20 else None    # output buffer reused: no fresh pages
    td = med(lambda: m.interp(xd, out=yd))
    r = {"queries": n, "host_buffers_us": round(th * 1e6, 1), "device_resident_us": round(td * 1e6, 1),
         "host_buffers_Mq_per_s": round(n / th / 1e6, 3), "device_resident_Mq_per_s": round(n / td / 1e6, 3)}
    if tho: r["host_buffers_reused_output_us"] = round(tho * 1e6, 1)
    rows.append(r); print(json.dumps(r), flush=True)
# the class-level call the reference's users make in adaptive loops: one point at a time
ti = dl.DelaunayInterp2(lambda x: x[0] ** 0.3 * x[1] ** 0.6)
for p in ([0.01, 0.01], [0.01, 5.], [5., 0.01], [5., 5.]): ti.insert(p, p[0] ** 0.3 * p[1] ** 0.6)
for _ in range(996): ti.insert_largest_error_point()
one = np.array([[1.234, 2.345]])
for _ in range(5): ti.interp(one)
ts = []
for _ in range(200):
    t0 = time.perf_counter(); ti.interp(one); ts.append(time.perf_counter() - t0)
rows.append({"what": "DelaunayInterp2.interp of ONE point, 1000-vertex adaptive mesh (device mesh in sync)",
             "us": round(float(np.median(ts)) * 1e6, 1)})
print(json.dumps(rows[-1]), flush=True)
os.makedirs(os.path.dirname(a.out), exist_ok=True)
json.dump({"workload": "cfg2 mesh, per-call latency by batch size", "rows": rows}, open(a.out, "w"), indent=1)
