"""Randomised differential run of the CUDA path against the CPU oracle: many small meshes (2D / 3D / 4D), random engine
options, coordinate scales and value functions, adversarial queries (vertices, edge / facet points, lattice points,
points outside the hull).  Same bar as the parity tests: statuses equal, the reported simplex contains the query
(exact check), values within 1e-12 relative.  usage: python scripts/fuzz_parity.py [seconds] [seed]"""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import helpers
import delaunay_linterp_b200 as dl
from oracle.oracle import OracleMesh

budget = float(sys.argv[1]) if len(sys.argv) > 1 else 120.0
seed0 = int(sys.argv[2]) if len(sys.argv) > 2 else 1
OPTS = {
    2: [{}, {"packed_faces": 2}, {"packed_faces": 3}, {"prefetch_ring": -1}, {"epilogue": "reference"}, {"packed_faces": 3, "hint_cells_per_simplex": 1.0},
        {"hint_cells_per_simplex": 0.5}],
    3: [{}, {"start_records": 1}, {"prefetch_ring": -1}, {"epilogue": "reference"}, {"hint_cells_per_simplex": 1.0}],
    4: [{}, {"start_records": -1}, {"packed_faces": 5}, {"prefetch_ring": -1}, {"epilogue": "reference"}, {"hint_cells_per_simplex": 2.0}],
}
t_end = time.time() + budget
it = fails = skipped = 0
worst = 0.0
while time.time() < t_end:
    rng = np.random.default_rng(seed0 * 100003 + it)
    dim = int(rng.choice([2, 2, 3, 4]))
    nv = int(rng.integers(dim + 2, {2: 6000, 3: 2500, 4: 900}[dim]))
    kind = int(rng.integers(0, 4))
    if kind == 0:
        pts = rng.random((nv, dim))
    elif kind == 1:                                   # clustered
        pts = np.concatenate([rng.random((nv // 2, dim)), 0.5 + 0.01 * rng.standard_normal((nv - nv // 2, dim))])
    elif kind == 2:                                   # jittered lattice (near-degenerate)
        g = int(round(nv ** (1.0 / dim))) + 1
        pts = np.stack(np.meshgrid(*([np.arange(g) / (g - 1.0)] * dim), indexing="ij"), -1).reshape(-1, dim)
        pts = pts + 1e-9 * rng.standard_normal(pts.shape)
    else:                                             # anisotropic
        pts = rng.random((nv, dim)) * np.array([1.0, 1e-3, 30.0, 1.0])[:dim]
    scale = float(rng.choice([1.0, 1.0, 1e-6, 3e5, 7.0]))
    shift = float(rng.choice([0.0, 0.0, 1.0, -250.0]))
    pts = np.unique(pts * scale + shift * scale, axis=0)
    if len(pts) < dim + 2:
        it += 1
        continue
    fk = int(rng.integers(0, 4))
    if fk == 0:
        f = helpers.cobb_douglas(np.abs(pts / scale - shift) + 0.1)
    elif fk == 1:
        f = np.sin(3.0 * pts[:, 0] / scale) * np.cos(2.0 * pts[:, 1] / scale)     # sign changes
    elif fk == 2:
        f = 1.0 + pts @ rng.standard_normal(dim) / scale                              # affine
    else:
        f = rng.standard_normal(len(pts)); f[::7] = 0.0
    try:
        sv, sn = helpers.scipy_mesh(pts)
    except Exception:
        it += 1
        continue
    opt = dict(OPTS[dim][int(rng.integers(0, len(OPTS[dim])))])
    lo, hi = pts.min(0), pts.max(0)
    nq = int(rng.integers(1, 40000))
    q = lo + (hi - lo) * (rng.random((nq, dim)) * 1.1 - 0.05)
    k = min(nq, 200)
    q[:k] = pts[rng.integers(0, len(pts), k)]                                    # vertices
    k2 = min(max(nq - k, 0), 200)
    if k2:
        s = sv[rng.integers(0, len(sv), k2)]
        w = rng.random((k2, 2)); w /= w.sum(1, keepdims=True)
        q[k:k + k2] = w[:, :1] * pts[s[:, 0]] + w[:, 1:] * pts[s[:, 1]]          # points on edges (rounded)
    try:
        m = dl.Mesh(pts, f, sv, sn, **opt)
    except dl.DliError as e:                                                      # degenerate Qhull output is rejected: fine
        it += 1
        continue
    ys, sx, st = m.interp(q, return_simplex=True, return_status=True)
    om = OracleMesh(pts, f, sv, sn)
    yo, so, sto = om.interp(q, mode=1, nthreads=8)
    # Is the INPUT a triangulation at all?  Qhull returns thousands of flat, inconsistently linked simplices on jittered
    # lattices far from the origin; on such a mesh the answer depends on where a walk starts, and the oracle disagrees
    # with itself (walk from a fixed face vs walk from the previous result).  Those inputs are skipped.
    if not np.array_equal(om.interp(q, mode=0, nthreads=8)[2], sto):
        skipped += 1
        it += 1
        continue
    bad = []
    if not np.array_equal(st, sto):
        bad.append("status %d differ" % int((st != sto).sum()))
    ok = (sto == 0) & (st == 0)
    cg, _ = om.classify(q, sx)
    if not cg[ok].all():
        bad.append("simplex does not contain the query: %d" % int((~cg[ok]).sum()))
    # values: the 1e-12 relative bar where both sides evaluated the SAME simplex (inside the hull).  A query on a shared
    # facet may be answered from either incident simplex, and a projected query (2D, outside) from a point an ulp apart:
    # equal up to the rounding of the data, not relative to a value that may cross zero -- bar: 1e-9 of max|f| there.
    fin = (sto <= 1) & (st == sto)
    same = fin & (sx == so) & (sto == 0)
    rel = float(np.max(np.abs(ys[same] - yo[same]) / np.maximum(np.abs(yo[same]), 1e-300))) if same.any() else 0.0
    worst = max(worst, rel)
    if not rel <= 1e-12:
        bad.append("rel %.3e on %d same-simplex queries" % (rel, int(same.sum())))
    other = fin & ~same
    if other.any():
        a = float(np.max(np.abs(ys[other] - yo[other]))) / max(1e-300, float(np.abs(f).max()))
        if not a <= 1e-9:
            bad.append("abs/fmax %.3e on %d boundary / projected queries" % (a, int(other.sum())))
    if not np.array_equal(np.isnan(ys), np.isnan(yo)):
        bad.append("nan pattern")
    if bad:
        fails += 1
        print(json.dumps({"iter": it, "seed": seed0, "dim": dim, "nv": len(pts), "kind": kind, "scale": scale, "shift": shift, "f": fk,
                          "opt": opt, "nq": nq, "bad": bad}), flush=True)
    it += 1
print(json.dumps({"iterations": it, "skipped_invalid_inputs": skipped, "failures": fails, "worst_rel": worst, "seconds": budget, "seed": seed0}))
sys.exit(1 if fails else 0)
