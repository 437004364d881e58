"""Native d-dimensional host triangulation (csrc/dli_hostmesh_d.cpp, `dli_dd_*` in include/dli_host.h): stands where the
reference keeps its patched CGAL Delaunay_d (modified_cgal/Delaunay_d.h).  CPU only."""
import math

import numpy as np
import pytest
from scipy.spatial import Delaunay

import delaunay_linterp_b200 as dl


def _simplex_set(t, pts):
    """simplices of the native triangulation as sorted tuples of INPUT row indices (vertex ids follow insertion order)"""
    co, _, sv, sn = t.export()
    idx = {tuple(r): i for i, r in enumerate(pts.tolist())}
    m = np.array([idx[tuple(r)] for r in co.tolist()], dtype=np.int64)
    return set(map(tuple, np.sort(m[sv], axis=1).tolist())), sv, sn


@pytest.mark.parametrize("d,n", [(2, 60), (2, 3000), (3, 60), (3, 3000), (4, 60), (4, 1500)])
@pytest.mark.parametrize("sort", [True, False])
def test_native_triangulation_equals_qhull(d, n, sort):
    """points in general position have ONE Delaunay triangulation: native == Qhull, simplex by simplex"""
    pts = np.random.default_rng(100 * d + n).random((n, d))
    t = dl.Delaunay_incremental_interp_d(d)
    t.insert_many(pts, np.zeros(n), spatial_sort=sort, use_gpu=False)
    assert t.validate() == 0
    have, sv, sn = _simplex_set(t, pts)
    want = set(map(tuple, np.sort(Delaunay(pts).simplices, axis=1).tolist()))
    assert have == want
    # neighbour table: k-th neighbour is opposite vertex k, -1 across the hull, symmetric
    for s in range(0, len(sv), max(1, len(sv) // 500)):
        for k in range(d + 1):
            nb = sn[s, k]
            if nb >= 0:
                assert s in sn[nb]
                assert set(sv[s]) - {sv[s, k]} <= set(sv[nb])


@pytest.mark.parametrize("d,g", [(2, 7), (3, 5), (4, 4)])
def test_regular_grids_are_valid_delaunay(d, g):
    """regular grids (every cell cospherical -- the README example's shape): any valid Delaunay triangulation is
    accepted; the simplices tile the box exactly and satisfy the empty-sphere property"""
    ax = np.linspace(0.01, 5.0, g)
    pts = np.array(np.meshgrid(*([ax] * d), indexing="ij")).reshape(d, -1).T.copy()
    for order in (np.arange(len(pts)), np.random.default_rng(d).permutation(len(pts))):
        t = dl.Delaunay_incremental_interp_d(d)
        for p in pts[order]:
            t.insert(p.tolist(), 1.0)
        assert t.validate() == 0
        co, _, sv, _ = t.export()
        vol = sum(abs(np.linalg.det(co[s[1:]] - co[s[0]])) for s in sv) / math.factorial(d)
        assert abs(vol - 4.99 ** d) < 1e-9 * 4.99 ** d
        assert len(sv) >= (g - 1) ** d * (2 if d == 2 else 5 if d == 3 else 16)


@pytest.mark.parametrize("d", [2, 3, 4])
def test_dimension_jump_duplicates_and_outside_insertions(d):
    """Delaunay_d grows its current dimension point by point (modified_cgal/Delaunay_d.h: dcur); so does this one"""
    t = dl.Delaunay_incremental_interp_d(d)
    assert t.current_dimension() == -1 and t.number_of_simplices() == 0
    e = np.eye(d)
    t.insert([0.0] * d, 1.0)
    assert t.current_dimension() == 0
    t.insert([0.0] * d, 2.0)                       # duplicate before any simplex exists: value overwritten
    assert t.number_of_vertices() == 1
    for k in range(d - 1):
        t.insert(e[k].tolist(), 1.0)
        t.insert((0.5 * e[k]).tolist(), 1.0)       # in the current affine hull: no dimension jump
        assert t.current_dimension() == k + 1 and t.number_of_simplices() == 0
    t.insert(e[d - 1].tolist(), 1.0)
    assert t.current_dimension() == d and t.number_of_simplices() >= 1 and t.validate() == 0
    nv = t.number_of_vertices()
    rng = np.random.default_rng(d)
    for p in rng.normal(size=(200, d)) * 3:        # mostly outside the current hull
        t.insert(p.tolist(), 0.0)
    t.insert(e[0].tolist(), 9.0)                   # duplicate of an existing vertex
    assert t.number_of_vertices() == nv + 200 and t.validate() == 0
    co, f, sv, _ = t.export()
    assert f[np.flatnonzero((co == e[0]).all(1))[0]] == 9.0
    # the axis points are collinear / cospherical (no unique triangulation to compare with): the simplices must tile
    # the convex hull exactly
    from scipy.spatial import ConvexHull
    vol = sum(abs(np.linalg.det(co[s[1:]] - co[s[0]])) for s in sv) / math.factorial(d)
    assert abs(vol - ConvexHull(co).volume) < 1e-9 * vol


def test_near_degenerate_points_need_the_exact_predicates():
    """cospherical and nearly cospherical points: the fp64 filter cannot decide, the expansion fallback must"""
    rng = np.random.default_rng(5)
    for d in (2, 3, 4):
        u = rng.normal(size=(300, d))
        u /= np.linalg.norm(u, axis=1, keepdims=True)        # on the unit sphere up to rounding
        pts = np.vstack([u, np.zeros((1, d)), u[:50] * (1 + 1e-15)])
        pts = np.unique(pts, axis=0)
        t = dl.Delaunay_incremental_interp_d(d)
        for p in pts:
            t.insert(p.tolist(), 0.0)
        assert t.validate() == 0 and t.number_of_vertices() == len(pts)


def test_per_point_interp_is_answered_on_the_host():
    """one point at a time (the reference's own loops) never needs the GPU: exact walk + reference-order epilogue"""
    from oracle.oracle import OracleMesh
    for d in (2, 3, 4):
        rng = np.random.default_rng(20 + d)
        pts = rng.random((400, d))
        fv = np.prod(pts ** np.array([0.3, 0.6, 0.2, 0.1])[:d], axis=1)
        t = dl.Delaunay_incremental_interp_d(d)
        t.insert_many(pts, fv, use_gpu=False)
        co, f, sv, sn = t.export()
        q = rng.random((300, d)) * 1.1 - 0.05
        ys = np.array([t.interp(p.tolist()) for p in q])
        yo, so, sto = OracleMesh(co, f, sv, sn).interp(q)
        inside = sto == 0
        assert np.array_equal(np.isnan(ys), ~inside) and inside.sum() > 100
        assert np.max(np.abs(ys[inside] - yo[inside]) / np.abs(yo[inside])) <= 1e-12
        # vertices (closed simplices: found, not NaN).  The reference-order arithmetic is not exact there in dD: a
        # determinant with two equal columns is only ~1e-17 after the pivoting LU
        assert all(abs(t.interp(p.tolist()) - v) <= 4e-16 * abs(v) for p, v in zip(co[:20], f[:20]))


@pytest.mark.parametrize("d,n", [(2, 120), (3, 120), (4, 70)])
def test_simplex_ids_are_stable_and_compact_across_insertions(d, n):
    """what the incremental device update (dli_mesh_update, dD) relies on: an insertion hands the ids of the simplices it
    removes to the ones it creates and appends the rest -- every simplex that does not contain the new vertex keeps its
    id -- and simplex_out of the per-point path speaks the same ids"""
    rng = np.random.default_rng(300 + d)
    pts = rng.random((n, d))
    t = dl.Delaunay_incremental_interp_d(d)
    prev = None
    stable_steps = 0
    for i, p in enumerate(pts):
        t.insert(p.tolist(), float(p.sum()))
        if t.number_of_simplices() == 0:
            continue
        co, f, sv, sn = t.export()
        assert len(sv) == t.number_of_simplices() and sv.min() >= 0 and sv.max() < len(co)
        assert sn.max() < len(sv) and sn.min() >= -1
        vnew = len(co) - 1
        if prev is not None and len(sv) >= len(prev):
            old = ~(sv == vnew).any(axis=1)                  # rows without the new vertex existed before ...
            ids = np.nonzero(old)[0]
            assert len(ids) == 0 or ids.max() < len(prev)
            assert np.array_equal(np.sort(sv[ids], axis=1), np.sort(prev[ids], axis=1))   # ... under the same id
            stable_steps += 1
        prev = sv
    assert stable_steps > n // 2 and t.validate() == 0
    q = pts[:40] * 0.5 + 0.25
    ys, sx, st = t.interp(q[:8], return_simplex=True, return_status=True)     # 8 queries: answered on the host
    co, f, sv, sn = t.export()
    for k in range(8):
        if st[k] == 0:
            # barycentric coordinates of the query in the reported simplex are all >= 0
            A = np.vstack([co[sv[sx[k]]].T, np.ones(d + 1)])
            lam = np.linalg.solve(A, np.append(q[k], 1.0))
            assert lam.min() > -1e-9


@pytest.mark.parametrize("d", [2, 3, 4])
def test_adaptive_insertion_on_a_general_hull_always_adds_a_vertex(d):
    """On a hull that is not an axis-aligned box the ROUNDED centroid of a hull sub-simplex lands an ulp off the hull
    when inserted; the sub-simplex survives and proposes the same point again.  The reference then returns the same
    tuple forever (src/delaunay_d_interp.h:1028-1034, src/delaunay_2_interp.h:498-504: a no-op value overwrite); here a
    tuple whose point already is a vertex is dropped, so every call refines the mesh.  Both classes."""
    al = np.array([0.3, 0.6, 0.2, 0.1])[:d]
    fn = lambda x: float(np.prod(np.asarray(x) ** al))   # noqa: E731
    pts = np.random.default_rng(3).random((1500, d)) * 4.99 + 0.01
    fv = np.array([fn(p) for p in pts])
    classes = [dl.Delaunay_incremental_interp_d(d, fn)]
    if d == 2:
        classes.append(dl.DelaunayInterp2(fn))
    for t in classes:
        t.insert_many(pts, fv, use_gpu=False)
        nv = lambda: t.number_of_vertices() if hasattr(t, "number_of_vertices") else t.n_vertices   # noqa: E731
        seen = set()
        for r in range(120):
            n0 = nv()
            e = t.insert_largest_error_point()
            assert nv() == n0 + 1, (r, e)
            assert tuple(e[2]) not in seen
            seen.add(tuple(e[2]))
        q = t.get_error_queue()
        assert all(q[i][0] >= q[i + 1][0] for i in range(len(q) - 1))


@pytest.mark.parametrize("d", [2, 3, 4])
def test_dd_refine_k_at_a_time(d):
    """dli_dd_refine: k = 1 reproduces insert_largest_error_point tuple by tuple (same queue afterwards); a batched
    round inserts several points with ONE callback, leaves a valid Delaunay triangulation and a queue equal to a
    from-scratch evaluation of every simplex"""
    al = np.array([0.3, 0.6, 0.2, 0.1])[:d]
    fn = lambda x: float(np.prod(np.asarray(x) ** al))                 # noqa: E731
    fnb = lambda a: np.prod(a ** al, axis=1)                            # noqa: E731
    lo, hi = 0.0, 4.0                                                   # box hull with exactly representable faces
    corners = np.array(np.meshgrid(*([[lo, hi]] * d), indexing="ij")).reshape(d, -1).T
    inner = np.random.default_rng(40 + d).random((60, d)) * 3.0 + 0.5
    pts = np.concatenate([corners, inner])
    fv = np.array([fn(p) for p in pts])

    def fresh():
        t = dl.Delaunay_incremental_interp_d(d, fn)
        for p, v in zip(pts, fv):
            t.insert(p.tolist(), float(v))
        return t
    a, b = fresh(), fresh()
    for _ in range(15):
        ta = a.insert_largest_error_point()
        tb = b.insert_largest_error_points(1, fn_batch=fnb)
        assert len(tb) == 1 and tb[0][2] == ta[2] and tb[0][0] == ta[0]
    qa, qb = a.get_error_queue(), b.get_error_queue()
    assert len(qa) == len(qb) and all(x[0] == y[0] and x[2] == y[2] for x, y in zip(qa, qb))
    calls = []
    c = fresh()
    n0 = c.number_of_vertices()

    def counted(arr):
        calls.append(len(arr))
        return fnb(arr)
    got = c.insert_largest_error_points(8, fn_batch=counted, rel_threshold=0.0)
    assert 1 <= len(got) <= 8 and c.number_of_vertices() == n0 + len(got) and len(calls) == 1
    assert c.validate() == 0
    errs = sorted(e[0] for e in c.get_error_queue())
    # from-scratch evaluation of the same triangulation: insert the same points into a new object
    co = np.array(c.get_all_vertices())
    r = dl.Delaunay_incremental_interp_d(d, fn)
    for row in co:
        r.insert(row[:d].tolist(), float(row[d]))
    ref = sorted(e[0] for e in r.get_error_queue())
    assert len(errs) == len(ref) and np.allclose(errs, ref, rtol=1e-9, atol=1e-18)
