"""CPU: the native host triangulation + adaptive queue (include/dli_host.h) that stand where the reference keeps
CGAL, pinned by BOTH README known answers (README.md:130 and :131)."""
import numpy as np

import helpers
from oracle.oracle import OracleMesh


def _fn(x):
    return x[0] ** 0.3 * x[1] ** 0.6


def _tri_key(tri, coords):
    return sorted(tuple(sorted(map(tuple, coords[t].tolist()))) for t in tri)


def test_regular_grid_via_insert_matches_readme():
    """README.md:55-60 + :130: 100 insert() calls on the 10x10 grid, SSE 5.800002; the cocircular tie-break
    (CGAL's symbolic perturbation) yields the (x1,y0)-(x0,y1) diagonals."""
    import delaunay_linterp_b200 as dl
    for lo, want in ((0.0, "5.800002"), (0.01, "1.909975")):
        g = helpers.linspace_ref(lo, 5.0, 10)
        t = dl.DelaunayInterp2()
        for i in range(10):
            for j in range(10):
                t.insert([g[i], g[j]], _fn([g[i], g[j]]))
        c, f, tri, nb = t.export()
        assert t.n_vertices == 100 and t.n_faces == 162
        c2, f2, t2 = helpers.grid_mesh(lo)
        assert _tri_key(tri, c) == _tri_key(t2, c2)
        q = helpers.grid_queries(lo)
        ys, _, st = OracleMesh(c, f, tri, nb).interp(q)
        assert "%.6f" % float(np.sum((helpers.cobb_douglas(q) - ys) ** 2)) == want


def test_insertion_order_independence_on_cocircular_grid():
    import delaunay_linterp_b200 as dl
    g = helpers.linspace_ref(0.0, 5.0, 7)
    pts = [[g[i], g[j]] for i in range(7) for j in range(7)]
    keys = []
    for seed in (0, 1, 2):
        order = np.random.default_rng(seed).permutation(len(pts))
        t = dl.DelaunayInterp2()
        for k in order:
            t.insert(pts[k], 0.0)
        c, f, tri, nb = t.export()
        keys.append(_tri_key(tri, c))
    assert keys[0] == keys[1] == keys[2]


def test_adaptive_insertion_matches_readme():
    """README.md:62-75 + :131: 4 corners + 96 x insert_largest_error_point, 'adaptive grid: sum of squared
    errors: 0.044019'."""
    import delaunay_linterp_b200 as dl
    a = dl.DelaunayInterp2(_fn)
    for p in ([0.0, 0.0], [0.0, 5.0], [5.0, 0.0], [5.0, 5.0]):
        a.insert(p, _fn(p))
    # 2 faces + 4 hull edges + 1 stale entry: the 4th corner is inserted OUTSIDE the hull of the first three, and
    # the face behind the formerly-hull hypotenuse is not in the new vertex's star, so its edge entry survives --
    # exactly what the reference does (src/delaunay_2_interp.h:459-474 erases star faces only); the README's
    # 0.044019 depends on it (its first adaptive point is that stale midpoint (2.5, 2.5)).
    q0 = a.get_error_queue()
    assert len(q0) == 2 + 4 + 1 and q0[0][2] == [2.5, 2.5]
    top = a.get_largest_error_tuple()
    got = a.insert_largest_error_point()
    assert got == top
    for _ in range(95):
        a.insert_largest_error_point()
    assert a.n_vertices == 100
    c, f, tri, nb = a.export()
    q = helpers.grid_queries(0.0)
    ys, _, _ = OracleMesh(c, f, tri, nb).interp(q)
    assert "%.6f" % float(np.sum((helpers.cobb_douglas(q) - ys) ** 2)) == "0.044019"
    # queue order is non-increasing and its entries belong to the current triangulation
    errs = [e[0] for e in a.get_error_queue()]
    assert all(errs[i] >= errs[i + 1] for i in range(len(errs) - 1))
    n_hull = int((nb < 0).sum())
    assert len(errs) == a.n_faces + n_hull


def test_random_points_match_qhull():
    import delaunay_linterp_b200 as dl
    from scipy.spatial import Delaunay
    pts = np.random.default_rng(5).random((5000, 2))
    for sort in (True, False):
        t = dl.DelaunayInterp2()
        t.insert_many(pts, np.zeros(len(pts)), spatial_sort=sort)
        c, f, tri, nb = t.export()
        idx = {tuple(p): i for i, p in enumerate(pts.tolist())}
        mine = sorted(tuple(sorted(idx[tuple(c[v])] for v in tr)) for tr in tri.tolist())
        ref = sorted(tuple(sorted(tr)) for tr in Delaunay(pts).simplices.tolist())
        assert mine == ref
        OracleMesh(c, f, tri, nb)            # valid orientation / manifold
        # neighbour table is symmetric
        for s in range(0, len(tri), 97):
            for k in range(3):
                o = nb[s, k]
                if o >= 0:
                    assert s in nb[o]


def test_degenerate_sequences():
    import delaunay_linterp_b200 as dl
    t = dl.DelaunayInterp2()
    assert t.dimension == -1
    t.insert([0.0, 0.0], 1.0)
    t.insert([1.0, 1.0], 2.0)
    t.insert([2.0, 2.0], 3.0)            # collinear
    t.insert([1.0, 1.0], 5.0)            # duplicate: value overwritten
    assert t.dimension == 1 and t.n_vertices == 3 and t.n_faces == 0
    t.insert([0.5, 0.5], 4.0)            # still collinear
    t.insert([0.0, 1.0], 0.0)            # first non-collinear point -> fan of 3 triangles
    assert t.dimension == 2 and t.n_faces == 3
    t.insert([1.5, 1.5], 9.0)            # on a hull edge
    t.insert([3.0, 3.0], 7.0)            # collinear with the hull line, outside
    t.insert([-1.0, -5.0], 7.0)          # far outside
    c, f, tri, nb = t.export()
    assert f[1] == 5.0
    OracleMesh(c, f, tri, nb)
    area = 0.0
    for a, b, cc in tri:
        area += 0.5 * ((c[b, 0] - c[a, 0]) * (c[cc, 1] - c[a, 1]) - (c[cc, 0] - c[a, 0]) * (c[b, 1] - c[a, 1]))
    from scipy.spatial import ConvexHull
    assert abs(area - ConvexHull(c).volume) < 1e-12


def test_export_ids_are_stable_and_compact_across_insertions():
    """an insertion changes only the ids of its own star (ids of removed faces are handed to new
    ones, the rest appended), the id space stays 0..n_faces-1 and the export stays a consistent triangulation."""
    import delaunay_linterp_b200 as dl
    rng = np.random.default_rng(5)
    t = dl.DelaunayInterp2()
    pts = rng.random((400, 2))
    t.insert_many(pts, helpers.cobb_douglas(pts), use_gpu=False)
    c0, f0, tri0, nb0 = t.export()
    extra = np.concatenate([rng.random((150, 2)), rng.random((30, 2)) * 3.0 - 1.0])   # inside and outside the hull
    for p in extra:
        c_prev, _, tri_prev, _ = t.export()
        t.insert([float(p[0]), float(p[1])], 1.0)
        c, f, tri, nb = t.export()
        assert tri.shape[0] == t.n_faces and tri.shape[0] > tri_prev.shape[0]
        same = np.all(tri[: tri_prev.shape[0]] == tri_prev, axis=1)
        changed = int((~same).sum()) + tri.shape[0] - tri_prev.shape[0]
        star = int(np.sum(np.any(tri == c.shape[0] - 1, axis=1)))
        assert changed == star                   # exactly the faces of the new vertex's star carry new content
    # neighbour links consistent: nb[t][k] shares the edge opposite vertex k
    for s in range(tri.shape[0]):
        for k in range(3):
            n = nb[s, k]
            if n >= 0:
                e = {tri[s, (k + 1) % 3], tri[s, (k + 2) % 3]}
                assert e <= set(tri[n].tolist()) and s in nb[n]
    sv, _ = helpers.scipy_mesh(c)
    assert _tri_key(tri, c) == _tri_key(sv, c)


def _corners(t):
    for p in ([0.01, 0.01], [0.01, 5.0], [5.0, 0.01], [5.0, 5.0]):
        t.insert(p, _fn(p))


def test_refine_with_k_1_is_insert_largest_error_point():
    """the batched refinement with k = 1 reproduces the sequential loop tuple by tuple and leaves
    the same queue (same entries, same order)."""
    import delaunay_linterp_b200 as dl
    a, b = dl.DelaunayInterp2(_fn), dl.DelaunayInterp2(_fn)
    _corners(a); _corners(b)
    for _ in range(150):
        ta = a.insert_largest_error_point()
        tb = b.insert_largest_error_points(1)
        assert [ta] == tb
    assert a.get_error_queue() == b.get_error_queue()
    assert np.array_equal(a.export()[2], b.export()[2])


def test_refine_k_at_a_time_batched_callback():
    import delaunay_linterp_b200 as dl
    calls = []

    def fb(x):
        calls.append(x.shape[0])
        return x[:, 0] ** 0.3 * x[:, 1] ** 0.6

    t = dl.DelaunayInterp2(_fn)
    _corners(t)
    n_ins = rounds = 0
    while n_ins < 1500:
        got = t.insert_largest_error_points(64, fn_batch=fb)
        assert 1 <= len(got) <= 64
        errs = [g[0] for g in got]
        assert errs == sorted(errs, reverse=True)          # taken in queue order
        assert errs[-1] >= 0.25 * errs[0]                  # default rel_threshold
        n_ins += len(got)
        rounds += 1
    assert len(calls) == rounds and t.n_vertices == 4 + n_ins   # ONE callback per round
    assert rounds * 8 < n_ins                                   # rounds really are batched
    # the triangulation is the Delaunay triangulation of its vertices, and the queue holds exactly the entries a
    # from-scratch evaluation of this triangulation gives (every face + every hull edge, nothing stale)
    c, f, tri, nb = t.export()
    sv, _ = helpers.scipy_mesh(c)
    assert _tri_key(tri, c) == _tri_key(sv, c)
    q = sorted((round(e, 18), round(x[0], 12), round(x[1], 12)) for e, _, x in t.get_error_queue())
    u = dl.DelaunayInterp2(_fn)
    u.insert_many(c, f, use_gpu=False)
    qu = sorted((round(e, 18), round(x[0], 12), round(x[1], 12)) for e, _, x in u.get_error_queue())
    assert len(q) == len(qu) == tri.shape[0] + int((nb < 0).sum())
    assert np.allclose([a[0] for a in q], [a[0] for a in qu], rtol=1e-9, atol=1e-300)
    # quality: same budget of points, error of the same order as the one-at-a-time loop
    s = dl.DelaunayInterp2(_fn)
    _corners(s)
    for _ in range(n_ins):
        s.insert_largest_error_point()
    q = np.random.default_rng(1).random((20000, 2)) * 4.99 + 0.01
    sse = []
    for m in (t, s):
        cm, fm, trim, nbm = m.export()
        ys, _, _ = OracleMesh(cm, fm, trim, nbm).interp(q, nthreads=4)
        sse.append(float(np.sum((helpers.cobb_douglas(q) - ys) ** 2)))
    assert sse[0] <= 1.5 * sse[1], sse


def test_one_point_host_path_honours_outside_nan():
    """a handle configured with outside_nan=1 answers NaN / OUTSIDE_NAN for a query outside the hull whatever
    the batch size -- the one-point host path used to project it (src/delaunay_d_interp.h:1005-1008 is the behaviour the
    option stands for; src/delaunay_2_interp.h:530-545 the default).  No GPU involved: batches this small stay on the host."""
    import delaunay_linterp_b200 as dl
    g = helpers.linspace_ref(0.0, 5.0, 6)
    pts = np.array([[a, b] for a in g for b in g])
    outside = np.array([[-1.0, 2.0], [6.0, 7.0], [2.5, -0.5]])
    inside = np.array([[1.3, 2.2], [5.0, 5.0], [0.0, 2.5]])
    for nan in (0, 1):
        t = dl.DelaunayInterp2(outside_nan=nan) if nan else dl.DelaunayInterp2()
        t.insert_many(pts, helpers.cobb_douglas(pts), use_gpu=False)
        y, s, st = t.interp(outside, return_status=True, return_simplex=True)
        if nan:
            assert np.isnan(y).all() and (st == dl.STATUS_OUTSIDE_NAN).all() and (s == -1).all()
        else:
            assert np.isfinite(y).all() and (st == dl.STATUS_OUTSIDE_PROJECTED).all() and (s >= 0).all()
        yi, sti = t.interp(inside, return_status=True)
        assert (sti == 0).all() and np.allclose(yi, helpers.cobb_douglas(inside), atol=0.5)
    ya = dl.DelaunayInterp2()
    ya.insert_many(pts, helpers.cobb_douglas(pts), use_gpu=False)
    assert np.array_equal(ya.interp(inside), yi)      # inside the hull the option changes nothing
