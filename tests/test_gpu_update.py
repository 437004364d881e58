"""GPU: incremental device update after insert (the reference's dirty set is the star returned by
insert_and_get_modifications, src/delaunay_2_interp.h:434-448).  The patched device mesh must answer exactly like a
mesh flattened from scratch from the same triangulation, and like the CPU oracle."""
import numpy as np
import pytest

import helpers
from oracle.oracle import OracleMesh

pytestmark = pytest.mark.gpu


def _queries(n, seed, lo=-0.03, hi=1.03):
    return np.random.default_rng(seed).random((n, 2)) * (hi - lo) + lo


def _check_against_fresh(t, q, dl, bit_exact=False):
    ys, sx, st = t.interp(q, return_status=True, return_simplex=True)
    c, f, tri, nb = t.export()
    fresh = dl.Mesh(c, f, tri, nb, **({"epilogue": 1} if bit_exact else {}))
    yf, sf, stf = fresh.interp(q, return_simplex=True, return_status=True)
    assert np.array_equal(st, stf)
    same = sx == sf
    if bit_exact:   # reference-order epilogue: same face -> same bits.  (In auto mode a query evaluates the face's
        # affine form in the fast pass and the reference order in the straggler pass; which pass it takes depends on
        # its start hint, and two builds of the hint grid need not agree cell by cell: 1e-12 below covers both.)
        assert np.array_equal(ys[same].view(np.uint64), yf[same].view(np.uint64))
    assert same.mean() > 0.999                                                  # others sit exactly on an edge
    rel = np.abs(ys - yf) / np.maximum(np.abs(yf), 1e-300)
    assert np.nanmax(rel) <= 1e-12
    yo, so, sto = OracleMesh(c, f, tri, nb).interp(q, nthreads=4)
    assert np.array_equal(st, sto)
    assert np.nanmax(np.abs(ys - yo) / np.maximum(np.abs(yo), 1e-300)) <= 1e-12
    inside = st == 0
    assert np.array_equal(sx[inside & (sx == so)], so[inside & (sx == so)]) and (sx[inside] == so[inside]).mean() > 0.999
    return fresh


@pytest.mark.parametrize("epilogue", ["auto", "reference"])
def test_patched_mesh_equals_fresh_mesh(gpu, epilogue):
    import delaunay_linterp_b200 as dl
    rng = np.random.default_rng(11)
    pts = rng.random((20000, 2))
    t = dl.DelaunayInterp2(**({} if epilogue == "auto" else {"epilogue": 1}))
    t.insert_many(pts, helpers.cobb_douglas(pts), use_gpu=False)
    q = _queries(200000, 12)
    _check_against_fresh(t, q, dl, epilogue == "reference")
    assert t.sync_counts() == (1, 0)
    lo, hi = pts.min(0), pts.max(0)
    for rnd in range(6):
        new = rng.random((40, 2)) * 0.98 + 0.01                  # interior stars
        if rnd % 2 == 1:                                          # + points outside the hull but inside the hint grid
            corners = np.array([[lo[0], lo[1]], [hi[0], lo[1]], [lo[0], hi[1]], [hi[0], hi[1]]])
            new = np.concatenate([new, corners + (0.5 - (corners > 0.5)) * 2e-4 * (rnd + 1)])
        for p in new:
            t.insert([float(p[0]), float(p[1])], float(helpers.cobb_douglas(p[None, :])[0]))
        _check_against_fresh(t, q, dl, epilogue == "reference")
        assert t.sync_counts() == (1, rnd + 1)                    # patched in place, never re-flattened


def test_value_overwrite_and_growth(gpu):
    import delaunay_linterp_b200 as dl
    rng = np.random.default_rng(21)
    pts = rng.random((1500, 2))
    t = dl.DelaunayInterp2()
    t.insert_many(pts, helpers.cobb_douglas(pts), use_gpu=False)
    q = _queries(50000, 22)
    t.interp(q)
    # overwrite the value of an existing vertex (CGAL insert returns the existing vertex, :446)
    v = 700
    t.insert([float(pts[v, 0]), float(pts[v, 1])], 123.5)
    assert t.interp([float(pts[v, 0]), float(pts[v, 1])]) == 123.5
    _check_against_fresh(t, q, dl)
    assert t.sync_counts() == (1, 1)
    # grow past the initial capacity of the device buffers (3x the faces), refreshing as we go; once half of the mesh
    # has been patched the class re-flattens (the start-hint grid is not patched), so both paths alternate
    for rnd in range(10):
        new = rng.random((300, 2)) * 0.96 + 0.02
        for p in new:
            t.insert([float(p[0]), float(p[1])], float(helpers.cobb_douglas(p[None, :])[0]))
        t.interp(q[:100])
    fresh = _check_against_fresh(t, q, dl)
    assert t.n_vertices == 4500
    n_full, n_patch = t.sync_counts()
    assert 2 <= n_full <= 6 and n_patch >= 6
    # the error evaluator sees the patched connectivity
    m = dl.Mesh._from_handle(dl.lib().dli_di2_device_mesh(t._h), 2)
    try:
        ca, sa, va = m.simplex_error_items()
        cb, sb, vb = fresh.simplex_error_items()
        assert np.array_equal(sa, sb) and np.array_equal(va, vb) and np.array_equal(ca, cb)
    finally:
        m._h = None   # owned by t


def test_unpatchable_updates_fall_back_to_a_full_flatten(gpu):
    import delaunay_linterp_b200 as dl
    rng = np.random.default_rng(31)
    pts = rng.random((3000, 2))
    t = dl.DelaunayInterp2()
    t.insert_many(pts, helpers.cobb_douglas(pts), use_gpu=False)
    q = _queries(20000, 32, -0.5, 1.5)
    t.interp(q)
    t.insert([1.4, -0.3], 2.0)                       # outside the bounding box of the start-hint grid
    _check_against_fresh(t, q, dl)
    assert t.sync_counts() == (2, 0)
    t.insert([0.5, 0.5], 2.0)
    _check_against_fresh(t, q, dl)
    assert t.sync_counts() == (2, 1)


def test_mesh_update_argument_checks(gpu):
    import delaunay_linterp_b200 as dl
    pts, f, sv, sn = helpers.random_mesh(2, 500, 41)
    m = dl.Mesh(pts, f, sv, sn)
    ns = m.n_simplices
    with pytest.raises(dl.DliError):                 # a new simplex id that is not listed
        m.update(np.zeros((0, 2)), [], ns + 1, [0], [sv[0]], [sn[0]])
    with pytest.raises(dl.DliError):                 # vertex index out of range
        m.update(np.zeros((0, 2)), [], ns, [0], [[0, 1, 500]], [sn[0]])
    assert m.update(np.zeros((0, 2)), [], ns, [], np.zeros((0, 3)), np.zeros((0, 3))) is True   # empty update
    assert m.update([[5.0, 5.0]], [1.0], ns, [], np.zeros((0, 3)), np.zeros((0, 3))) is False   # outside the grid
    hdr, gp = m.header(), None
    m3 = dl.Mesh(*helpers.random_mesh(3, 300, 42))
    assert m3.update(np.zeros((0, 3)), [], m3.n_simplices, [], np.zeros((0, 4)), np.zeros((0, 4))) is True   # dD: supported
    p4, f4, sv4, sn4 = helpers.random_mesh(4, 200, 43)
    m4 = dl.Mesh(p4, f4, sv4, sn4, packed_faces=5)
    assert m4.update(np.zeros((0, 4)), [], m4.n_simplices, [], np.zeros((0, 5)), np.zeros((0, 5))) is False  # split 4D records: not patchable
    q = _queries(1000, 43, 0.1, 0.9)
    y0 = m.interp(q)
    assert np.array_equal(y0, dl.Mesh(pts, f, sv, sn).interp(q))


@pytest.mark.parametrize("d,nv", [(3, 4000), (4, 1500)])
def test_dd_patched_mesh_equals_fresh_mesh(gpu, d, nv):
    """dD form of the incremental update (the reference's dirty list: modified_cgal/Convex_hull_d.h:1172-1178 consumed at
    src/delaunay_d_interp.h:939-965): after insert / value overwrite / insert_largest_error_point the device mesh is
    patched (new simplices + the ring around them + the stars of overwritten vertices) and answers like a mesh flattened
    from scratch and like the oracle.  The start copies of 4D meshes are dropped by the first patch."""
    import delaunay_linterp_b200 as dl
    rng = np.random.default_rng(500 + d)
    pts = rng.random((nv, d))
    fn = lambda x: float(np.prod(np.asarray(x) ** np.array([0.3, 0.6, 0.2, 0.1])[:d]) + 1.0)   # noqa: E731
    t = dl.Delaunay_incremental_interp_d(d, fn)
    t.insert_many(pts, np.array([fn(p) for p in pts]), use_gpu=False)
    q = rng.random((120000, d)) * 1.04 - 0.02

    def check():
        ys, sx, st = t.interp(q, return_status=True, return_simplex=True)
        c, f, sv, sn = t.export()
        fresh = dl.Mesh(c, f, sv, sn)
        yf, sf, stf = fresh.interp(q, return_simplex=True, return_status=True)
        assert np.array_equal(st, stf)
        same = sx == sf
        assert same.mean() > 0.999
        ok = st == 0
        rel = np.abs(ys[ok] - yf[ok]) / np.abs(yf[ok])
        assert rel.max() <= 1e-12 and np.array_equal(np.isnan(ys), np.isnan(yf))
        bits = ys[same & ok] == yf[same & ok]
        assert bits.mean() > 0.99          # same simplex, same record builder -> same bits (but for exact-pass stragglers)
        om = OracleMesh(c, f, sv, sn)
        yo, so, sto = om.interp(q, mode=1, nthreads=8)
        assert np.array_equal(st, sto)
        assert np.max(np.abs(ys[ok] - yo[ok]) / np.abs(yo[ok])) <= 1e-12
        cg, _ = om.classify(q, sx)
        assert cg[ok].all()

    check()
    assert t.sync_counts() == (1, 0)
    for rnd in range(4):
        for p in rng.random((12, d)) * 0.9 + 0.05:                  # interior insertions
            t.insert(p.tolist(), fn(p))
        if rnd == 1:                                                 # value overwrite of existing vertices
            for v in (3, 77):
                t.insert(pts[v].tolist(), fn(pts[v]) + 0.25)
        if rnd >= 2:
            for _ in range(5):
                t.insert_largest_error_point()
        check()
        full, patched = t.sync_counts()
        assert full + patched == rnd + 2 and patched >= 1, (full, patched)
        if d == 3:
            assert (full, patched) == (1, rnd + 1)    # patched in place, never re-flattened
        # (4D: an insertion dirties ~800 simplices of 35k; once half of the mesh the start hints were built for has been
        # replaced the class re-flattens on purpose -- in the second round here)
    # a vertex outside the start-hint grid cannot be patched in: full re-flatten, same answers
    full0 = t.sync_counts()[0]
    t.insert([1.5] * d, fn([1.5] * d))
    check()
    assert t.sync_counts()[0] == full0 + 1
