"""The reference-facing classes (DelaunayInterp2, Delaunay_incremental_interp_d): host logic on CPU, interp on GPU."""
import math

import numpy as np
import pytest

import helpers
from oracle.oracle import OracleMesh


def _fn3(x):
    return x[0] ** 0.3 * x[1] ** 0.6 * x[2] ** 0.2


def test_dd_class_host_side_queue():
    """src/delaunay_d_test2.cpp-style adaptive loop in 3D: corners of the cube, then insert_largest_error_point."""
    import delaunay_linterp_b200 as dl
    t = dl.Delaunay_incremental_interp_d(3, _fn3)
    assert t.get_dimension() == 3
    corners = [[a, b, c] for a in (0.01, 2.0) for b in (0.01, 2.0) for c in (0.01, 2.0)]
    for p in corners:
        t.insert(p, _fn3(p))
    q = t.get_error_queue()
    assert len(q) > 0 and all(q[i][0] >= q[i + 1][0] for i in range(len(q) - 1))
    top = t.get_largest_error_tuple()
    got = t.insert_largest_error_point()
    assert got == top and len(got[2]) == 3
    for _ in range(20):
        t.insert_largest_error_point()
    pts, f, sv, sn = t.export()
    assert len(pts) == 8 + 21
    # every live entry belongs to a current simplex; the mesh is valid for the oracle
    OracleMesh(pts, f, sv, sn)
    # per-insert host tuples == the oracle's SVD-form tuples of the final mesh (as multisets of errors)
    om = OracleMesh(pts, f, sv, sn)
    o_simplex, o_nv, o_verts, o_cent = om.error_items()
    want = np.sort(om.error_eval(o_nv, o_verts, o_cent, np.array([_fn3(c) for c in o_cent])))
    have = np.sort(np.array([e[0] for e in t.get_error_queue()]))
    assert len(want) == len(have)
    assert np.allclose(want, have, rtol=1e-8, atol=1e-18)
    assert dl.get_lifted_point([1.0, 2.0]) == [1.0, 2.0, 5.0]


@pytest.mark.gpu
def test_interp2_class_on_gpu(gpu):
    import delaunay_linterp_b200 as dl
    fn = lambda x: x[0] ** 0.3 * x[1] ** 0.6
    a = dl.DelaunayInterp2(fn)
    for p in ([0.0, 0.0], [0.0, 5.0], [5.0, 0.0], [5.0, 5.0]):
        a.insert(p, fn(p))
    for _ in range(96):
        a.insert_largest_error_point()
    q = helpers.grid_queries(0.0)
    ys = a.interp(q)
    assert "%.6f" % float(np.sum((helpers.cobb_douglas(q) - ys) ** 2)) == "0.044019"    # README.md:131
    assert a.interp([2.5, 2.5]) == ys[np.flatnonzero((q == [2.5, 2.5]).all(1))[0]] if ((q == [2.5, 2.5]).all(1)).any() else True
    assert isinstance(a.interp([1.0, 1.0]), float)
    # interleaving insert and interp re-flattens lazily
    y0 = a.interp([1.234, 2.345])
    a.insert([1.234, 2.345], 7.0)
    assert a.interp([1.234, 2.345]) == 7.0 and y0 != 7.0
    # empty triangulation: the scalar form raises like the reference (std::invalid_argument)
    with pytest.raises(ValueError):
        dl.DelaunayInterp2().interp([0.5, 0.5])
    # bulk build + GPU-batched queue rebuild == per-insert host queue
    pts = np.random.default_rng(3).random((3000, 2)) * 4.9 + 0.05
    b1, b2 = dl.DelaunayInterp2(fn), dl.DelaunayInterp2(fn)
    b1.insert_many(pts, helpers.cobb_douglas(pts), use_gpu=True)
    b2.insert_many(pts, helpers.cobb_douglas(pts), use_gpu=False)
    e1 = np.array([e[0] for e in b1.get_error_queue()])
    e2 = np.array([e[0] for e in b2.get_error_queue()])
    assert len(e1) == len(e2) and np.allclose(e1, e2, rtol=1e-10, atol=1e-20)
    assert b1.get_largest_error_tuple()[2] == b2.get_largest_error_tuple()[2]


@pytest.mark.gpu
def test_dd_class_on_gpu(gpu):
    import delaunay_linterp_b200 as dl
    t = dl.Delaunay_incremental_interp_d(3, _fn3)
    corners = [[a, b, c] for a in (0.01, 2.0) for b in (0.01, 2.0) for c in (0.01, 2.0)]
    for p in corners:
        t.insert(p, _fn3(p))
    for _ in range(40):
        t.insert_largest_error_point()
    assert math.isnan(t.interp([3.0, 3.0, 3.0]))                      # outside: NaN (:1005-1008)
    q = np.random.default_rng(1).random((5000, 3)) * 1.9 + 0.05
    ys = t.interp(q)
    pts, f, sv, sn = t.export()
    yo, _, _ = OracleMesh(pts, f, sv, sn).interp(q)
    assert np.max(np.abs(ys - yo) / np.maximum(np.abs(yo), 1e-300)) <= 1e-12
    before = sorted(e[0] for e in t.get_error_queue())
    t.rebuild_error_queue()                                            # GPU batch == per-insert host tuples
    after = sorted(e[0] for e in t.get_error_queue())
    assert len(before) == len(after) and np.allclose(before, after, rtol=1e-8, atol=1e-18)


@pytest.mark.gpu
def test_interp2_class_gradient_on_gpu(gpu):
    """DelaunayInterp2.gradient (src/delaunay_2_python.cpp:231 -> src/delaunay_2_interp.h:608-642)."""
    import delaunay_linterp_b200 as dl
    fn = lambda x: 1.0 + 2.0 * x[0] - 0.5 * x[1]
    a = dl.DelaunayInterp2(fn)
    pts = np.random.default_rng(8).random((500, 2))
    for p in pts.tolist() + [[0.0, 0.0], [0.0, 1.0], [1.0, 0.0], [1.0, 1.0]]:
        a.insert(p, fn(p))
    g = a.gradient([0.4, 0.6])
    assert isinstance(g, list) and len(g) == 2 and abs(g[0] - 2.0) < 1e-9 and abs(g[1] + 0.5) < 1e-9
    assert all(math.isnan(v) for v in a.gradient([2.0, 2.0]))            # outside the hull: NaN, not projected
    gs, st = a.gradient(np.vstack([pts[:10], [[0.5, 0.5]]]), return_status=True)
    assert (st[:10] == dl.STATUS_ON_VERTEX).all() and np.abs(gs - [2.0, -0.5]).max() < 1e-8
    # parity with the oracle on the exported triangulation
    c, f, t, nb = a.export()
    q = np.random.default_rng(9).random((20000, 2)) * 1.2 - 0.1
    gq, sq = a.gradient(q, return_status=True)
    go, _, so = OracleMesh(c, f, t, nb).gradient(q)
    assert np.array_equal(sq, so) and np.array_equal(gq[so != 2], go[so != 2])


def test_reference_module_names_resolve():
    """`import _delaunay_2_python` / `import _delaunay_d_python` (delaunay_2.py:22, delaunay_d_test.py:4) and the
    reference's own test1() body (delaunay_d_test.py:34-44) without the plotting"""
    import _delaunay_2_python as _delaunay2
    import _delaunay_d_python as _delaunayd
    fn = lambda x: x[0] ** 0.3 * x[1] ** 0.6
    triang = _delaunayd.Delaunay_incremental_interp_d(2, fn)
    for x in [[0.01, 0.01], [0.01, 3.0], [3.0, 0.01], [3.0, 3.0]]:
        triang.insert(x, fn(x))
    for _ in range(50):
        triang.insert_largest_error_point()
    segs, lo, hi = triang.get_line_segments()
    assert len(triang.get_all_vertices()) == 54 and len(segs) > 54 and lo[0] == 0.01 and hi[1] == 3.0
    assert triang.fn is fn and triang.triang.dimension() == 2
    assert abs(triang.interp([1.0, 2.0]) - fn([1.0, 2.0])) < 0.05
    assert math.isnan(triang.interp([5.0, 5.0]))
    t2 = _delaunay2.DelaunayInterp2(fn)
    for x in [[0.01, 0.01], [0.01, 3.0], [3.0, 0.01], [3.0, 3.0]]:
        t2.insert(x, fn(x))
    for _ in range(50):
        t2.insert_largest_error_point()
    # one point at a time is answered on the host: no GPU needed; outside the hull is projected (2D class)
    assert abs(t2.interp([1.0, 2.0]) - fn([1.0, 2.0])) < 0.05
    assert t2.interp([5.0, 2.0]) == t2.interp([3.0, 2.0])
