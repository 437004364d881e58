"""CPU: the oracle's gradient() restatement (src/delaunay_2_interp.h:608-642) against exact rationals."""
from fractions import Fraction as Fr

import numpy as np

import helpers
from oracle import pyoracle
from oracle.oracle import OracleMesh, OK, ON_EDGE, ON_VERTEX, OUTSIDE_NAN


def _check_against_exact(c, f, t, q, gs, status):
    for i in range(len(q)):
        ge, hits = pyoracle.gradient_exact(c, f, t, q[i])
        if ge is None:
            assert status[i] == OUTSIDE_NAN and np.isnan(gs[i]).all()
            continue
        want = (OK, ON_EDGE, ON_VERTEX)  # 1 face: interior, a hull edge, or a hull corner with a one-face fan
        if len(hits) == 2:
            assert status[i] in (ON_EDGE, ON_VERTEX)   # interior edge, or a hull vertex with a 2-face fan
        elif len(hits) > 2:
            assert status[i] == ON_VERTEX
        else:
            assert status[i] in want
        for k in range(2):
            ref = float(ge[k])
            assert abs(gs[i, k] - ref) <= 1e-12 * max(abs(ref), abs(float(ge[1 - k])), 1e-300), (i, gs[i], ge)


def test_gradient_grid_vertices_edges_faces():
    # config-1 grid with lo = 0.01 (f > 0 everywhere): queries hit vertices, edges and face interiors
    c, f, t = helpers.grid_mesh(0.01, 5.0, 6)
    g = helpers.linspace_ref(0.01, 5.0, 6)
    q = [[g[i], g[j]] for i in range(6) for j in range(6)]                       # vertices (corner fans are open)
    q += [[(g[i] + g[i + 1]) / 2, g[j]] for i in range(5) for j in range(6)]     # edge midpoints (incl. hull edges)
    q += [[g[i] + 0.3 * (g[i + 1] - g[i]), g[j] + 0.2 * (g[j + 1] - g[j])] for i in range(5) for j in range(5)]
    q += [[-1.0, 2.0], [2.0, 6.0]]                                               # outside: NaN, not projected
    q = np.array(q)
    om = OracleMesh(c, f, t)
    gs, s, st = om.gradient(q)
    assert (st[:36] == ON_VERTEX).all()
    assert (st[-2:] == OUTSIDE_NAN).all() and (s[-2:] == -1).all()
    assert (st == OK).sum() >= 25
    _check_against_exact(c, f, t, q, gs, st)
    # un-hinted mode gives the same answers
    gs0, _, st0 = om.gradient(q, mode=0)
    assert np.array_equal(st0, st) and np.array_equal(gs0[np.isfinite(gs0)], gs[np.isfinite(gs)])


def test_gradient_affine_function_is_constant():
    pts, _, sv, sn = helpers.random_mesh(2, 400, 3)
    f = 2.0 + 3.0 * pts[:, 0] - 1.5 * pts[:, 1]
    om = OracleMesh(pts, f, sv, sn)
    rng = np.random.default_rng(5)
    q = np.vstack([rng.random((500, 2)) * 0.8 + 0.1, pts[:50]])   # interior points + exact vertices
    gs, _, st = om.gradient(q)
    ok = st != OUTSIDE_NAN
    assert ok[:500].all() and (st[500:] == ON_VERTEX).all()
    assert np.abs(gs[ok] - np.array([3.0, -1.5])).max() < 1e-9


def test_gradient_random_mesh_matches_exact():
    pts, f, sv, sn = helpers.random_mesh(2, 60, 9)
    rng = np.random.default_rng(10)
    mids = (pts[sv[:20, 0]] + pts[sv[:20, 1]]) / 2    # edge midpoints (exactly on the edge only if representable)
    q = np.vstack([rng.random((100, 2)), pts[:20], mids])
    om = OracleMesh(pts, f, sv, sn)
    gs, s, st = om.gradient(q)
    _check_against_exact(pts, f, sv, q, gs, st)
