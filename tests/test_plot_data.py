"""Plot data: get_line_segments + is_edge_concave (src/delaunay_2_interp.h:842-862, :766-803).
CPU: the oracle restatement against analytic cases.  GPU: the batched kernel against the oracle, edge by edge."""
import numpy as np
import pytest

import helpers
from oracle.oracle import OracleMesh


def _rows(seg, cc):
    a = np.concatenate([seg.reshape(-1, 6), cc.reshape(-1, 1).astype(np.float64)], axis=1)
    return a[np.lexsort(a.T[::-1])]


def _paraboloid_mesh(sign, n=900, seed=3):
    rng = np.random.default_rng(seed)
    pts = rng.random((n, 2)) * 2.0 - 1.0
    sv, sn = helpers.scipy_mesh(pts)
    return pts, sign * (pts[:, 0] ** 2 + pts[:, 1] ** 2), sv, sn


def test_oracle_edges_euler_and_minmax():
    c, f, t = helpers.grid_mesh(0.0)
    seg, cc, mn, mx = OracleMesh(c, f, t).line_segments()
    assert seg.shape[0] == 100 + 162 - 1                       # Euler: E = V + F - 1
    assert np.array_equal(mn, [0.0, 0.0, 0.0]) and np.allclose(mx, [5.0, 5.0, helpers.cobb_douglas(np.array([[5.0, 5.0]]))[0]])
    und = {tuple(sorted([tuple(s[0][:2]), tuple(s[1][:2])])) for s in seg.tolist()}
    assert len(und) == seg.shape[0]                            # every edge once


def test_oracle_concavity_flags_on_paraboloids():
    """The interpolant of a strictly concave f on a Delaunay mesh... is not concave edge by edge in general, but for
    f = -(x^2 + y^2) the Delaunay triangulation IS the upper hull of the lifted points, so every interior edge is a
    concave crease: flag 1 (both gradient-jump components <= 0, :798); for f = +(x^2 + y^2) every interior edge is a
    convex crease and the flag is 0 unless the edge is axis-parallel in one component; hull edges are always 1."""
    for sign in (-1.0, 1.0):
        pts, f, sv, sn = _paraboloid_mesh(sign)
        seg, cc, mn, mx = OracleMesh(pts, f, sv, sn).line_segments()
        n_hull = int((sn < 0).sum())
        assert seg.shape[0] == (3 * sv.shape[0] + n_hull) // 2
        if sign < 0:
            assert cc.sum() == cc.shape[0]
        else:
            assert cc.sum() == n_hull
        assert np.allclose(mn[:2], pts.min(0)) and np.allclose(mx[:2], pts.max(0))


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["grid", "cobb", "concave", "convex", "noise"])
def test_gpu_line_segments_match_oracle(gpu, case):
    import delaunay_linterp_b200 as dl
    if case == "grid":
        c, f, t = helpers.grid_mesh(0.0)
        sv, sn = t, None
    elif case == "cobb":
        c, f, sv, sn = helpers.random_mesh(2, 30000, 5)
    elif case in ("concave", "convex"):
        c, f, sv, sn = _paraboloid_mesh(-1.0 if case == "concave" else 1.0, 20000, 6)
    else:
        c, _, sv, sn = helpers.random_mesh(2, 30000, 7)
        f = np.random.default_rng(8).standard_normal(c.shape[0])
    m = dl.Mesh(c, f, sv, sn)
    seg, cc, mn, mx = m.line_segments()
    so, co, mno, mxo = OracleMesh(c, f, sv, sn).line_segments()
    assert seg.shape == so.shape
    assert np.array_equal(_rows(seg, cc), _rows(so, co))       # same edges, same end point order, same flags
    assert np.array_equal(mn, mno) and np.array_equal(mx, mxo)
    if case == "noise":
        assert 0.05 < cc.mean() < 0.95                         # the flags carry information


@pytest.mark.gpu
def test_python_class_returns_the_reference_shape(gpu):
    import delaunay_linterp_b200 as dl
    fn = lambda x: x[0] ** 0.3 * x[1] ** 0.6
    t = dl.DelaunayInterp2(fn)
    for p in ([0.01, 0.01], [0.01, 5.0], [5.0, 0.01], [5.0, 5.0]):
        t.insert(p, fn(p))
    for _ in range(96):
        t.insert_largest_error_point()
    all_segments, min_coords, max_coords = t.get_line_segments()          # delaunay_2.py:65
    assert len(all_segments) == 100 + t.n_faces - 1 == len(t.get_edges())
    p, q, flag = all_segments[0]
    assert len(p) == 3 and len(q) == 3 and flag in (0, 1)
    assert min_coords[:2] == [0.01, 0.01] and max_coords[:2] == [5.0, 5.0]
    # the incremental device update keeps the plot data in step
    t.insert_largest_error_point()
    assert len(t.get_line_segments()[0]) == 101 + t.n_faces - 1 and t.sync_counts()[1] >= 1


def test_dd_class_line_segments_shape():
    """src/delaunay_d_interp.h:1103-1142: consecutive vertex pairs of every simplex, once, flag 1, min / max."""
    import delaunay_linterp_b200 as dl
    t = dl.Delaunay_incremental_interp_d(3)
    pts = np.random.default_rng(2).random((40, 3))
    for p in pts:
        t.insert([float(v) for v in p], float(p.sum()))
    segs, lo, hi = t.get_line_segments()
    assert all(len(a) == 4 and len(b) == 4 and fl == 1 for a, b, fl in segs)
    assert np.allclose(lo[:3], pts.min(0)) and np.allclose(hi[:3], pts.max(0))
    keys = {tuple(sorted([tuple(a), tuple(b)])) for a, b, _ in segs}
    assert len(keys) == len(segs)
