"""GPU parity: dD (3D, 4D) interp and the per-simplex error evaluator vs the CPU oracle."""
import numpy as np
import pytest

import helpers
from oracle.oracle import OracleMesh

pytestmark = pytest.mark.gpu
REL_TOL = 1e-12


@pytest.mark.parametrize("dim,nv,nq,seed", [(3, 3000, 60000, 61), (4, 800, 30000, 62), (3, 40000, 200000, 63)])
def test_dd_random_mesh(gpu, dim, nv, nq, seed):
    import delaunay_linterp_b200 as dl
    pts, f, sv, sn = helpers.random_mesh(dim, nv, seed)
    rng = np.random.default_rng(seed + 100)
    q = rng.random((nq, dim)) * 1.06 - 0.03            # some outside the hull -> NaN
    q[:100] = pts[:100]                                 # vertices
    q[100:200] = 0.5 * (pts[sv[:100, 0]] + pts[sv[:100, 1]])   # edge midpoints (rounded)
    m = dl.Mesh(pts, f, sv, sn)
    ys, s, st = m.interp(q, return_simplex=True, return_status=True)
    om = OracleMesh(pts, f, sv, sn)
    yo, so, sto = om.interp(q, mode=1, nthreads=8)
    assert np.array_equal(st, sto)
    out = sto == 2
    assert out.sum() > 0 and np.isnan(ys[out]).all() and (s[out] == -1).all()
    inside = ~out
    cg, _ = om.classify(q, s)
    _, strict_o = om.classify(q, so)
    assert cg[inside].all()
    same = s == so
    assert same[inside & strict_o].all()
    rel = np.abs(ys[inside] - yo[inside]) / np.maximum(np.abs(yo[inside]), 1e-300)
    assert rel.max() <= REL_TOL, rel.max()
    assert np.array_equal(ys[inside & same], yo[inside & same])   # same simplex -> same operation order -> bit-exact
    # derived neighbours, closed-form epilogue
    y2 = dl.Mesh(pts, f, sv, None).interp(q)
    assert np.array_equal(ys, y2, equal_nan=True)
    y3 = dl.Mesh(pts, f, sv, sn, epilogue="closed_form").interp(q)
    rel3 = np.abs(y3[inside] - yo[inside]) / np.maximum(np.abs(yo[inside]), 1e-300)
    assert rel3.max() < 1e-8 and np.median(rel3) < 1e-14


@pytest.mark.parametrize("dim", [3, 4])
def test_dd_affine_reproduction(gpu, dim):
    import delaunay_linterp_b200 as dl
    pts, f, sv, sn = helpers.random_mesh(dim, 1500 if dim == 3 else 500, 70 + dim)
    w = np.arange(1, dim + 1, dtype=np.float64)
    m = dl.Mesh(pts, 1.0 + pts @ w, sv, sn)
    q = np.random.default_rng(71).random((20000, dim)) * 0.6 + 0.2
    ys, st = m.interp(q, return_status=True)
    ok = st == 0
    assert ok.mean() > 0.9
    assert np.max(np.abs(ys[ok] - (1.0 + q[ok] @ w))) < 1e-11


@pytest.mark.parametrize("dim", [3, 4])
def test_device_predicate_adversarial_dd(gpu, dim):
    import delaunay_linterp_b200 as dl
    from oracle import pyoracle
    rng = np.random.default_rng(80 + dim)
    n = 400
    pts = rng.random((n, dim + 1, dim))
    w = rng.random((n, dim))
    w /= w.sum(1, keepdims=True)
    pts[:, dim] = np.einsum("nk,nkd->nd", w, pts[:, :dim])      # last point (nearly) in the facet's hyperplane
    pts[:40, dim] = pts[:40, 1]                                  # exact duplicates -> 0
    pts[40:80, :, 0] = 0.5                                       # all points in the plane x0 = 0.5 -> exactly 0
    sign, used = dl.debug_orient(pts)
    want = np.array([pyoracle.orient(p.tolist()) for p in pts], dtype=np.int8)
    assert np.array_equal(sign, want)
    assert used.mean() > 0.5


@pytest.mark.parametrize("dim,nv", [(2, 5000), (3, 1500), (4, 300)])
def test_simplex_error_evaluator(gpu, dim, nv):
    """batched compute_simplex_error (+ boundary sub-simplices) vs the oracle's SVD form."""
    import delaunay_linterp_b200 as dl
    pts, f, sv, sn = helpers.random_mesh(dim, nv, 90 + dim)
    m = dl.Mesh(pts, f, sv, sn)
    cent, item_simplex, item_verts = m.simplex_error_items()
    om = OracleMesh(pts, f, sv, sn)
    o_simplex, o_nv, o_verts, o_cent = om.error_items()
    assert np.array_equal(item_simplex, o_simplex)
    assert np.array_equal(item_verts, o_verts)
    assert np.array_equal(cent, o_cent)                      # same accumulation order -> bit-exact
    phi = helpers.cobb_douglas(cent)
    err = m.simplex_error(phi)
    want = om.error_eval(o_nv, o_verts, o_cent, phi)
    scale = np.maximum(want, 1e-30)
    assert np.max(np.abs(err - want) / scale) < 1e-8
    assert np.median(np.abs(err - want) / scale) < 1e-12
    # hull facets produce sub-simplex items; the largest-error item is what insert_largest_error_point would pick
    assert len(err) > m.n_simplices
    assert int(np.argmax(err)) == int(np.argmax(want))
