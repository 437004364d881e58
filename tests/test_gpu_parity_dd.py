"""GPU parity: dD (3D, 4D) interp and the per-simplex error evaluator vs the CPU oracle."""
import numpy as np
import pytest

import helpers
from oracle.oracle import OracleMesh

pytestmark = pytest.mark.gpu
REL_TOL = 1e-12


@pytest.mark.parametrize("epilogue", ["auto", "reference"])   # auto = walk records + guarded affine form (default)
@pytest.mark.parametrize("dim,nv,nq,seed", [(3, 3000, 60000, 61), (4, 800, 30000, 62), (3, 40000, 200000, 63)])
def test_dd_random_mesh(gpu, dim, nv, nq, seed, epilogue):
    import delaunay_linterp_b200 as dl
    pts, f, sv, sn = helpers.random_mesh(dim, nv, seed)
    rng = np.random.default_rng(seed + 100)
    q = rng.random((nq, dim)) * 1.06 - 0.03            # some outside the hull -> NaN
    q[:100] = pts[:100]                                 # vertices
    q[100:200] = 0.5 * (pts[sv[:100, 0]] + pts[sv[:100, 1]])   # edge midpoints (rounded)
    m = dl.Mesh(pts, f, sv, sn, epilogue=epilogue)
    m.collect_stats(True)
    ys, s, st = m.interp(q, return_simplex=True, return_status=True)
    assert m.stats()["walk_steps"] >= (st == 0).sum()
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
    if epilogue == "reference":
        assert np.array_equal(ys[inside & same], yo[inside & same])   # same simplex -> same operation order -> bit-exact
    # derived neighbours, closed-form epilogue
    y2 = dl.Mesh(pts, f, sv, None, epilogue=epilogue).interp(q)
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


@pytest.mark.parametrize("dim", [3, 4])
def test_dd_walk_record_guards(gpu, dim):
    """sign-changing / vanishing f and a structured (Kuhn) grid with exactly cancelling cofactors: the auto engine
    must stay within 1e-12 of the oracle (flagged simplices take the reference-order pass)."""
    import itertools
    import delaunay_linterp_b200 as dl
    pts, _, sv, sn = helpers.random_mesh(dim, 2500 if dim == 3 else 600, 95 + dim)
    f = np.sin(5.0 * pts[:, 0]) * np.cos(3.0 * pts[:, 1]) + 0.1 * pts[:, 2]
    f[::53] = 0.0
    q = np.random.default_rng(96).random((60000, dim)) * 0.9 + 0.05
    for mesh_pts, vals, simp, nbr, qq in [(pts, f, sv, sn, q)]:
        m = dl.Mesh(mesh_pts, vals, simp, nbr)
        ys, s, st = m.interp(qq, return_simplex=True, return_status=True)
        yo, so, sto = OracleMesh(mesh_pts, vals, simp, nbr).interp(qq, mode=1, nthreads=8)
        assert np.array_equal(st, sto)
        ok = sto == 0
        rel = np.abs(ys[ok] - yo[ok]) / np.maximum(np.abs(yo[ok]), 1e-300)
        assert rel.max() <= REL_TOL, rel.max()
    # regular grid, every cell split into d! Kuhn simplices (axis-aligned edges: exact zeros in G)
    n = 6
    g = np.arange(n, dtype=np.float64) / (n - 1)
    grid = np.array(list(itertools.product(*([g] * dim))))
    vid = lambda idx: int(np.ravel_multi_index(idx, (n,) * dim))
    simp = []
    for cell in itertools.product(*([range(n - 1)] * dim)):
        for perm in itertools.permutations(range(dim)):
            cur = list(cell)
            verts = [vid(tuple(cur))]
            for ax in perm:
                cur[ax] += 1
                verts.append(vid(tuple(cur)))
            simp.append(verts)
    simp = np.array(simp, dtype=np.int32)
    vals = 1.0 + helpers.cobb_douglas(grid + 0.5)
    m = dl.Mesh(grid, vals, simp, None)
    qq = np.random.default_rng(97).random((40000, dim))
    qq[:500] = grid[np.random.default_rng(98).integers(0, len(grid), 500)]      # exact vertices
    ys, s, st = m.interp(qq, return_simplex=True, return_status=True)
    om = OracleMesh(grid, vals, simp, None)
    yo, so, sto = om.interp(qq, mode=1, nthreads=8)
    assert np.array_equal(st, sto)
    cg, _ = om.classify(qq, s)
    assert cg[sto == 0].all()
    rel = np.abs(ys - yo) / np.maximum(np.abs(yo), 1e-300)
    assert rel[sto == 0].max() <= REL_TOL, rel[sto == 0].max()


@pytest.mark.parametrize("dim", [3, 4])
def test_device_predicate_on_lattice_points(gpu, dim):
    """exact zeros on regular grids through the device predicate (fixed-point integer path of the exact fallback)."""
    import delaunay_linterp_b200 as dl
    from oracle import pyoracle
    pts = helpers.lattice_predicate_inputs(dim)
    sign, used_exact = dl.debug_orient(pts, on_device=True)
    want = np.array([pyoracle.orient(p.tolist()) for p in pts], dtype=np.int8)
    assert np.array_equal(sign, want)
    assert used_exact[want == 0].all()
