"""GPU parity: CUDA path (through the C ABI) vs the CPU oracle, 2D."""
import numpy as np
import pytest

import helpers
from oracle.oracle import OracleMesh

pytestmark = pytest.mark.gpu

REL_TOL = 1e-12  # north_star: values within 1e-12 relative in fp64


# engine variants: auto = 64-byte walk records + guarded gradient form + prefetch ring (default); reference = reference
# operation order everywhere (bit-exact vs the oracle); delta = the neighbour encoding meshes with >= 2^21-2 faces
# get; noring = walk records without the cp.async prefetch ring; cells = cell-indexed walk records (the layout meshes
# with >= 8M faces get: start record addressed by the hint cell)
VARIANTS = {"auto": {}, "reference": {"epilogue": "reference"}, "delta": {"packed_faces": 2}, "noring": {"prefetch_ring": -1},
            "cells": {"packed_faces": 3}, "cells_noring": {"packed_faces": 3, "prefetch_ring": -1}}
EPILOGUES = list(VARIANTS)


def _check_parity(mesh, om, q, ys, simplex, status, allow_outside=True, bit_exact=True):
    yo, so, sto = om.interp(q, mode=1, nthreads=8)
    assert np.array_equal(status, sto), "status codes differ"
    inside = sto == 0
    # located simplex: identical for strictly interior queries, any incident simplex otherwise
    contains, strict_o = om.classify(q, so)
    contains_g, _ = om.classify(q, simplex)
    assert contains_g[inside].all(), "GPU returned a simplex that does not contain the query"
    same = simplex == so
    assert same[inside & strict_o].all(), "interior query located in a different simplex"
    # values
    scale = np.maximum(np.abs(yo), 1e-300)
    rel = np.abs(ys - yo) / scale
    assert np.all(rel[np.isfinite(yo)] <= REL_TOL), "max rel diff %g" % rel[np.isfinite(yo)].max()
    # where the simplex is identical the arithmetic is the same operation order: bit-exact
    bit = same & np.isfinite(yo)
    if bit_exact:
        assert np.array_equal(ys[bit], yo[bit])
    return yo, so, sto


@pytest.mark.parametrize("epilogue", EPILOGUES)
def test_config1_readme_grid(gpu, epilogue):
    import delaunay_linterp_b200 as dl
    for lo, want in ((0.0, 5.800002), (0.01, 1.909975)):
        c, f, t = helpers.grid_mesh(lo)
        q = helpers.grid_queries(lo)
        m = dl.Mesh(c, f, t, **VARIANTS[epilogue])
        ys, s, st = m.interp(q, return_simplex=True, return_status=True)
        sse = float(np.sum((helpers.cobb_douglas(q) - ys) ** 2))
        assert abs(sse - want) < 5e-7, sse   # README.md:130 prints 5.800002
        om = OracleMesh(c, f, t)
        _check_parity(m, om, q, ys, s, st, bit_exact=(epilogue == "reference"))


@pytest.mark.parametrize("epilogue", EPILOGUES)
@pytest.mark.parametrize("nv,nq,seed", [(1000, 20000, 11), (100000, 400000, 12)])
def test_random_mesh_inside_and_outside(gpu, nv, nq, seed, epilogue):
    import delaunay_linterp_b200 as dl
    pts, f, sv, sn = helpers.random_mesh(2, nv, seed)
    rng = np.random.default_rng(seed + 100)
    q = rng.random((nq, 2)) * 1.1 - 0.05          # ~17% outside the hull
    # add exact vertices, edge midpoints and far-away points
    q[:200] = pts[:200]
    q[200:400] = 0.5 * (pts[sv[:200, 0]] + pts[sv[:200, 1]])
    q[400:410] = [[-3.0, 0.5], [4.0, 0.5], [0.5, -7.0], [0.5, 9.0], [-2, -2], [3, 3], [-2, 3], [3, -2], [0.5, 1e6], [-1e6, 0.5]]
    m = dl.Mesh(pts, f, sv, sn, **VARIANTS[epilogue])
    m.collect_stats(True)
    ys, s, st = m.interp(q, return_simplex=True, return_status=True)
    stats = m.stats()
    om = OracleMesh(pts, f, sv, sn)
    _check_parity(m, om, q, ys, s, st, bit_exact=(epilogue == "reference"))
    assert stats["outside"] == int((st == 1).sum())
    assert stats["walk_steps"] >= nq
    # neighbours derived by the library give the same answers
    m2 = dl.Mesh(pts, f, sv, None, **VARIANTS[epilogue])
    ys2 = m2.interp(q)
    assert np.array_equal(ys, ys2, equal_nan=True)
    # locate-only agrees with interp's simplex
    s2 = m.locate(q)
    assert np.array_equal(s, s2)


def test_device_pointer_path_matches_host_path(gpu):
    import torch
    import delaunay_linterp_b200 as dl
    pts, f, sv, sn = helpers.random_mesh(2, 5000, 21)
    q = np.random.default_rng(22).random((100000, 2))
    m = dl.Mesh(pts, f, sv, sn)
    ys_h = m.interp(q)
    xq = torch.from_numpy(q).cuda()
    ys_d, s_d, st_d = m.interp(xq, return_simplex=True, return_status=True)
    torch.cuda.synchronize()
    assert np.array_equal(ys_h, ys_d.cpu().numpy(), equal_nan=True)


@pytest.mark.parametrize("dim", [2, 3])
def test_pageable_host_buffers_take_the_staged_path(gpu, dim):
    """numpy arrays are pageable memory: batches of >= 2^18 queries go through the pinned staging slots with one
    worker thread per slot (several chunks, ragged last chunk, queries outside the hull included); every output must
    equal the device-pointer path bit for bit, for interp, simplex, status and (2D) gradient."""
    import torch
    import delaunay_linterp_b200 as dl
    pts, f, sv, sn = helpers.random_mesh(dim, 5000 if dim == 2 else 2000, 23)
    n = (1 << 19) * 3 + 12345
    q = np.random.default_rng(24).random((n, dim)) * 1.04 - 0.02
    m = dl.Mesh(pts, f, sv, sn)
    xq = torch.from_numpy(q).cuda()
    y_d, s_d, t_d = m.interp(xq, return_simplex=True, return_status=True)
    y_h, s_h, t_h = m.interp(q, return_simplex=True, return_status=True)
    assert np.array_equal(y_h, y_d.cpu().numpy(), equal_nan=True)
    assert np.array_equal(s_h, s_d.cpu().numpy()) and np.array_equal(t_h, t_d.cpu().numpy())
    assert np.array_equal(m.interp(q), y_h, equal_nan=True)          # values only (no simplex / status staging)
    if dim == 2:
        assert np.array_equal(m.gradient(q), m.gradient(xq).cpu().numpy(), equal_nan=True)
    # statistics calls stay on one worker and still count every query
    # statistics cover the whole call, however many chunks it is cut into
    m.collect_stats(True)
    k = (1 << 19) + 7
    _, t_k = m.interp(q[:k], return_status=True)
    st = m.stats()
    assert st["queries"] == k and st["outside"] == int(((t_k == 1) | (t_k == 2)).sum())
    m.collect_stats(False)


def test_closed_form_epilogue_within_tolerance(gpu):
    import delaunay_linterp_b200 as dl
    pts, f, sv, sn = helpers.random_mesh(2, 20000, 31)
    q = np.random.default_rng(32).random((200000, 2)) * 0.96 + 0.02
    y0 = dl.Mesh(pts, f, sv, sn, epilogue="reference").interp(q)
    y1 = dl.Mesh(pts, f, sv, sn, epilogue="closed_form").interp(q)
    rel = np.abs(y1 - y0) / np.maximum(np.abs(y0), 1e-300)
    assert rel.max() < 1e-10, rel.max()     # slivers: see SURVEY appendix A.2
    assert np.median(rel) < 1e-15


@pytest.mark.parametrize("opts", [{}, {"packed_faces": 2}, {"packed_faces": 3}])
def test_affine_reproduction_and_set_values(gpu, opts):
    import delaunay_linterp_b200 as dl
    pts, f, sv, sn = helpers.random_mesh(2, 3000, 41)
    m = dl.Mesh(pts, f, sv, sn, **opts)
    g = 2.0 + 3.0 * pts[:, 0] - 1.5 * pts[:, 1]
    m.set_values(g)
    q = np.random.default_rng(42).random((50000, 2)) * 0.9 + 0.05
    ys = m.interp(q)
    want = 2.0 + 3.0 * q[:, 0] - 1.5 * q[:, 1]
    assert np.max(np.abs(ys - want)) < 1e-12
    # vertex queries return the vertex value
    yv = m.interp(pts[:500])
    assert np.max(np.abs(yv - g[:500])) < 1e-13


def test_empty_and_tiny_inputs(gpu):
    import delaunay_linterp_b200 as dl
    pts = np.array([[0.0, 0.0], [1.0, 0.0], [0.0, 1.0]])
    m = dl.Mesh(pts, np.array([1.0, 2.0, 3.0]), np.array([[0, 1, 2]], dtype=np.int32))
    assert m.interp(np.zeros((0, 2))).shape == (0,)
    ys, st = m.interp(np.array([[0.25, 0.25], [2.0, 2.0], [0.0, 0.0]]), return_status=True)
    assert ys[0] == 1.0 + 0.25 * 1 + 0.25 * 2 and st[0] == 0
    assert st[1] == dl.STATUS_OUTSIDE_PROJECTED and abs(ys[1] - 2.5) < 1e-15   # projected on the hypotenuse midpoint
    assert ys[2] == 1.0
    # a mesh without simplices: every query reports NO_MESH and NaN
    m0 = dl.Mesh(pts[:2], np.array([1.0, 2.0]), np.zeros((0, 3), np.int32))
    y0, s0, st0 = m0.interp(np.array([[0.1, 0.1]]), return_simplex=True, return_status=True)
    assert np.isnan(y0[0]) and s0[0] == -1 and st0[0] == dl.STATUS_NO_MESH
    # degenerate simplex rejected
    with pytest.raises(dl.DliError):
        dl.Mesh(np.array([[0.0, 0.0], [1.0, 1.0], [2.0, 2.0]]), np.zeros(3), np.array([[0, 1, 2]], dtype=np.int32))


def test_device_predicate_adversarial(gpu):
    """orientation signs of near-degenerate triples vs exact rationals"""
    import delaunay_linterp_b200 as dl
    from oracle import pyoracle
    rng = np.random.default_rng(51)
    n = 3000
    a = rng.random((n, 2))
    b = rng.random((n, 2))
    t = rng.random(n)
    c = a + (b - a) * t[:, None]                    # nearly collinear (rounded)
    k = rng.integers(-2, 3, size=n)
    for i in range(n):                              # nudge by a few ulps
        for _ in range(abs(int(k[i]))):
            c[i, 1] = np.nextafter(c[i, 1], 2.0 if k[i] > 0 else -1.0)
    c[:50] = a[:50]                                 # exact duplicates -> 0
    pts = np.stack([a, b, c], axis=1)
    sign, used = dl.debug_orient(pts)
    want = np.array([pyoracle.orient(p.tolist()) for p in pts], dtype=np.int8)
    assert np.array_equal(sign, want)
    assert used.mean() > 0.1                        # these inputs do exercise the exact fallback
    sign_h = dl.debug_orient(pts, on_device=False)
    assert np.array_equal(sign_h, want)


def test_walk_record_guard_sign_changing_values(gpu):
    """auto epilogue: faces on which f changes sign / vanishes carry the slow flag and take the reference-order
    pass, so the 1e-12 relative bound holds even where y crosses zero; slivers likewise."""
    import delaunay_linterp_b200 as dl
    pts, _, sv, sn = helpers.random_mesh(2, 20000, 61)
    f = np.sin(7.0 * pts[:, 0]) * np.cos(5.0 * pts[:, 1])          # sign changes
    f[::97] = 0.0                                                   # exact zeros at vertices
    q = np.random.default_rng(62).random((300000, 2))
    m = dl.Mesh(pts, f, sv, sn)
    ys, s, st = m.interp(q, return_simplex=True, return_status=True)
    om = OracleMesh(pts, f, sv, sn)
    _check_parity(m, om, q, ys, s, st, bit_exact=False)
    # coordinates so small that the fp32 edge vectors of the walk records underflow: every face carries the slow
    # flag and the whole batch is served by the exact pass (power-of-two scaling keeps the mesh Delaunay)
    sc = 2.0 ** -140
    m2 = dl.Mesh(pts * sc, np.abs(f) + 1.0, sv, sn)
    q2 = q[:50000] * sc
    ys2, s2, st2 = m2.interp(q2, return_simplex=True, return_status=True)
    om2 = OracleMesh(pts * sc, np.abs(f) + 1.0, sv, sn)
    _check_parity(m2, om2, q2, ys2, s2, st2, bit_exact=True)


def test_large_mesh_delta_encoded_walk_records(gpu):
    """> 2^21 faces: walk records with delta-encoded neighbours over Morton-ordered faces (far neighbours read the
    canonical record); parity on a sample vs the oracle, and the full batch vs the reference-order engine."""
    import delaunay_linterp_b200 as dl
    rng = np.random.default_rng(71)
    pts = rng.random((1_150_000, 2))
    t = dl.DelaunayInterp2()
    t.insert_many(pts, helpers.cobb_douglas(pts), spatial_sort=True, use_gpu=False)
    c, f, sv, sn = t.export()
    assert len(sv) >= (1 << 21)
    m = dl.Mesh(c, f, sv, sn)
    q = rng.random((2_000_000, 2)) * 1.02 - 0.01
    ys, s, st = m.interp(q, return_simplex=True, return_status=True)
    mr = dl.Mesh(c, f, sv, sn, epilogue="reference")
    yr, sr, str_ = mr.interp(q, return_simplex=True, return_status=True)
    assert np.array_equal(st, str_) and np.array_equal(s, sr)
    assert np.max(np.abs(ys - yr) / np.maximum(np.abs(yr), 1e-300)) <= REL_TOL
    om = OracleMesh(c, f, sv, sn)
    k = 100_000
    _check_parity(m, om, q[:k], ys[:k], s[:k], st[:k], bit_exact=False)


def test_cell_indexed_records_on_a_graded_mesh(gpu):
    """cell-indexed walk records (packed_faces = 3) on an adaptively refined mesh: faces of very different sizes, so
    big faces own thousands of cells, tiny faces none (extra slots), and neighbour copies may be out of delta range
    (those queries go to the exact pass).  Same answers as the default layout, parity with the oracle."""
    import delaunay_linterp_b200 as dl
    fn = lambda x: x[0] ** 0.3 * x[1] ** 0.6
    t = dl.DelaunayInterp2(fn)
    for p in ([0.01, 0.01], [0.01, 5.0], [5.0, 0.01], [5.0, 5.0]):
        t.insert(p, fn(p))
    for _ in range(3000):
        t.insert_largest_error_point()
    c, f, sv, sn = t.export()
    q = np.random.default_rng(81).random((400000, 2)) * 5.2 - 0.1
    m0 = dl.Mesh(c, f, sv, sn)
    for cps in (2.0, 16.0):
        m = dl.Mesh(c, f, sv, sn, packed_faces=3, hint_cells_per_simplex=cps)
        m.collect_stats(True)
        ys, s, st = m.interp(q, return_simplex=True, return_status=True)
        stats = m.stats()
        y0, s0, st0 = m0.interp(q, return_simplex=True, return_status=True)
        assert np.array_equal(st, st0) and np.array_equal(s, s0) and np.array_equal(ys, y0)
        assert stats["straggler_queries"] < 0.2 * len(q)
        assert m.device_bytes > m0.device_bytes
    om = OracleMesh(c, f, sv, sn)
    _check_parity(m, om, q, ys, s, st, bit_exact=False)


def test_non_finite_queries_are_reported_not_walked(gpu):
    """NaN / Inf coordinates: status STEP_LIMIT ("no simplex found"), NaN value, simplex -1 -- immediately, not after
    max_walk_steps; the neighbouring queries of the batch are unaffected."""
    import torch
    import delaunay_linterp_b200 as dl
    pts, f, sv, sn = helpers.random_mesh(2, 20000, 91)
    q = np.random.default_rng(92).random((100000, 2)) * 0.9 + 0.05
    bad = np.arange(0, len(q), 7)
    qb = q.copy()
    qb[bad[0::3], 0] = np.nan
    qb[bad[1::3], 1] = np.inf
    qb[bad[2::3]] = -np.inf
    for opts in ({}, {"epilogue": "reference"}, {"packed_faces": 3}):
        m = dl.Mesh(pts, f, sv, sn, **opts)
        y0 = m.interp(q)
        ys, s, st = m.interp(qb, return_simplex=True, return_status=True)
        good = np.ones(len(q), bool); good[bad] = False
        assert np.array_equal(ys[good], y0[good]) and (st[good] == 0).all()
        assert np.isnan(ys[bad]).all() and (s[bad] == -1).all() and (st[bad] == dl.STATUS_STEP_LIMIT).all()
        g, gst = m.gradient(qb, return_status=True)
        assert np.isnan(g[bad]).all() and (gst[bad] == dl.STATUS_STEP_LIMIT).all() and np.isfinite(g[good]).all()
    p3, f3, sv3, sn3 = helpers.random_mesh(3, 3000, 93)
    q3 = np.random.default_rng(94).random((20000, 3)) * 0.8 + 0.1
    q3[::5, 1] = np.nan
    q3[1::5, 2] = np.inf
    y3, s3, st3 = dl.Mesh(p3, f3, sv3, sn3).interp(q3, return_simplex=True, return_status=True)
    assert (st3[::5] == dl.STATUS_STEP_LIMIT).all() and (st3[1::5] == dl.STATUS_STEP_LIMIT).all() and np.isnan(y3[::5]).all()
    assert (st3[2::5] == 0).all()
    # misaligned device pointer: rejected, not a misaligned-address fault
    m = dl.Mesh(pts, f, sv, sn)
    x = torch.rand(1001, 2, dtype=torch.float64, device="cuda")
    flat = x.view(-1)[1:2001]
    with pytest.raises(dl.DliError):
        m.interp_device_ptr(1000, flat.data_ptr(), torch.empty(1000, dtype=torch.float64, device="cuda").data_ptr())
