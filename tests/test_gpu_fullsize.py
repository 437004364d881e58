"""GPU, BASELINE.json's full sizes: size-independent properties of the batched interp path.

The oracle cannot finish 100M queries in seconds, so at full size the CUDA path is checked through properties the
domain offers (affine reproduction, batch-partition and permutation invariance, agreement of the two device
engines -- walk records vs reference operation order -- on every query) plus oracle parity on a 1M-query subsample.
"""
import numpy as np
import pytest

import helpers
from oracle.oracle import OracleMesh

pytestmark = pytest.mark.gpu

REL_TOL = 1e-12


def _native_mesh(nv, seed):
    import delaunay_linterp_b200 as dl
    pts = np.random.default_rng(seed).random((nv, 2))
    t = dl.DelaunayInterp2()
    t.insert_many(pts, helpers.cobb_douglas(pts), spatial_sort=True, use_gpu=False)
    return t.export()


def test_cfg2_full_size_properties(gpu):
    """configs[1]: 1M uniform-random vertices, 100M uniform queries in the hull, fp64."""
    import torch
    import delaunay_linterp_b200 as dl
    c, f, sv, sn = _native_mesh(1_000_000, 1001)
    nq = 100_000_000
    g = torch.Generator(device="cuda")
    g.manual_seed(2001)
    xq = torch.rand(nq, 2, dtype=torch.float64, device="cuda", generator=g) * 0.96 + 0.02
    m = dl.Mesh(c, f, sv, sn)
    ys, s, st = m.interp(xq, return_simplex=True, return_status=True)
    torch.cuda.synchronize()
    assert int(st.max().item()) == 0                       # every query inside the hull, none hit the step limit
    assert bool(torch.isfinite(ys).all())
    # (1) the two device engines agree on every one of the 100M queries: same face, values within 1e-12
    mr = dl.Mesh(c, f, sv, sn, epilogue="reference")
    yr, sr = mr.interp(xq, return_simplex=True)
    torch.cuda.synchronize()
    assert bool(torch.equal(s, sr))
    rel = ((ys - yr).abs() / yr.abs().clamp_min(1e-300)).max().item()
    assert rel <= REL_TOL, rel
    # (2) batch-partition invariance: ragged shards give bit-identical results
    cuts = [0, 1, 31, 33, 1_000_003, 37_000_001, 99_999_999, nq]
    for a, b in zip(cuts[:-1], cuts[1:]):
        assert bool(torch.equal(m.interp(xq[a:b]), ys[a:b]))
    # (3) permutation equivariance on a 20M slice (the engine is order-independent)
    k = 20_000_000
    perm = torch.randperm(k, device="cuda", generator=g)
    assert bool(torch.equal(m.interp(xq[:k][perm].contiguous()), ys[:k][perm]))
    # (4) interpolant bounds: every value lies between the smallest and largest f of its face
    svt = torch.from_numpy(sv.astype(np.int64)).cuda()
    ft = torch.from_numpy(f).cuda()
    fv = ft[svt[s.long()[:k]]]
    lo, hi = fv.min(dim=1).values, fv.max(dim=1).values
    assert bool(((ys[:k] >= lo * (1 - 1e-12)) & (ys[:k] <= hi * (1 + 1e-12))).all())
    del fv, lo, hi, perm, yr, sr
    # (5) affine reproduction over all 100M queries (set_values rebuilds the walk records)
    m.set_values(2.0 + 3.0 * c[:, 0] - 1.5 * c[:, 1])
    ya = m.interp(xq)
    want = 2.0 + 3.0 * xq[:, 0] - 1.5 * xq[:, 1]
    assert ((ya - want).abs() / want.abs()).max().item() <= REL_TOL
    del ya, want
    # (6) oracle parity on a 1M-query subsample (values, statuses, located faces)
    m.set_values(f)
    n1 = 1_000_000
    q1 = xq[:n1].cpu().numpy()
    yo, so, sto = OracleMesh(c, f, sv, sn).interp(q1, mode=1, nthreads=8)
    assert np.array_equal(sto, st[:n1].cpu().numpy())
    y1 = ys[:n1].cpu().numpy()
    assert np.max(np.abs(y1 - yo) / np.maximum(np.abs(yo), 1e-300)) <= REL_TOL
    # identical face wherever the oracle's face strictly contains the query (always, for random queries)
    assert np.mean(s[:n1].cpu().numpy() == so) > 0.999999


def test_cfg4_3d_full_size_properties(gpu):
    """configs[3], 3D: 200k random vertices, 20M queries; outside the hull -> NaN, simplex -1."""
    import torch
    import delaunay_linterp_b200 as dl
    pts, f, sv, sn = helpers.random_mesh(3, 200_000, 1003)
    m = dl.Mesh(pts, f, sv, sn)
    nq = 20_000_000
    g = torch.Generator(device="cuda")
    g.manual_seed(2004)
    xq = torch.rand(nq, 3, dtype=torch.float64, device="cuda", generator=g) * 0.8 + 0.1
    ys, s, st = m.interp(xq, return_simplex=True, return_status=True)
    torch.cuda.synchronize()
    assert int(st.max().item()) == 0 and bool(torch.isfinite(ys).all())
    for a, b in ((0, 7), (7, 5_000_001), (5_000_001, nq)):
        assert bool(torch.equal(m.interp(xq[a:b]), ys[a:b]))
    # affine reproduction
    m.set_values(1.0 + pts @ np.array([0.5, -2.0, 3.0]))
    ya = m.interp(xq)
    want = 1.0 + xq @ torch.tensor([0.5, -2.0, 3.0], dtype=torch.float64, device="cuda")
    assert ((ya - want).abs() / want.abs().clamp_min(1e-3)).max().item() <= 1e-12
    # outside queries
    out = torch.tensor([[2.0, 0.5, 0.5], [-1.0, -1.0, -1.0]], dtype=torch.float64, device="cuda")
    yo, so, sto = m.interp(out, return_simplex=True, return_status=True)
    assert bool(torch.isnan(yo).all()) and so.tolist() == [-1, -1] and sto.tolist() == [dl.STATUS_OUTSIDE_NAN] * 2
    # oracle parity on a 200k subsample
    m.set_values(f)
    n1 = 200_000
    yo, so, sto = OracleMesh(pts, f, sv, sn).interp(xq[:n1].cpu().numpy(), mode=1, nthreads=8)
    y1 = ys[:n1].cpu().numpy()
    assert np.max(np.abs(y1 - yo) / np.maximum(np.abs(yo), 1e-300)) <= REL_TOL


def test_cfg3_adaptive_full_size_properties(gpu):
    """configs[2]: 100k points placed by insert_largest_error_point, 50M unsorted queries, plus one batched
    per-simplex error evaluation over all faces + hull edges."""
    import torch
    import delaunay_linterp_b200 as dl
    fn = lambda x: x[0] ** 0.3 * x[1] ** 0.6
    t = dl.DelaunayInterp2(fn)
    for p in ([0.01, 0.01], [0.01, 5.0], [5.0, 0.01], [5.0, 5.0]):
        t.insert(p, fn(p))
    for _ in range(99_996):
        t.insert_largest_error_point()
    c, f, sv, sn = t.export()
    assert len(c) == 100_000
    m = dl.Mesh(c, f, sv, sn)
    nq = 50_000_000
    g = torch.Generator(device="cuda")
    g.manual_seed(2003)
    xq = torch.rand(nq, 2, dtype=torch.float64, device="cuda", generator=g) * 4.99 + 0.01
    ys, s, st = m.interp(xq, return_simplex=True, return_status=True)
    torch.cuda.synchronize()
    assert int(st.max().item()) == 0 and bool(torch.isfinite(ys).all())
    # the two device engines agree on all 50M queries (strongly graded mesh: tiny faces near the axes)
    yr, sr = dl.Mesh(c, f, sv, sn, epilogue="reference").interp(xq, return_simplex=True)
    assert bool(torch.equal(s, sr))
    assert ((ys - yr).abs() / yr.abs().clamp_min(1e-300)).max().item() <= REL_TOL
    for a, b in ((0, 1), (1, 12_345_679), (12_345_679, nq)):
        assert bool(torch.equal(m.interp(xq[a:b]), ys[a:b]))
    # the adaptive mesh interpolates f much better than a regular grid of the same size would: sanity on the error
    err = (ys - xq[:, 0] ** 0.3 * xq[:, 1] ** 0.6).abs().max().item()
    assert err < 0.05
    # oracle parity on a 500k subsample
    n1 = 500_000
    yo, so, sto = OracleMesh(c, f, sv, sn).interp(xq[:n1].cpu().numpy(), mode=1, nthreads=8)
    assert np.max(np.abs(ys[:n1].cpu().numpy() - yo) / np.maximum(np.abs(yo), 1e-300)) <= REL_TOL
    # batched error evaluation of every face + hull edge == the oracle's SVD form, and == the host queue's entries
    cent, isx, iv = m.simplex_error_items()
    errs = m.simplex_error(helpers.cobb_douglas(cent))
    om = OracleMesh(c, f, sv, sn)
    o_simplex, o_nv, o_verts, o_cent = om.error_items()
    assert len(errs) == len(o_nv) and np.array_equal(cent, o_cent)
    want = om.error_eval(o_nv, o_verts, o_cent, helpers.cobb_douglas(o_cent))
    assert np.allclose(errs, want, rtol=1e-7, atol=1e-22)
    assert np.all(np.diff([e[0] for e in t.get_error_queue()[:1000]]) <= 0)


def test_cfg4_4d_full_size_properties(gpu):
    """configs[3], 4D: 200k random vertices (6.2M simplices), 20M queries."""
    import torch
    import delaunay_linterp_b200 as dl
    pts, f, sv, sn = helpers.random_mesh(4, 200_000, 1004)
    m = dl.Mesh(pts, f, sv, sn)
    nq = 20_000_000
    g = torch.Generator(device="cuda")
    g.manual_seed(2005)
    xq = torch.rand(nq, 4, dtype=torch.float64, device="cuda", generator=g) * 0.8 + 0.1
    ys, s, st = m.interp(xq, return_simplex=True, return_status=True)
    torch.cuda.synchronize()
    assert int(st.max().item()) == 0 and bool(torch.isfinite(ys).all())
    for a, b in ((0, 33), (33, 9_000_001), (9_000_001, nq)):
        assert bool(torch.equal(m.interp(xq[a:b]), ys[a:b]))
    yr, sr = dl.Mesh(pts, f, sv, sn, epilogue="reference").interp(xq[:5_000_000], return_simplex=True)
    assert bool(torch.equal(s[:5_000_000], sr))
    assert ((ys[:5_000_000] - yr).abs() / yr.abs().clamp_min(1e-300)).max().item() <= REL_TOL
    m.set_values(1.0 + pts @ np.array([0.5, -2.0, 3.0, 0.25]))
    ya = m.interp(xq)
    want = 1.0 + xq @ torch.tensor([0.5, -2.0, 3.0, 0.25], dtype=torch.float64, device="cuda")
    assert ((ya - want).abs() / want.abs().clamp_min(1e-3)).max().item() <= 1e-12
    n1 = 100_000
    yo, so, sto = OracleMesh(pts, f, sv, sn).interp(xq[:n1].cpu().numpy(), mode=1, nthreads=8)
    assert np.max(np.abs(ys[:n1].cpu().numpy() - yo) / np.maximum(np.abs(yo), 1e-300)) <= REL_TOL


def test_cfg5_one_shard_full_size_properties(gpu):
    """configs[4], one GPU's share: 16M-vertex mesh (32M faces, cell-indexed walk records), 125M queries = 1B / 8."""
    import torch
    import delaunay_linterp_b200 as dl
    c, f, sv, sn = _native_mesh(16_000_000, 1005)
    m = dl.Mesh(c, f, sv, sn)
    nq = 125_000_000
    g = torch.Generator(device="cuda")
    g.manual_seed(2006)
    xq = torch.rand(nq, 2, dtype=torch.float64, device="cuda", generator=g) * 0.96 + 0.02
    ys, s, st = m.interp(xq, return_simplex=True, return_status=True)
    torch.cuda.synchronize()
    assert int(st.max().item()) == 0 and bool(torch.isfinite(ys).all())
    # the Morton-ordered layout (delta-encoded neighbours) gives the same faces and bit-identical values (same hint
    # density as the cell-indexed default, so that a query takes the same pass -- fast or exact -- in both)
    m2 = dl.Mesh(c, f, sv, sn, packed_faces=2, hint_cells_per_simplex=4.0)
    k = 40_000_000
    y2, s2 = m2.interp(xq[:k], return_simplex=True)
    assert bool(torch.equal(s[:k], s2)) and bool(torch.equal(ys[:k], y2))
    del m2, y2, s2
    for a, b in ((0, 5), (5, 60_000_001), (60_000_001, nq)):
        assert bool(torch.equal(m.interp(xq[a:b]), ys[a:b]))
    m.set_values(2.0 + 3.0 * c[:, 0] - 1.5 * c[:, 1])
    ya = m.interp(xq)
    want = 2.0 + 3.0 * xq[:, 0] - 1.5 * xq[:, 1]
    assert ((ya - want).abs() / want.abs()).max().item() <= REL_TOL
    del ya, want
    n1 = 300_000
    yo, so, sto = OracleMesh(c, f, sv, sn).interp(xq[:n1].cpu().numpy(), mode=1, nthreads=8)
    assert np.max(np.abs(ys[:n1].cpu().numpy() - yo) / np.maximum(np.abs(yo), 1e-300)) <= REL_TOL
    assert np.mean(s[:n1].cpu().numpy() == so) > 0.99999


@pytest.mark.parametrize("dim", [2, 3, 4])
def test_degenerate_lattice_batches(gpu, dim):
    """The README example scaled up: lattice queries on a regular-grid mesh, every query exactly on a vertex, edge or
    facet (the filter can certify nothing: the whole batch takes the exact pass) -- statuses, containment and values
    against the oracle on a subsample, finiteness and engine agreement on the whole batch."""
    import torch
    import delaunay_linterp_b200 as dl
    rng = np.random.default_rng(70 + dim)
    if dim == 2:
        n = 1000
        c, f, sv = helpers.regular_grid_mesh_2d(n)
        nq = 8_000_000
        q = rng.integers(0, 4 * (n - 1) + 1, (nq, 2)) * (5.0 / (4 * (n - 1)))
    else:
        n = 40 if dim == 3 else 12
        c, f, sv = helpers.kuhn_grid_mesh(dim, n)
        nq = 2_000_000 if dim == 3 else 1_000_000
        q = rng.integers(0, 4 * (n - 1) + 1, (nq, dim)) / (4.0 * (n - 1))
    m = dl.Mesh(c, f, sv, None)
    xq = torch.tensor(q, device="cuda")
    m.collect_stats(True)
    ys, s, st = m.interp(xq, return_simplex=True, return_status=True)
    torch.cuda.synchronize()
    stats = m.stats()
    m.collect_stats(False)
    assert int(st.max().item()) == 0 and bool(torch.isfinite(ys).all())
    # 62.5 % of the 2D lattice queries sit exactly on a vertex / edge.  The 2D fast pass certifies the ones on
    # axis-aligned edges through vertex 0 of a walk record itself (closed-face test o >= e with both products exact
    # zeros); the ones on cell diagonals (18.75 %) need the exact predicate.  dD: every on-facet query takes the exact pass.
    need = 0.15 if dim == 2 else 0.5
    assert stats["straggler_queries"] > need * nq and stats["exact_fallbacks"] > 0.6 * need * nq    # really the exact pass
    m2 = dl.Mesh(c, f, sv, None, epilogue="reference")                                         # reference-order engine
    y2 = m2.interp(xq)
    assert float(((ys - y2).abs() / y2.abs().clamp_min(1e-300)).max().item()) <= REL_TOL
    sub = rng.choice(nq, 100_000, replace=False)
    om = OracleMesh(c, f, sv, None)
    yo, so, sto = om.interp(q[sub], mode=1, nthreads=8)
    assert np.array_equal(st[sub].cpu().numpy(), sto)
    contains, _ = om.classify(q[sub], s[sub].cpu().numpy())
    assert contains.all()                                  # a closed simplex containing the query (any incident one)
    yg = ys[sub].cpu().numpy()
    assert np.max(np.abs(yg - yo) / np.maximum(np.abs(yo), 1e-300)) <= REL_TOL
