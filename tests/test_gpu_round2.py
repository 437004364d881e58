"""GPU tests added in round 2: bounded straggler queue (overflow marks), outside_nan, empty-mesh gradient, a third
opinion (scipy LinearNDInterpolator) for 2D / 3D / 4D, concurrent calls on one mesh, k-at-a-time refinement followed by
the device mesh, the native dD class on the GPU in every dimension."""
import math
import threading

import numpy as np
import pytest

import helpers
from oracle.oracle import OracleMesh

pytestmark = pytest.mark.gpu
REL_TOL = 1e-12


def _rel(a, b):
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-300)))


@pytest.mark.parametrize("dim,nv", [(2, 20000), (3, 5000), (4, 1200)])
@pytest.mark.parametrize("via", ["host", "device"])
def test_straggler_queue_overflow_marks(gpu, dim, nv, via):
    """more than 1/16 of the batch needs the exact pass (queries far outside the hull, queries on vertices): the queue
    overflows, the rest is marked in the result slots and found again by the exact pass -- results as without overflow"""
    import torch
    import delaunay_linterp_b200 as dl
    pts, f, sv, sn = helpers.random_mesh(dim, nv, 70 + dim)
    rng = np.random.default_rng(71)
    nq = 1 << 20
    q = rng.random((nq, dim)) * 2.0 - 0.5                      # ~ 1 - 2^-dim of the queries are outside the hull
    q[: nq // 8] = pts[rng.integers(0, nv, nq // 8)]           # exactly on vertices: the filter cannot certify them
    m = dl.Mesh(pts, f, sv, sn)
    m.collect_stats(True)
    if via == "device":
        ys, s, st = m.interp(torch.from_numpy(q).cuda(), return_simplex=True, return_status=True)
        ys, s, st = ys.cpu().numpy(), s.cpu().numpy(), st.cpu().numpy()
    else:
        ys, s, st = m.interp(q, return_simplex=True, return_status=True)
    stats = m.stats()
    assert stats["straggler_queries"] > nq // 4                # far beyond the queue capacity (nq / 16)
    yo, so, sto = OracleMesh(pts, f, sv, sn).interp(helpers_sorted(q), mode=1, nthreads=8)
    order = helpers_order(q)
    assert np.array_equal(st[order], sto)
    ok = sto != 2
    assert _rel(ys[order][ok], yo[ok]) <= REL_TOL and np.array_equal(np.isnan(ys[order]), np.isnan(yo))
    assert (s[order][sto == 2] == -1).all()
    # dli_locate (no value array: the mark goes into the simplex slot)
    s2, st2 = m.locate(q, return_status=True)
    assert np.array_equal(st2, st) and np.array_equal(s2 >= 0, s >= 0)
    if dim == 2:
        g, sg = m.gradient(q, return_status=True)
        assert np.isnan(g[sg == 2]).all() and np.isfinite(g[sg != 2]).all()


def helpers_order(q, bits=8):
    lo, hi = q.min(0), q.max(0)
    g = np.minimum(((q - lo) / (hi - lo) * (1 << bits)).astype(np.int64), (1 << bits) - 1)
    key = np.zeros(len(q), np.int64)
    for b in range(bits):
        for c in range(q.shape[1]):
            key |= ((g[:, c] >> b) & 1) << (b * q.shape[1] + c)
    return np.argsort(key, kind="stable")


def helpers_sorted(q):
    return q[helpers_order(q)]


def test_outside_nan_option_2d(gpu):
    """Delaunay_incremental_interp_d(2) answers NaN outside the hull (src/delaunay_d_interp.h:1005-1008); the 2D engine
    projects by default (src/delaunay_2_interp.h:530-545)"""
    import delaunay_linterp_b200 as dl
    pts, f, sv, sn = helpers.random_mesh(2, 5000, 81)
    q = np.random.default_rng(82).random((200000, 2)) * 1.4 - 0.2
    y0, s0, st0 = dl.Mesh(pts, f, sv, sn).interp(q, return_simplex=True, return_status=True)
    y1, s1, st1 = dl.Mesh(pts, f, sv, sn, outside_nan=1).interp(q, return_simplex=True, return_status=True)
    out = st0 == dl.STATUS_OUTSIDE_PROJECTED
    assert out.sum() > 50000 and (st1[out] == dl.STATUS_OUTSIDE_NAN).all() and np.isnan(y1[out]).all() and (s1[out] == -1).all()
    assert np.array_equal(y0[~out], y1[~out]) and np.array_equal(st0[~out], st1[~out]) and np.array_equal(s0[~out], s1[~out])
    # the dD class in two dimensions, batch large enough for the GPU
    t = dl.Delaunay_incremental_interp_d(2)
    t.insert_many(pts, f, use_gpu=False)
    yd, std = t.interp(q, return_status=True)
    assert np.array_equal(np.isnan(yd), out) and (std[out] == dl.STATUS_OUTSIDE_NAN).all()
    assert _rel(yd[~out], y0[~out]) <= REL_TOL


def test_small_batches_equal_large_batches_with_outside_nan(gpu):
    """the host path for small batches follows the handle's options: with outside_nan=1 a one-point call
    and a 100k-query call agree on status and value for the same queries (NaN outside the hull, not a projection)."""
    import delaunay_linterp_b200 as dl
    pts, f, sv, sn = helpers.random_mesh(2, 3000, 83)
    q = np.random.default_rng(84).random((100000, 2)) * 1.4 - 0.2
    t = dl.DelaunayInterp2(outside_nan=1)
    t.insert_many(pts, f, use_gpu=False)
    yb, stb = t.interp(q, return_status=True)                     # device
    assert (stb == dl.STATUS_OUTSIDE_NAN).sum() > 20000 and not (stb == dl.STATUS_OUTSIDE_PROJECTED).any()
    for i in range(0, 400):
        y1, st1 = t.interp(q[i:i + 1], return_status=True)        # host
        assert st1[0] == stb[i], (i, st1[0], stb[i])
        assert (np.isnan(y1[0]) and np.isnan(yb[i])) or abs(y1[0] - yb[i]) <= REL_TOL * max(abs(yb[i]), 1e-300)


@pytest.mark.parametrize("dim,nv", [(2, 4000), (3, 1500), (4, 500)])
def test_error_evaluator_on_device_pointers(gpu, dim, nv):
    """prepare -> f evaluated by torch on the device -> finish, nothing staged through the host: bit-identical to the
    host-pointer path (src/delaunay_2_interp.h:725-741, 683-699 are what the two kernels replace)"""
    import torch
    import delaunay_linterp_b200 as dl
    pts, f, sv, sn = helpers.random_mesh(dim, nv, 60 + dim)
    m = dl.Mesh(pts, f, sv, sn)
    cent_h, isx, iv = m.simplex_error_items()
    cent_d = m.simplex_error_centroids_device()
    torch.cuda.synchronize()
    assert cent_d.shape == cent_h.shape and np.array_equal(cent_d.cpu().numpy(), cent_h)
    alpha = torch.tensor([0.3, 0.6, 0.2, 0.1][:dim], dtype=torch.float64, device="cuda")
    phi_d = torch.prod(torch.pow(cent_d, alpha), dim=1)          # the caller's f, on the GPU
    err_d = m.simplex_error(phi_d)
    torch.cuda.synchronize()
    err_h = m.simplex_error(phi_d.cpu().numpy())                 # the same values through the host-pointer path
    assert np.array_equal(err_d.cpu().numpy(), err_h)
    assert err_h.shape == (len(isx),) and np.isfinite(err_h).all() and err_h.max() > 0
    # mixed host / device pointers are refused
    from delaunay_linterp_b200._capi import lib
    bad = np.empty(len(isx))
    assert lib().dli_simplex_error_finish(m._h, len(isx), phi_d.data_ptr(), bad.ctypes.data) == -1


def test_empty_mesh_gradient_is_all_nan(gpu):
    """the no-mesh fill must cover 2 n doubles for gradients"""
    import torch
    import delaunay_linterp_b200 as dl
    m = dl.Mesh(np.array([[0.0, 0.0], [1.0, 1.0]]), np.zeros(2), np.zeros((0, 3), np.int32))
    q = np.random.default_rng(1).random((5000, 2))
    g, st = m.gradient(q, return_status=True)
    assert np.isnan(g).all() and (st == dl.STATUS_NO_MESH).all()
    gd = torch.full((5000, 2), 7.0, dtype=torch.float64, device="cuda")
    dl.lib().dli_gradient(m._h, 5000, torch.from_numpy(q).cuda().data_ptr(), gd.data_ptr(), None, None, None)
    torch.cuda.synchronize()
    assert bool(torch.isnan(gd).all())
    a = dl.DelaunayInterp2()
    a.insert([0.0, 0.0], 1.0)
    assert np.isnan(a.gradient(q)).all()


@pytest.mark.parametrize("dim,nv", [(2, 20000), (3, 6000), (4, 1500)])
def test_third_opinion_scipy_linear_nd_interpolator(gpu, dim, nv):
    """an implementation that shares no code with this repo or its oracle (Qhull + scipy's own barycentric
    transform arithmetic).  Well-shaped simplices only: scipy's affine transforms lose digits on slivers."""
    from scipy.interpolate import LinearNDInterpolator
    from scipy.spatial import Delaunay
    import delaunay_linterp_b200 as dl
    rng = np.random.default_rng(90 + dim)
    pts = rng.random((nv, dim))
    f = 1.0 + helpers.cobb_douglas(pts)
    tri = Delaunay(pts)
    q = rng.random((200000, dim)) * 0.9 + 0.05
    want = LinearNDInterpolator(tri, f)(q)
    t = dl.Delaunay_incremental_interp_d(dim)             # native triangulation + GPU engine
    t.insert_many(pts, f, use_gpu=False)
    got = t.interp(q)
    assert np.array_equal(np.isnan(got), np.isnan(want))
    # condition of the containing simplex: skip the worst 0.1 % (slivers)
    s = tri.find_simplex(q)
    T = tri.transform[s, :dim]
    cond = np.linalg.norm(T.reshape(len(q), -1), axis=1)
    keep = (~np.isnan(want)) & (cond <= np.quantile(cond, 0.999))
    assert _rel(got[keep], want[keep]) <= 1e-11


def test_concurrent_calls_on_one_mesh(gpu):
    """include/dli.h: interp / locate calls on the same mesh may be issued concurrently from several host threads and
    streams.  4 threads x (device-pointer calls on own streams + host-pointer calls), results bit-identical."""
    import torch
    import delaunay_linterp_b200 as dl
    pts, f, sv, sn = helpers.random_mesh(2, 50000, 95)
    m = dl.Mesh(pts, f, sv, sn)
    rng = np.random.default_rng(96)
    qs = [rng.random((300000, 2)) * 1.1 - 0.05 for _ in range(4)]
    want = [m.interp(q) for q in qs]
    errors = []

    def work(i):
        try:
            torch.cuda.set_device(0)
            stream = torch.cuda.Stream()
            xq = torch.from_numpy(qs[i]).cuda()
            for rep in range(8):
                with torch.cuda.stream(stream):
                    yd = m.interp(xq, stream=stream.cuda_stream)
                stream.synchronize()
                if not np.array_equal(yd.cpu().numpy(), want[i]):
                    errors.append("device path, thread %d rep %d" % (i, rep))
                if not np.array_equal(m.interp(qs[i]), want[i]):
                    errors.append("host path, thread %d rep %d" % (i, rep))
        except Exception as e:  # noqa: BLE001
            errors.append(repr(e))

    th = [threading.Thread(target=work, args=(i,)) for i in range(4)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert not errors, errors[:3]


def test_refine_rounds_keep_the_device_mesh_in_step(gpu):
    """k-at-a-time refinement (dli_di2_refine) + incremental device update; after every round the
    device mesh answers like the oracle on the exported triangulation, and the patch path (not a re-flatten) ran"""
    import delaunay_linterp_b200 as dl
    fn = lambda x: x[0] ** 0.3 * x[1] ** 0.6   # noqa: E731
    fnb = lambda xs: helpers.cobb_douglas(xs)   # noqa: E731
    t = dl.DelaunayInterp2(fn)
    for p in ([0.01, 0.01], [0.01, 5.0], [5.0, 0.01], [5.0, 5.0]):
        t.insert(p, fn(p))
    for _ in range(500):
        t.insert_largest_error_point()
    q = np.random.default_rng(97).random((100000, 2)) * 5.2 - 0.1
    t.interp(q)
    full0, inc0 = t.sync_counts()
    total = 0
    for rnd in range(12):
        total += len(t.insert_largest_error_points(64, fn_batch=fnb))
        ys, st = t.interp(q, return_status=True)
        c, f, tri, nb = t.export()
        yo, so, sto = OracleMesh(c, f, tri, nb).interp(helpers_sorted(q), mode=1, nthreads=8)
        o = helpers_order(q)
        assert np.array_equal(st[o], sto) and _rel(ys[o], yo) <= REL_TOL
    full1, inc1 = t.sync_counts()
    assert total > 200 and inc1 - inc0 >= 10 and full1 - full0 <= 2
    # the GPU evaluator over the whole refined mesh == the queue the rounds left behind (nothing stale, nothing missing)
    m = dl.Mesh(c, f, tri, nb)
    cent, isx, iv = m.simplex_error_items()
    err = np.sort(m.simplex_error(fnb(cent)))
    have = np.sort(np.array([e[0] for e in t.get_error_queue()]))
    assert len(err) == len(have) and np.allclose(err, have, rtol=1e-9, atol=1e-18)


@pytest.mark.parametrize("dim", [2, 3, 4])
def test_native_dd_class_on_gpu(gpu, dim):
    """Delaunay_incremental_interp_d on the native triangulation: adaptive loop on the host, batched interp on the GPU,
    per-point interp on the host -- same values; GPU-rebuilt queue == per-insert host queue"""
    import delaunay_linterp_b200 as dl
    a = np.array([0.3, 0.6, 0.2, 0.1])[:dim]
    fn = lambda x: float(np.prod(np.power(x, a)))   # noqa: E731
    t = dl.Delaunay_incremental_interp_d(dim, fn)
    for c in range(1 << dim):
        p = [2.0 if (c >> k) & 1 else 0.01 for k in range(dim)]
        t.insert(p, fn(p))
    for _ in range(150):
        t.insert_largest_error_point()
    assert t.validate() == 0
    q = np.random.default_rng(98).random((50000, dim)) * 2.2 - 0.1
    ys, st = t.interp(q, return_status=True)
    pts, f, sv, sn = t.export()
    o = helpers_order(q)
    yo, so, sto = OracleMesh(pts, f, sv, sn).interp(q[o], mode=1, nthreads=8)
    assert np.array_equal(st[o], np.where(sto == 1, 2, sto)) or np.array_equal(st[o], sto)
    ok = sto == 0
    assert _rel(ys[o][ok], yo[ok]) <= REL_TOL and np.isnan(ys[o][~ok]).all()
    one = np.array([t.interp(p.tolist()) for p in q[:300]])
    assert np.array_equal(np.isnan(one), np.isnan(ys[:300])) and _rel(one[~np.isnan(one)], ys[:300][~np.isnan(one)]) <= REL_TOL
    before = np.sort([e[0] for e in t.get_error_queue()])
    t.rebuild_error_queue()
    after = np.sort([e[0] for e in t.get_error_queue()])
    assert len(before) == len(after) and np.allclose(before, after, rtol=1e-8, atol=1e-18)
    assert math.isnan(t.interp([9.0] * dim))


def test_4d_packed_records_equal_split_records(gpu):
    """4D walk records: the packed one-line layout (block-floating-point barycentric gradients, default) answers like the
    split layout (fp32 gradients + separate interpolant record; packed_faces=5) -- same statuses, same simplex for
    every query both fast passes certify, bit-identical values on the same simplex (same fp64 interpolant, same
    operation order) -- and like the oracle; the wider filter band of the packed mantissas only sends a few more
    queries to the exact pass."""
    import delaunay_linterp_b200 as dl
    pts, f, sv, sn = helpers.random_mesh(4, 3000, 81)
    rng = np.random.default_rng(82)
    q = rng.random((200000, 4)) * 1.04 - 0.02
    q[:200] = pts[:200]
    res = {}
    for name, opt in (("packed", {}), ("split", {"packed_faces": 5})):
        m = dl.Mesh(pts, f, sv, sn, **opt)
        m.collect_stats(True)
        ys, s, st = m.interp(q, return_simplex=True, return_status=True)
        res[name] = (ys, s, st, m.stats())
    (y1, s1, t1, st1), (y2, s2, t2, st2) = res["packed"], res["split"]
    assert np.array_equal(t1, t2)
    same = s1 == s2
    assert same.mean() > 0.999
    # the same fp64 interpolant in the same operation order -> the same bits, except for the queries only ONE layout's
    # filter could not certify (the exact pass evaluates those in the reference's operation order)
    bits = (y1[same] == y2[same]) | (np.isnan(y1[same]) & np.isnan(y2[same]))
    assert bits.mean() > 0.99
    fin = same & np.isfinite(y2)
    assert np.array_equal(np.isnan(y1), np.isnan(y2)) and _rel(y1[fin], y2[fin]) <= REL_TOL
    assert st1["straggler_queries"] < 0.02 * len(q) + st2["straggler_queries"]
    om = OracleMesh(pts, f, sv, sn)
    yo, so, sto = om.interp(q, mode=1, nthreads=8)
    assert np.array_equal(t1, sto)
    ok = sto == 0
    cg, _ = om.classify(q, s1)
    assert cg[ok].all()
    assert _rel(y1[ok], yo[ok]) <= REL_TOL
    # coordinates at another scale: row exponents far from 0
    for scale in (1e-30, 3e25):
        m = dl.Mesh(pts * scale, f, sv, sn)
        ys, st = m.interp(q * scale, return_status=True)
        agree = (st == sto) & ok      # (a scale that is not a power of two re-rounds the coordinates)
        assert np.mean(st == sto) > 0.9999 and _rel(ys[agree], yo[agree]) <= 1e-10


@pytest.mark.parametrize("dim,nv", [(3, 6000), (4, 2500)])
def test_start_copies_equal_hinted_start(gpu, dim, nv):
    """dD start copies (a copy of the hinted simplex's record in every hint cell: the first record of a walk costs one
    memory access instead of hint + record; default in 4D): the same walks, so the same simplices, statuses and bits as
    with the hint table -- except where the second step's facet test against the facet the walk came through is
    uncertain (a copy does not know its own id), which only moves that query to the exact pass."""
    import delaunay_linterp_b200 as dl
    pts, f, sv, sn = helpers.random_mesh(dim, nv, 90 + dim)
    rng = np.random.default_rng(91)
    q = rng.random((300000, dim)) * 1.04 - 0.02
    q[:300] = pts[:300]
    on = dl.Mesh(pts, f, sv, sn, start_records=1)
    off = dl.Mesh(pts, f, sv, sn, start_records=-1)
    assert on.device_bytes > off.device_bytes
    y1, s1, t1 = on.interp(q, return_simplex=True, return_status=True)
    y0, s0, t0 = off.interp(q, return_simplex=True, return_status=True)
    assert np.array_equal(t1, t0)
    same = s1 == s0
    assert same.mean() > 0.999
    bits = (y1[same] == y0[same]) | (np.isnan(y1[same]) & np.isnan(y0[same]))
    assert bits.mean() > 0.999
    om = OracleMesh(pts, f, sv, sn)
    yo, so, sto = om.interp(q, mode=1, nthreads=8)
    ok = sto == 0
    assert np.array_equal(t1, sto) and _rel(y1[ok], yo[ok]) <= REL_TOL
    cg, _ = om.classify(q, s1)
    assert cg[ok].all()
    # locate and new values go through the copies too
    s2, t2 = on.locate(q, return_status=True)
    assert np.array_equal(t2, t1) and np.array_equal(s2 >= 0, s1 >= 0)
    on.set_values(2.0 * f + 1.0)
    y3 = on.interp(q)
    assert _rel(y3[ok], 2.0 * yo[ok] + 1.0) <= 1e-11


@pytest.mark.parametrize("d", [3, 4])
def test_dd_refine_rounds_keep_the_device_mesh_in_step(gpu, d):
    """dD form of the test above: dli_dd_refine rounds + the dD incremental device update; after every round the device
    mesh answers like the oracle on the exported triangulation, and the GPU evaluator over the refined mesh equals the
    queue the rounds left behind"""
    import delaunay_linterp_b200 as dl
    al = np.array([0.3, 0.6, 0.2, 0.1])[:d]
    fn = lambda x: float(np.prod(np.asarray(x) ** al) + 1.0)   # noqa: E731
    fnb = lambda a: np.prod(a ** al, axis=1) + 1.0              # noqa: E731
    corners = np.array(np.meshgrid(*([[0.0, 4.0]] * d), indexing="ij")).reshape(d, -1).T
    pts = np.concatenate([corners, np.random.default_rng(130 + d).random((1500 if d == 3 else 500, d)) * 3.8 + 0.1])
    t = dl.Delaunay_incremental_interp_d(d, fn)
    t.insert_many(pts, fnb(pts), use_gpu=False)
    q = np.random.default_rng(131).random((60000, d)) * 4.2 - 0.1
    t.interp(q)
    full0, inc0 = t.sync_counts()
    total = 0
    for rnd in range(6):
        total += len(t.insert_largest_error_points(16, fn_batch=fnb, rel_threshold=0.0))
        ys, sx, st = t.interp(q, return_simplex=True, return_status=True)
        c, f, sv, sn = t.export()
        om = OracleMesh(c, f, sv, sn)
        yo, so, sto = om.interp(q, mode=1, nthreads=8)
        ok = sto == 0
        assert np.array_equal(st, sto) and _rel(ys[ok], yo[ok]) <= REL_TOL
        cg, _ = om.classify(q, sx)
        assert cg[ok].all()
    full1, inc1 = t.sync_counts()
    assert total >= 20 and inc1 - inc0 >= 3 and (full1 - full0) + (inc1 - inc0) == 6   # at least half of the six syncs are in-place patches (3D: 3, 4D: more; a dD insertion that leaves holes in the id space re-flattens)
    m = dl.Mesh(c, f, sv, sn)
    cent, isx, iv = m.simplex_error_items()
    err = np.sort(m.simplex_error(fnb(cent)))
    have = np.sort(np.array([e[0] for e in t.get_error_queue()]))
    assert len(err) == len(have) and np.allclose(err, have, rtol=1e-9, atol=1e-18)
