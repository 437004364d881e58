"""dli_interp_noalloc: the interp entry point for callers that cannot allocate device memory.

Who needs it: a program that handed the whole GPU to its own memory pool before it loaded this library (the driver
has nothing left for the queue dli_interp allocates per call), a loop that cannot afford the driver lock or the
device-wide synchronisation an allocation may bring, a driver set-up without stream-ordered pools.  The entry point
keeps its straggler queue in an array that is part of libdli.so's device image instead.

What is checked: through the C ABI (ctypes, raw device pointers), twice in a row on the same stream, the values,
simplex ids and status codes equal dli_interp's on the same queries -- the second call is the one that finds the
array dirty from the first.  2D (README-shaped lattice queries that fill the queue, and random queries with ~17 %
outside the hull) and 4D.  Also: a batch above DLI_NOALLOC_MAX_QUERIES and host pointers are refused with a message.

The file name sorts last on purpose: pytest -x stops at the first failure and this test must not hide the others.
torch only provides the device buffers here; they are allocated before the calls under test.
"""
import ctypes as C

import numpy as np
import pytest

import helpers

pytestmark = pytest.mark.gpu

REL_TOL = 1e-12  # the tolerance of the parity tests (tests/test_gpu_parity_2d.py)


def _device_buffers(q):
    import torch
    n = q.shape[0]
    xs = torch.from_numpy(q).cuda()
    mk = lambda dt: [torch.empty(n, dtype=dt, device="cuda") for _ in range(3)]
    return xs, mk(torch.float64), mk(torch.int32), mk(torch.uint8)


def _compare(mesh, q):
    import torch
    import delaunay_linterp_b200 as dl
    from delaunay_linterp_b200._capi import check
    L = dl.lib()
    n = q.shape[0]
    xs, ys, ss, ts = _device_buffers(q)
    for y in ys:
        y.fill_(-12345.0)
    st = torch.cuda.current_stream().cuda_stream
    # [0] the existing entry point, [1] and [2] two no-allocation calls in a row
    check(L.dli_interp(mesh._h, n, xs.data_ptr(), ys[0].data_ptr(), ss[0].data_ptr(), ts[0].data_ptr(), st))
    for k in (1, 2):
        check(L.dli_interp_noalloc(mesh._h, n, xs.data_ptr(), ys[k].data_ptr(), ss[k].data_ptr(), ts[k].data_ptr(), st))
    torch.cuda.synchronize()
    y0, s0, t0 = ys[0].cpu().numpy(), ss[0].cpu().numpy(), ts[0].cpu().numpy()
    for k in (1, 2):
        yk, sk, tk = ys[k].cpu().numpy(), ss[k].cpu().numpy(), ts[k].cpu().numpy()
        assert np.array_equal(tk, t0), "call %d: status codes differ" % k
        assert np.array_equal(sk, s0), "call %d: simplex ids differ" % k
        assert np.array_equal(np.isnan(yk), np.isnan(y0))
        fin = np.isfinite(y0)
        rel = np.abs(yk[fin] - y0[fin]) / np.maximum(np.abs(y0[fin]), 1e-300)
        assert rel.size == 0 or rel.max() <= REL_TOL, "call %d: max rel diff %g" % (k, rel.max())
    return y0, t0


def test_noalloc_matches_dli_interp_2d(gpu):
    import delaunay_linterp_b200 as dl
    # README-shaped: regular grid, lattice queries -- a third or more of them go through the queue
    c, f, t = helpers.regular_grid_mesh_2d(64)
    g = np.linspace(0.0, 5.0, 190)
    q = np.stack(np.meshgrid(g, g, indexing="ij"), -1).reshape(-1, 2)
    m = dl.Mesh(c, f, t)
    m.collect_stats(True)
    y, st = _compare(m, q)
    assert m.stats()["straggler_queries"] > 0, "the batch never used the queue: the test would prove nothing"
    assert np.isfinite(y).all() and (st == 0).all()
    # random mesh, ~17 % of the queries outside the hull (projected by the exact pass)
    pts, f, sv, sn = helpers.random_mesh(2, 20000, 5)
    q = np.random.default_rng(6).random((300000, 2)) * 1.1 - 0.05
    m2 = dl.Mesh(pts, f, sv, sn)
    y, st = _compare(m2, q)
    assert (st == 1).sum() > 1000
    # and the first mesh again: the array now holds the other mesh's entries
    _compare(m, np.ascontiguousarray(q[:50000] * 4.0 + 0.3))


def test_noalloc_matches_dli_interp_4d(gpu):
    import delaunay_linterp_b200 as dl
    pts, f, sv, sn = helpers.random_mesh(4, 400, 9)
    q = np.random.default_rng(10).random((60000, 4))
    m = dl.Mesh(pts, f, sv, sn)
    y, st = _compare(m, q)
    assert (st == 2).any() and (st == 0).any()     # outside the hull -> NaN in dD


def test_noalloc_refuses_what_it_cannot_hold(gpu):
    import torch
    import delaunay_linterp_b200 as dl
    from delaunay_linterp_b200 import _capi
    L = dl.lib()
    c, f, t = helpers.grid_mesh()
    m = dl.Mesh(c, f, t)
    n = _capi.DLI_NOALLOC_MAX_QUERIES + 1
    xs = torch.zeros((n, 2), dtype=torch.float64, device="cuda")
    ys = torch.full((n,), 7.0, dtype=torch.float64, device="cuda")
    rc = L.dli_interp_noalloc(m._h, n, xs.data_ptr(), ys.data_ptr(), None, None, None)
    assert rc == -1 and b"DLI_NOALLOC_MAX_QUERIES" in L.dli_last_error()
    torch.cuda.synchronize()
    assert float(ys.min()) == 7.0 and float(ys.max()) == 7.0      # nothing was launched
    # exactly the maximum is accepted
    rc = L.dli_interp_noalloc(m._h, n - 1, xs.data_ptr(), ys.data_ptr(), None, None, None)
    assert rc == 0, L.dli_last_error()
    torch.cuda.synchronize()
    # host pointers are refused
    hq = np.zeros((8, 2))
    hy = np.zeros(8)
    rc = L.dli_interp_noalloc(m._h, 8, hq.ctypes.data, hy.ctypes.data, None, None, None)
    assert rc == -1 and b"device pointers only" in L.dli_last_error()
    # per-launch timing creates events: refused while it is on, fine again once off
    m.collect_stats(True, timing=True)
    rc = L.dli_interp_noalloc(m._h, 8, xs.data_ptr(), ys.data_ptr(), None, None, None)
    assert rc == -5
    m.collect_stats(False)
    assert L.dli_interp_noalloc(m._h, 8, xs.data_ptr(), ys.data_ptr(), None, None, None) == 0
    torch.cuda.synchronize()
