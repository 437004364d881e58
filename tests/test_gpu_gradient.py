"""GPU parity: batched gradient (dli_gradient through the C ABI) vs the CPU oracle.
Reference: Delaunay_incremental_interp_2::gradient, src/delaunay_2_interp.h:608-642."""
import numpy as np
import pytest

import helpers
from oracle.oracle import OracleMesh, OK, ON_EDGE, ON_VERTEX, OUTSIDE_NAN

pytestmark = pytest.mark.gpu

VARIANTS = {"auto": {}, "reference": {"epilogue": "reference"}, "delta": {"packed_faces": 2}, "noring": {"prefetch_ring": -1},
            "cells": {"packed_faces": 3}, "cells_noring": {"packed_faces": 3, "prefetch_ring": -1}}


def _check(om, q, gs, simplex, status):
    go, so, sto = om.gradient(q)
    assert np.array_equal(status, sto), "status codes differ"
    fin = np.isfinite(go).all(axis=1)
    assert np.array_equal(np.isfinite(gs).all(axis=1), fin)
    assert (simplex[sto == OUTSIDE_NAN] == -1).all()
    # located face: identical strictly inside a face, any incident face on an edge / vertex
    interior = sto == OK
    assert np.array_equal(simplex[interior], so[interior])
    contains, _ = om.classify(q, simplex)
    assert contains[fin].all()
    # same operations in the same (canonical) order on both sides: bit-identical
    assert np.array_equal(gs[fin], go[fin]), "max abs diff %g" % np.abs(gs[fin] - go[fin]).max()


@pytest.mark.parametrize("variant", list(VARIANTS))
def test_gradient_grid_vertices_edges_faces(gpu, variant):
    import delaunay_linterp_b200 as dl
    c, f, t = helpers.grid_mesh(0.01, 5.0, 10)
    g = helpers.linspace_ref(0.01, 5.0, 10)
    q = [[g[i], g[j]] for i in range(10) for j in range(10)]
    q += [[(g[i] + g[i + 1]) / 2, g[j]] for i in range(9) for j in range(10)]
    q += [[g[j], (g[i] + g[i + 1]) / 2] for i in range(9) for j in range(10)]
    q += [[-1.0, 2.0], [2.0, 6.0], [5.5, 5.5]]
    q = np.vstack([np.array(q), helpers.grid_queries(0.01)])
    m = dl.Mesh(c, f, t, **VARIANTS[variant])
    gs, s, st = m.gradient(q, return_simplex=True, return_status=True)
    assert (st[:100] == ON_VERTEX).all() and (st == ON_EDGE).sum() >= 100 and (st == OUTSIDE_NAN).sum() == 3
    _check(OracleMesh(c, f, t), q, gs, s, st)


@pytest.mark.parametrize("variant", list(VARIANTS))
def test_gradient_random_mesh(gpu, variant):
    import delaunay_linterp_b200 as dl
    pts, f, sv, sn = helpers.random_mesh(2, 50000, 21)
    rng = np.random.default_rng(22)
    q = np.vstack([rng.random((200000, 2)) * 1.1 - 0.05, pts[:2000]])   # inside, outside, and exact vertices
    m = dl.Mesh(pts, f, sv, sn, **VARIANTS[variant])
    gs, s, st = m.gradient(q, return_simplex=True, return_status=True)
    assert (st[-2000:] == ON_VERTEX).all() and (st == OUTSIDE_NAN).sum() > 100
    _check(OracleMesh(pts, f, sv, sn), q, gs, s, st)


def test_gradient_device_pointers_and_affine(gpu):
    import torch
    import delaunay_linterp_b200 as dl
    pts, _, sv, sn = helpers.random_mesh(2, 20000, 23)
    f = 2.0 + 3.0 * pts[:, 0] - 1.5 * pts[:, 1]
    m = dl.Mesh(pts, f, sv, sn)
    xq = torch.rand(500000, 2, dtype=torch.float64, device="cuda:0") * 0.9 + 0.05
    gs, st = m.gradient(xq, return_status=True)
    torch.cuda.synchronize()
    assert (st == 0).all()
    assert (gs - torch.tensor([3.0, -1.5], dtype=torch.float64, device="cuda:0")).abs().max().item() < 1e-8
    # host-pointer path == device-pointer path
    gh = m.gradient(xq.cpu().numpy())
    assert np.array_equal(gh, gs.cpu().numpy())


def test_gradient_rejects_dd_and_empty(gpu):
    import delaunay_linterp_b200 as dl
    pts, f, sv, sn = helpers.random_mesh(3, 200, 5)
    m3 = dl.Mesh(pts, f, sv, sn)
    with pytest.raises(ValueError):
        m3.gradient(np.zeros((4, 3)))
    with pytest.raises(dl.DliError):
        from delaunay_linterp_b200._capi import check
        out = np.empty((4, 3))
        check(dl.lib().dli_gradient(m3._h, 4, out.ctypes.data, out.ctypes.data, None, None, None))
    c = np.array([[0.0, 0.0], [1.0, 0.0]])
    m0 = dl.Mesh(c, np.zeros(2), np.zeros((0, 3), np.int32))
    gs, s, st = m0.gradient(np.array([[0.5, 0.0], [2.0, 2.0]]), return_simplex=True, return_status=True)
    assert np.isnan(gs).all() and (s == -1).all() and (st == 4).all()
