"""GPU: mesh snapshot (dli_mesh_save / dli_mesh_load) -- a loaded replica answers bit-identically."""
import numpy as np
import pytest

import helpers

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dim,opts", [(2, {}), (2, {"epilogue": "reference"}), (2, {"packed_faces": 2}), (2, {"packed_faces": 3}), (3, {}), (4, {}),
                                      (3, {"start_records": 1}), (4, {"start_records": -1}), (4, {"packed_faces": 5})])
def test_save_load_roundtrip(gpu, tmp_path, dim, opts):
    import delaunay_linterp_b200 as dl
    pts, f, sv, sn = helpers.random_mesh(dim, 3000 if dim < 4 else 600, 40 + dim)
    m = dl.Mesh(pts, f, sv, sn, **opts)
    path = tmp_path / "mesh.dli"
    m.save(path)
    r = dl.Mesh.load(path, **opts)
    assert r.dim == dim and r.n_simplices == m.n_simplices and r.n_vertices == m.n_vertices
    q = np.random.default_rng(7).random((50000, dim)) * 1.1 - 0.05
    y0, s0, t0 = m.interp(q, return_simplex=True, return_status=True)
    y1, s1, t1 = r.interp(q, return_simplex=True, return_status=True)
    assert np.array_equal(t0, t1) and np.array_equal(s0, s1)
    assert np.array_equal(y0[np.isfinite(y0)], y1[np.isfinite(y0)]) and np.array_equal(np.isnan(y0), np.isnan(y1))
    if dim == 2:
        g0, g1 = m.gradient(q), r.gradient(q)
        assert np.array_equal(np.isnan(g0), np.isnan(g1)) and np.array_equal(g0[np.isfinite(g0)], g1[np.isfinite(g1)])
    assert r.stats()["slow_simplices"] == m.stats()["slow_simplices"]
    if dim == 2:
        a, b = m.line_segments(), r.line_segments()       # plot data works on a replica (hull size travels in the header)
        assert all(np.array_equal(u, v) for u, v in zip(a, b))
        # a replica has no host connectivity: it cannot be patched, the caller re-creates it
        assert r.update(np.zeros((0, 2)), [], r.n_simplices, [], np.zeros((0, 3)), np.zeros((0, 3))) is False
    with pytest.raises(dl.DliError):
        r.simplices()          # replicas carry no host connectivity


def test_load_rejects_garbage(gpu, tmp_path):
    import delaunay_linterp_b200 as dl
    p = tmp_path / "bad.dli"
    p.write_bytes(b"not a mesh snapshot at all" * 10)
    with pytest.raises(dl.DliError):
        dl.Mesh.load(p)
    with pytest.raises(dl.DliError):
        dl.Mesh.load(tmp_path / "missing.dli")
    # truncated file
    pts, f, sv, sn = helpers.random_mesh(2, 500, 3)
    m = dl.Mesh(pts, f, sv, sn)
    good = tmp_path / "good.dli"
    m.save(good)
    data = good.read_bytes()
    (tmp_path / "trunc.dli").write_bytes(data[: len(data) // 2])
    with pytest.raises(dl.DliError):
        dl.Mesh.load(tmp_path / "trunc.dli")
    # damage inside the DERIVED buffers (walk records: the neighbour ids the fast pass follows first) is caught
    # by the per-buffer checksums -- flip one byte at a few places across the file, every copy must be refused
    assert dl.Mesh.load(good).n_simplices == m.n_simplices
    for frac in (0.3, 0.55, 0.75, 0.9, 0.98):
        bad = bytearray(data)
        bad[int(len(bad) * frac)] ^= 0x40
        (tmp_path / "flip.dli").write_bytes(bytes(bad))
        with pytest.raises(dl.DliError):
            dl.Mesh.load(tmp_path / "flip.dli")


@pytest.mark.parametrize("opts", [{"outside_nan": 1}, {"epilogue": "closed_form"}, {"outside_nan": 1, "epilogue": "reference"},
                                  {"max_walk_steps": 3}])
def test_replica_and_snapshot_answer_like_their_source(gpu, tmp_path, opts):
    """the options that change what a query returns travel in the mesh header, so a replica built with
    default options (what dli_mesh_replicate does on the receiving ranks) and a snapshot loaded with default options
    answer like the mesh they came from.  1-GPU form of the replicate path: header -> dli_mesh_create_from_header ->
    dli_mesh_buffer_read / _write through a device bounce buffer (what parallel.replicate_mesh does between ranks)."""
    import torch
    import delaunay_linterp_b200 as dl
    pts, f, sv, sn = helpers.random_mesh(2, 3000, 47)
    q = np.random.default_rng(8).random((80000, 2)) * 1.3 - 0.15
    m = dl.Mesh(pts, f, sv, sn, **opts)
    header, gp = m.header()
    r = dl.Mesh.from_header(header, gp)                      # no options on the receiving side
    assert len(r.buffers()) == len(m.buffers())
    for i, ((_, nb), (_, nb_r)) in enumerate(zip(m.buffers(), r.buffers())):
        assert nb == nb_r, "buffer %d: %d vs %d bytes" % (i, nb, nb_r)
        if nb == 0:
            continue
        buf = torch.empty(nb, dtype=torch.uint8, device="cuda")
        m.buffer_read(i, buf.data_ptr())
        r.buffer_write(i, buf.data_ptr())
    torch.cuda.synchronize()
    path = tmp_path / "mesh.dli"
    m.save(path)
    s = dl.Mesh.load(path)                                   # default options here too
    y0, s0, t0 = m.interp(q, return_simplex=True, return_status=True)
    if "outside_nan" in opts:
        assert (t0 == dl.STATUS_OUTSIDE_NAN).sum() > 5000
    if "max_walk_steps" in opts:
        assert (t0 == dl.STATUS_STEP_LIMIT).sum() > 1000
    for other in (r, s):
        y1, s1, t1 = other.interp(q, return_simplex=True, return_status=True)
        assert np.array_equal(t0, t1) and np.array_equal(s0, s1)
        assert np.array_equal(np.isnan(y0), np.isnan(y1)) and np.array_equal(y0[np.isfinite(y0)], y1[np.isfinite(y1)])
