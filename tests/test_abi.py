"""CPU: the C-ABI library loads and exports every symbol include/dli.h declares; host-side helpers work."""
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    names = set()
    for hdr in ("dli.h", "dli_host.h"):
        src = open(os.path.join(ROOT, "include", hdr)).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        names |= set(re.findall(r"\b(dli_[a-z0-9_]+)\s*\(", src))
    names.discard("dli_fn")
    return sorted(names)


def test_library_exports_every_declared_symbol():
    import ctypes
    from delaunay_linterp_b200 import LIB_PATH, _capi
    assert os.path.exists(LIB_PATH), "libdli.so not built (python -m delaunay_linterp_b200.build)"
    L = ctypes.CDLL(LIB_PATH)
    names = _declared_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(L, n), "libdli.so does not export %s" % n
    bound = {s[0] for s in _capi.SYMBOLS + _capi.HOST_SYMBOLS}
    assert set(names) == bound, set(names) ^ bound


def test_options_defaults_and_host_predicate():
    import ctypes as C
    import delaunay_linterp_b200 as dl
    from delaunay_linterp_b200._capi import Options
    o = Options()
    dl.lib().dli_options_init(C.byref(o))
    assert o.struct_size == C.sizeof(Options) and o.device == -1 and o.epilogue == 0
    from oracle import pyoracle
    rng = np.random.default_rng(0)
    for d in (2, 3, 4):
        pts = rng.random((200, d + 1, d))
        w = rng.random((200, d))
        w /= w.sum(1, keepdims=True)
        pts[:, d] = np.einsum("nk,nkd->nd", w, pts[:, :d])     # near-degenerate
        pts[:20, d] = pts[:20, 0]
        sign = dl.debug_orient(pts, on_device=False)
        want = np.array([pyoracle.orient(p.tolist()) for p in pts], dtype=np.int8)
        assert np.array_equal(sign, want)


def test_no_gpu_fails_loudly():
    import delaunay_linterp_b200 as dl
    if dl.lib().dli_device_count() > 0:
        pytest.skip("GPU present")
    with pytest.raises(dl.DliError):
        dl.Mesh(np.array([[0.0, 0.0], [1.0, 0.0], [0.0, 1.0]]), np.zeros(3), np.array([[0, 1, 2]], dtype=np.int32))


def test_product_does_not_touch_the_oracle():
    """the oracle is test infrastructure: nothing under the package or include/ may reference it."""
    bad = []
    for base in ("delaunay_linterp_b200", "include"):
        for dp, _, files in os.walk(os.path.join(ROOT, base)):
            for fn in files:
                if fn.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".hpp")):
                    txt = open(os.path.join(dp, fn), errors="replace").read()
                    if re.search(r"\boracle\b", txt) or "dlo_" in txt:
                        bad.append(os.path.join(dp, fn))
    assert not bad, bad


def test_host_predicate_on_lattice_points_fixed_point_path():
    """exact zeros and near-zeros on regular grids: the exact fallback's fixed-point integer path (coordinates on a
    common 2^g grid, 256-bit determinant) against exact rationals, including inputs that do NOT fit its bit budget."""
    import delaunay_linterp_b200 as dl
    from oracle import pyoracle
    import helpers
    for d in (3, 4):
        pts = helpers.lattice_predicate_inputs(d)
        sign = dl.debug_orient(pts, on_device=False)
        want = np.array([pyoracle.orient(p.tolist()) for p in pts], dtype=np.int8)
        assert np.array_equal(sign, want)
        assert (want == 0).sum() >= 100 and (want != 0).sum() >= 100


def test_host_predicate_2d_projected_points():
    """a point projected on the line of a segment (inexact construction) tested against that segment: the filter
    cannot decide, the differences are not exact -> fixed-point integer path of the 2D exact predicate."""
    import delaunay_linterp_b200 as dl
    from oracle import pyoracle
    rng = np.random.default_rng(9)
    P, Q, X = rng.random((3000, 2)), rng.random((3000, 2)), rng.random((3000, 2)) * 1.2 - 0.1
    d = Q - P
    t = ((X - P) * d).sum(1) / (d * d).sum(1)
    R = P + t[:, None] * d                                   # on the line PQ up to rounding
    pts = np.stack([P, Q, R], axis=1)
    pts[:200, 2] = 0.5 * (P[:200] + Q[:200])                 # midpoints: often exactly collinear
    pts[200:300, 0, 0] = 1e-40                               # out of the fixed-point budget: expansion path
    sign = dl.debug_orient(pts, on_device=False)
    want = np.array([pyoracle.orient(p.tolist()) for p in pts], dtype=np.int8)
    assert np.array_equal(sign, want)
    assert (want == 0).sum() > 20 and (want > 0).sum() > 500 and (want < 0).sum() > 500
