"""ctypes loader for the CPU oracle (oracle/dli_oracle.c).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module; the product package never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_port", "libdli_oracle.so")

OK, OUTSIDE_PROJECTED, OUTSIDE_NAN, STEP_LIMIT, NO_MESH, ON_EDGE, ON_VERTEX = 0, 1, 2, 3, 4, 5, 6


def build(force=False):
    src = os.path.join(_HERE, "dli_oracle.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        L = C.CDLL(_LIB_PATH)
        dp, ip, bp = C.POINTER(C.c_double), C.POINTER(C.c_int32), C.POINTER(C.c_uint8)
        L.dlo_mesh_create.restype = C.c_void_p
        L.dlo_mesh_create.argtypes = [C.c_int, C.c_int64, dp, dp, C.c_int64, ip, ip]
        L.dlo_mesh_destroy.argtypes = [C.c_void_p]
        L.dlo_mesh_set_values.argtypes = [C.c_void_p, dp]
        L.dlo_mesh_start.restype = C.c_int32
        L.dlo_mesh_start.argtypes = [C.c_void_p]
        L.dlo_mesh_get_simplices.argtypes = [C.c_void_p, ip, ip]
        L.dlo_interp.argtypes = [C.c_void_p, C.c_int64, dp, dp, ip, bp, C.c_int, C.c_int, C.POINTER(C.c_int64)]
        L.dlo_gradient.argtypes = [C.c_void_p, C.c_int64, dp, dp, ip, bp, C.c_int]
        L.dlo_classify.argtypes = [C.c_void_p, C.c_int64, dp, ip, bp, bp]
        L.dlo_orient_exact.argtypes = [C.c_int, dp]
        L.dlo_orient_filtered.argtypes = [C.c_int, dp]
        L.dlo_error_items.restype = C.c_int64
        L.dlo_error_items.argtypes = [C.c_void_p, ip, ip, ip, dp]
        L.dlo_line_segments.restype = C.c_int64
        L.dlo_line_segments.argtypes = [C.c_void_p, dp, ip, dp, dp]
        L.dlo_error_eval.argtypes = [C.c_void_p, C.c_int64, ip, ip, dp, dp, dp]
        _lib = L
    return _lib


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _ip(a):
    return a.ctypes.data_as(C.POINTER(C.c_int32))


def _bp(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint8))


class OracleMesh:
    """CPU oracle over a flattened simplex mesh (same inputs as dli_mesh_create)."""

    def __init__(self, coords, f, simplices, neighbors=None):
        coords = np.ascontiguousarray(coords, dtype=np.float64)
        f = np.ascontiguousarray(f, dtype=np.float64)
        simplices = np.ascontiguousarray(simplices, dtype=np.int32)
        self.dim = coords.shape[1]
        self.nv = coords.shape[0]
        self.ns = simplices.shape[0]
        assert simplices.shape[1] == self.dim + 1 and f.shape == (self.nv,)
        nb = None
        if neighbors is not None:
            nb = np.ascontiguousarray(neighbors, dtype=np.int32)
            assert nb.shape == simplices.shape
        self._h = lib().dlo_mesh_create(self.dim, self.nv, _dp(coords), _dp(f), self.ns, _ip(simplices),
                                        _ip(nb) if nb is not None else None)
        if not self._h:
            raise ValueError("oracle: invalid mesh (degenerate simplex, bad index or non-manifold facet)")

    def __del__(self):
        if getattr(self, "_h", None):
            lib().dlo_mesh_destroy(self._h)
            self._h = None

    @property
    def start(self):
        return lib().dlo_mesh_start(self._h)

    def simplices(self):
        sv = np.empty((self.ns, self.dim + 1), np.int32)
        sn = np.empty((self.ns, self.dim + 1), np.int32)
        lib().dlo_mesh_get_simplices(self._h, _ip(sv), _ip(sn))
        return sv, sn

    def set_values(self, f):
        f = np.ascontiguousarray(f, dtype=np.float64)
        assert f.shape == (self.nv,)
        lib().dlo_mesh_set_values(self._h, _dp(f))

    def interp(self, xs, mode=1, nthreads=1, return_steps=False):
        """mode 0 = reference-faithful un-hinted walk; mode 1 = walk from the previous result."""
        xs = np.ascontiguousarray(xs, dtype=np.float64).reshape(-1, self.dim)
        n = xs.shape[0]
        ys = np.empty(n, np.float64)
        simplex = np.empty(n, np.int32)
        status = np.empty(n, np.uint8)
        steps = C.c_int64(0)
        lib().dlo_interp(self._h, n, _dp(xs), _dp(ys), _ip(simplex), _bp(status), mode, nthreads, C.byref(steps))
        if return_steps:
            return ys, simplex, status, steps.value
        return ys, simplex, status

    def gradient(self, xs, mode=1):
        """2D gradient queries (src/delaunay_2_interp.h:608-642): (grads (n, 2), simplex, status)."""
        xs = np.ascontiguousarray(xs, dtype=np.float64).reshape(-1, 2)
        n = xs.shape[0]
        gs = np.empty((n, 2), np.float64)
        simplex = np.empty(n, np.int32)
        status = np.empty(n, np.uint8)
        rc = lib().dlo_gradient(self._h, n, _dp(xs), _dp(gs), _ip(simplex), _bp(status), mode)
        if rc != 0:
            raise ValueError("oracle gradient: 2D meshes only")
        return gs, simplex, status

    def classify(self, xs, simplex):
        """(contains, strictly_inside) of each query w.r.t. the given simplex ids; exact."""
        xs = np.ascontiguousarray(xs, dtype=np.float64).reshape(-1, self.dim)
        simplex = np.ascontiguousarray(simplex, dtype=np.int32)
        n = xs.shape[0]
        c = np.empty(n, np.uint8)
        s = np.empty(n, np.uint8)
        lib().dlo_classify(self._h, n, _dp(xs), _ip(simplex), _bp(c), _bp(s))
        return c.astype(bool), s.astype(bool)

    def line_segments(self):
        """get_line_segments (src/delaunay_2_interp.h:842-862) + is_edge_concave (:766-803):
        (segments (ne, 2, 3), concave (ne,), min (3,), max (3,))."""
        ne = lib().dlo_line_segments(self._h, None, None, None, None)
        if ne < 0:
            raise ValueError("oracle line_segments: 2D meshes only")
        seg = np.empty((ne, 2, 3), np.float64)
        cc = np.empty(ne, np.int32)
        mn, mx = np.empty(3), np.empty(3)
        lib().dlo_line_segments(self._h, _dp(seg), _ip(cc), _dp(mn), _dp(mx))
        return seg, cc, mn, mx

    def error_items(self):
        n = lib().dlo_error_items(self._h, None, None, None, None)
        k = self.dim + 1
        item_simplex = np.empty(n, np.int32)
        item_nv = np.empty(n, np.int32)
        item_verts = np.empty((n, k), np.int32)
        cent = np.empty((n, self.dim), np.float64)
        lib().dlo_error_items(self._h, _ip(item_simplex), _ip(item_nv), _ip(item_verts), _dp(cent))
        return item_simplex, item_nv, item_verts, cent

    def error_eval(self, item_nv, item_verts, centroids, phi):
        n = item_nv.shape[0]
        phi = np.ascontiguousarray(phi, dtype=np.float64)
        err = np.empty(n, np.float64)
        lib().dlo_error_eval(self._h, n, _ip(item_nv), _ip(item_verts), _dp(centroids), _dp(phi), _dp(err))
        return err


def orient_exact(pts):
    pts = np.ascontiguousarray(pts, dtype=np.float64)
    return lib().dlo_orient_exact(pts.shape[1], _dp(pts))


def orient_filtered(pts):
    pts = np.ascontiguousarray(pts, dtype=np.float64)
    return lib().dlo_orient_filtered(pts.shape[1], _dp(pts))
