"""Delaunay_incremental_interp_d: Python mirror of the reference's `_delaunay_d_python.Delaunay_incremental_interp_d`
(src/delaunay_d_python.cpp:364-373; C++ class src/delaunay_d_interp.h:753-1142) for d = 2, 3, 4.

Same method names and conventions (points are lists of floats, ErrorTuple -> (err, f, [x...]), interp outside the
hull returns NaN, :1005-1008), plus batched interp(ndarray (n, d)) -> ndarray (n,) on the GPU.

Host side: the reference builds on the patched CGAL Delaunay_d (modified_cgal/Delaunay_d.h), which is unavailable;
the d-dimensional triangulation and the adaptive queue are native (csrc/dli_hostmesh_d.cpp behind the `dli_dd_*` C ABI
of include/dli_host.h -- the same object the C++ drop-in header include/delaunay_d_interp.h wraps).
"""
import ctypes as C

import numpy as np

from ._capi import DLI_FN, DLI_FN_BATCH, DliError, check, lib
from .mesh import _is_torch_tensor, _make_options


class Delaunay_incremental_interp_d:
    def __init__(self, n_dims, fn=None, **options):
        if not 2 <= n_dims <= 4:
            raise ValueError("n_dims must be 2, 3 or 4")
        self.N = int(n_dims)
        self.fn = fn                                  # read-only attribute of the reference binding (:366)
        if fn is not None:
            def _tramp(n, x, _user):
                return float(fn([x[i] for i in range(n)]))   # callback receives a list (src/delaunay_d_python.cpp:233-241)
            self._cb = DLI_FN(_tramp)
        else:
            self._cb = C.cast(None, DLI_FN)
        self._h = C.c_void_p(lib().dli_dd_create(self.N, self._cb, None))
        if not self._h:
            raise ValueError("dli_dd_create failed")
        if options:
            opt = _make_options(**options)
            check(lib().dli_dd_set_options(self._h, C.byref(opt)))

    def set_options(self, **options):
        """engine options for the device mesh (see Mesh); the next interp re-flattens the mesh from scratch."""
        opt = _make_options(**options)
        check(lib().dli_dd_set_options(self._h, C.byref(opt)))

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                lib().dli_dd_destroy(self._h)
                self._h = None
        except Exception:
            pass

    @property
    def triang(self):
        """the reference exposes its CGAL Delaunay_d here (:365); this object plays both roles"""
        return self

    def get_dimension(self):
        return self.N

    def dimension(self):
        return self.N

    def current_dimension(self):
        return lib().dli_dd_current_dimension(self._h)

    def number_of_vertices(self):
        return lib().dli_dd_num_vertices(self._h)

    def number_of_simplices(self):
        return lib().dli_dd_num_simplices(self._h)

    # ---- construction (host)
    def _point(self, x):
        if len(x) != self.N:
            raise TypeError("a point is a list of %d floats" % self.N)
        return (C.c_double * self.N)(*[float(v) for v in x])

    def insert(self, x, f):
        v = lib().dli_dd_insert(self._h, self._point(x), float(f))
        if v < 0:
            raise DliError(v, "insert: bad point")
        return None

    def insert_many(self, xs, fs, spatial_sort=True, use_gpu=True):
        xs = np.ascontiguousarray(xs, dtype=np.float64).reshape(-1, self.N)
        fs = np.ascontiguousarray(fs, dtype=np.float64).reshape(-1)
        assert xs.shape[0] == fs.shape[0]
        check(lib().dli_dd_insert_many(self._h, xs.shape[0], xs.ctypes.data, fs.ctypes.data, int(spatial_sort),
                                       int(use_gpu)))

    def validate(self, check_delaunay=True):
        """0 when the triangulation is structurally valid (and Delaunay); test hook (dli_dd_validate)."""
        return lib().dli_dd_validate(self._h, int(check_delaunay))

    def get_largest_error_tuple(self):
        e, f, x = C.c_double(), C.c_double(), (C.c_double * self.N)()
        rc = lib().dli_dd_get_largest_error_tuple(self._h, C.byref(e), C.byref(f), x)
        assert rc == 0, "empty queue or no callback"          # the reference asserts (:1029-1030)
        return (e.value, f.value, list(x))

    def insert_largest_error_point(self):
        e, f, x = C.c_double(), C.c_double(), (C.c_double * self.N)()
        rc = lib().dli_dd_insert_largest_error_point(self._h, C.byref(e), C.byref(f), x)
        if rc != 0:
            raise AssertionError("insert_largest_error_point: empty queue or no callback")
        return (e.value, f.value, list(x))

    def get_error_queue(self):
        cap = lib().dli_dd_error_queue_size(self._h)
        err, f, xs = np.empty(max(cap, 1)), np.empty(max(cap, 1)), np.empty((max(cap, 1), self.N))
        n = lib().dli_dd_get_error_queue(self._h, cap, err.ctypes.data, f.ctypes.data, xs.ctypes.data)
        if n < 0:
            raise DliError(n, "get_error_queue")
        return [(float(err[i]), float(f[i]), xs[i].tolist()) for i in range(n)]

    def insert_largest_error_points(self, k, fn_batch=None, rel_threshold=0.25):
        """k-at-a-time adaptive refinement (dli_dd_refine): up to k largest-error tuples on simplices that share no
        vertex are inserted, then the items of all new simplices are evaluated in one batch.  fn_batch(ndarray (n, d))
        -> ndarray (n,) is called ONCE per round; None: the scalar callback per item.  k = 1 is
        insert_largest_error_point (with one batched callback instead of hundreds of scalar ones).  Returns the
        inserted tuples [(err, f, [x...]), ...]."""
        k = int(k)
        err, f, xs = np.empty(max(k, 1)), np.empty(max(k, 1)), np.empty((max(k, 1), self.N))
        if fn_batch is not None:
            def _tramp(nd, n, x, out, _user):
                a = np.ctypeslib.as_array(x, shape=(n, nd))
                np.ctypeslib.as_array(out, shape=(n,))[:] = np.asarray(fn_batch(a.copy()), dtype=np.float64).reshape(n)
            cb = DLI_FN_BATCH(_tramp)
        else:
            cb = C.cast(None, DLI_FN_BATCH)
        n = lib().dli_dd_refine(self._h, k, float(rel_threshold), cb, None, err.ctypes.data, f.ctypes.data, xs.ctypes.data)
        if n < 0:
            raise AssertionError("insert_largest_error_points: no callback")
        return [(float(err[i]), float(f[i]), xs[i].tolist()) for i in range(n)]

    def rebuild_error_queue(self):
        """recompute every tuple of the current triangulation in one batch on the GPU (dli_simplex_error_*)."""
        assert self.fn is not None
        check(lib().dli_dd_insert_many(self._h, 0, None, None, 0, 1))

    # ---- the hot path (GPU; one point at a time is answered on the host, see include/dli_host.h)
    def sync_counts(self):
        """(full re-flattens, in-place patches) of the device mesh so far: after insert / insert_largest_error_point the
        next interp re-sends only the simplices the insertions touched (dli_mesh_update)."""
        a, b = C.c_int64(), C.c_int64()
        lib().dli_dd_sync_counts(self._h, C.byref(a), C.byref(b))
        return a.value, b.value

    def interp(self, x, return_status=False, return_simplex=False):
        """interp([x...]) -> float (NaN outside the hull) or interp(ndarray (n, d) / torch CUDA tensor) -> array.
        return_simplex (numpy form): also the located simplex, in the numbering of export()."""
        if _is_torch_tensor(x):
            import torch
            xs = x.contiguous().view(-1, self.N)
            ys = torch.empty(xs.shape[0], dtype=torch.float64, device=xs.device)
            st = torch.empty(xs.shape[0], dtype=torch.uint8, device=xs.device) if return_status else None
            check(lib().dli_dd_interp(self._h, xs.shape[0], xs.data_ptr(), ys.data_ptr(), None,
                                      st.data_ptr() if st is not None else None,
                                      torch.cuda.current_stream(xs.device).cuda_stream))
            return (ys, st) if return_status else ys
        scalar = isinstance(x, (list, tuple)) and len(x) == self.N and not isinstance(x[0], (list, tuple, np.ndarray))
        xs = np.ascontiguousarray(x, dtype=np.float64).reshape(-1, self.N)
        n = xs.shape[0]
        ys = np.empty(n, np.float64)
        st = np.empty(n, np.uint8)
        sx = np.empty(n, np.int32) if return_simplex else None
        check(lib().dli_dd_interp(self._h, n, xs.ctypes.data, ys.ctypes.data, sx.ctypes.data if return_simplex else None,
                                  st.ctypes.data, None))
        if scalar:
            return float(ys[0])
        if return_simplex:
            return (ys, sx, st) if return_status else (ys, sx)
        return (ys, st) if return_status else ys

    # ---- plot data
    def get_all_vertices(self):
        nv = lib().dli_dd_get_all_vertices(self._h, None)
        out = np.empty((max(nv, 1), self.N + 1))
        lib().dli_dd_get_all_vertices(self._h, out.ctypes.data)
        return out[:nv].tolist()

    def get_line_segments(self):
        """(all_segments, min_coords, max_coords) as the reference returns it (src/delaunay_d_interp.h:1115-1142): the
        segments of a simplex are its consecutive vertex pairs (i, i+1 mod N+1) (:1103-1112), every pair once, flag 1."""
        n1 = self.N + 1
        ne = lib().dli_dd_get_line_segments(self._h, None, None, None)
        if ne < 0:
            raise DliError(ne, "get_line_segments")
        seg = np.empty((max(ne, 1), 2, n1))
        lo, hi = np.empty(n1), np.empty(n1)
        lib().dli_dd_get_line_segments(self._h, seg.ctypes.data, lo.ctypes.data, hi.ctypes.data)
        return [(seg[i, 0].tolist(), seg[i, 1].tolist(), 1) for i in range(ne)], lo.tolist(), hi.tolist()

    def export(self):
        nv, ns, k = self.number_of_vertices(), self.number_of_simplices(), self.N + 1
        pts, f = np.empty((nv, self.N)), np.empty(nv)
        sv, sn = np.empty((ns, k), np.int32), np.empty((ns, k), np.int32)
        check(lib().dli_dd_export(self._h, pts.ctypes.data if nv else None, f.ctypes.data if nv else None,
                                  sv.ctypes.data if ns else None, sn.ctypes.data if ns else None))
        return pts, f, sv, sn


def get_lifted_point(x):
    """free function of the reference module (src/delaunay_d_interp.h:1145-1149): (x, |x|^2)."""
    x = [float(v) for v in x]
    return x + [sum(v * v for v in x)]
