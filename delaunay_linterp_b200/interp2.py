"""DelaunayInterp2: Python mirror of the reference's `_delaunay_2_python.DelaunayInterp2`
(src/delaunay_2_python.cpp:200-231; C++ class Delaunay_incremental_interp_2, src/delaunay_2_interp.h:402-864).

Same method names and argument conventions (a point is a list of floats; an ErrorTuple converts to
(err, f, [x, y]); the callback receives a list), plus the new batched form: interp(ndarray (n, 2)) -> ndarray (n,).
Triangulation construction and the adaptive queue run on the host (native, include/dli_host.h); every interp
goes through the CUDA engine.
"""
import ctypes as C

import numpy as np

from ._capi import DLI_FN, DLI_FN_BATCH, DliError, check, lib
from .mesh import _is_torch_tensor, _make_options


class DelaunayInterp2:
    def __init__(self, fn=None, **options):
        self._fn = fn
        if fn is not None:
            def _tramp(n, x, _user):
                return float(fn([x[i] for i in range(n)]))   # callback receives a list (:189-197)
            self._cb = DLI_FN(_tramp)
        else:
            self._cb = C.cast(None, DLI_FN)
        self._h = C.c_void_p(lib().dli_di2_create(self._cb, None))
        if options:
            opt = _make_options(**options)
            check(lib().dli_di2_set_options(self._h, C.byref(opt)))

    def set_options(self, **options):
        """engine options for the device mesh (see Mesh); the next interp re-flattens the mesh from scratch."""
        opt = _make_options(**options)
        check(lib().dli_di2_set_options(self._h, C.byref(opt)))

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                lib().dli_di2_destroy(self._h)
                self._h = None
        except Exception:
            pass

    @staticmethod
    def _point(x):
        # the reference's converter accepts a list of Python floats only (src/delaunay_2_python.cpp:131-167)
        if len(x) != 2:
            raise TypeError("a point is a list of 2 floats")
        return float(x[0]), float(x[1])

    # ---- construction (host)
    def insert(self, x, f):
        px, py = self._point(x)
        v = lib().dli_di2_insert(self._h, px, py, float(f))
        if v < 0:
            raise DliError(v, "insert: bad point")
        return None

    def insert_many(self, xs, fs, spatial_sort=True, use_gpu=True):
        xs = np.ascontiguousarray(xs, dtype=np.float64).reshape(-1, 2)
        fs = np.ascontiguousarray(fs, dtype=np.float64).reshape(-1)
        assert xs.shape[0] == fs.shape[0]
        check(lib().dli_di2_insert_many(self._h, xs.shape[0], xs.ctypes.data, fs.ctypes.data, int(spatial_sort),
                                        int(use_gpu)))

    def insert_largest_error_point(self):
        e, f, xy = C.c_double(), C.c_double(), (C.c_double * 2)()
        rc = lib().dli_di2_insert_largest_error_point(self._h, C.byref(e), C.byref(f), xy)
        if rc != 0:
            raise AssertionError("insert_largest_error_point: empty queue or no callback")   # the reference asserts
        return (e.value, f.value, [xy[0], xy[1]])

    def insert_largest_error_points(self, k, fn_batch=None, rel_threshold=0.25):
        """k-at-a-time adaptive refinement (dli_di2_refine): up to k largest-error tuples on pairwise non-adjacent faces
        are inserted, then the errors of all new faces are evaluated in one batch.  fn_batch(ndarray (n, 2)) ->
        ndarray (n,) is called ONCE per round; None: the scalar callback is called per item.  A round stops at the first
        tuple whose error is below rel_threshold x the round's largest error.  k = 1 is
        insert_largest_error_point.  Returns the inserted tuples [(err, f, [x, y]), ...]."""
        k = int(k)
        err, f, xy = np.empty(max(k, 1)), np.empty(max(k, 1)), np.empty((max(k, 1), 2))
        if fn_batch is not None:
            def _tramp(nd, n, xs, out, _user):
                a = np.ctypeslib.as_array(xs, shape=(n, nd))
                np.ctypeslib.as_array(out, shape=(n,))[:] = np.asarray(fn_batch(a.copy()), dtype=np.float64).reshape(n)
            cb = DLI_FN_BATCH(_tramp)
        else:
            cb = C.cast(None, DLI_FN_BATCH)
        n = lib().dli_di2_refine(self._h, k, float(rel_threshold), cb, None, err.ctypes.data, f.ctypes.data, xy.ctypes.data)
        if n < 0:
            raise AssertionError("insert_largest_error_points: no callback")
        return [(float(err[i]), float(f[i]), [float(xy[i, 0]), float(xy[i, 1])]) for i in range(n)]

    def get_largest_error_tuple(self):
        e, f, xy = C.c_double(), C.c_double(), (C.c_double * 2)()
        rc = lib().dli_di2_get_largest_error_tuple(self._h, C.byref(e), C.byref(f), xy)
        if rc != 0:
            raise AssertionError("get_largest_error_tuple: empty queue or no callback")
        return (e.value, f.value, [xy[0], xy[1]])

    def get_error_queue(self):
        n = lib().dli_di2_error_queue_size(self._h)
        err, f, xy = np.empty(n), np.empty(n), np.empty((n, 2))
        check(lib().dli_di2_get_error_queue(self._h, err.ctypes.data, f.ctypes.data, xy.ctypes.data))
        return [(float(err[i]), float(f[i]), [float(xy[i, 0]), float(xy[i, 1])]) for i in range(n)]

    # ---- the hot path (GPU)
    def interp(self, x, return_status=False, return_simplex=False):
        """interp([x, y]) -> float (reference form) or interp(ndarray (n, 2) / torch CUDA tensor) -> array.
        return_simplex (numpy form): also the located face, in the numbering of export()."""
        if _is_torch_tensor(x):
            import torch
            xs = x.contiguous().view(-1, 2)
            ys = torch.empty(xs.shape[0], dtype=torch.float64, device=xs.device)
            st = torch.empty(xs.shape[0], dtype=torch.uint8, device=xs.device) if return_status else None
            check(lib().dli_di2_interp(self._h, xs.shape[0], xs.data_ptr(), ys.data_ptr(), None,
                                       st.data_ptr() if st is not None else None,
                                       torch.cuda.current_stream(xs.device).cuda_stream))
            return (ys, st) if return_status else ys
        scalar = isinstance(x, (list, tuple)) and len(x) == 2 and not isinstance(x[0], (list, tuple, np.ndarray))
        xs = np.ascontiguousarray(x, dtype=np.float64).reshape(-1, 2)
        n = xs.shape[0]
        ys = np.empty(n, np.float64)
        st = np.empty(n, np.uint8)
        sx = np.empty(n, np.int32) if return_simplex else None
        check(lib().dli_di2_interp(self._h, n, xs.ctypes.data, ys.ctypes.data, sx.ctypes.data if return_simplex else None,
                                   st.ctypes.data, None))
        if scalar:
            if st[0] >= 3:   # no simplex contains the point: the reference throws std::invalid_argument (:546-551)
                raise ValueError("interp: no simplex containing the point could be found")
            return float(ys[0])
        if return_simplex:
            return (ys, sx, st) if return_status else (ys, sx)
        return (ys, st) if return_status else ys

    def gradient(self, x, return_status=False):
        """gradient([x, y]) -> [gx, gy] (reference form, src/delaunay_2_python.cpp:231 / delaunay_2_interp.h:608-642)
        or gradient(ndarray (n, 2) / torch CUDA tensor) -> (n, 2).  NaN outside the hull."""
        if _is_torch_tensor(x):
            import torch
            xs = x.contiguous().view(-1, 2)
            gs = torch.empty((xs.shape[0], 2), dtype=torch.float64, device=xs.device)
            st = torch.empty(xs.shape[0], dtype=torch.uint8, device=xs.device) if return_status else None
            check(lib().dli_di2_gradient(self._h, xs.shape[0], xs.data_ptr(), gs.data_ptr(), None,
                                         st.data_ptr() if st is not None else None,
                                         torch.cuda.current_stream(xs.device).cuda_stream))
            return (gs, st) if return_status else gs
        scalar = isinstance(x, (list, tuple)) and len(x) == 2 and not isinstance(x[0], (list, tuple, np.ndarray))
        xs = np.ascontiguousarray(x, dtype=np.float64).reshape(-1, 2)
        n = xs.shape[0]
        gs = np.empty((n, 2), np.float64)
        st = np.empty(n, np.uint8)
        check(lib().dli_di2_gradient(self._h, n, xs.ctypes.data, gs.ctypes.data, None, st.ctypes.data, None))
        if scalar:
            return [float(gs[0, 0]), float(gs[0, 1])]
        return (gs, st) if return_status else gs

    # ---- introspection / plot data
    def get_all_vertices(self):
        n = lib().dli_di2_get_all_vertices(self._h, None)
        a = np.empty((n, 3))
        lib().dli_di2_get_all_vertices(self._h, a.ctypes.data)
        return a.tolist()

    def get_line_segments(self):
        """(all_segments, min_coords, max_coords) as the reference returns it (src/delaunay_2_interp.h:842-862,
        delaunay_2.py:65): all_segments = [([x, y, f], [x, y, f], is_concave), ...] with the flag of is_edge_concave
        (:766-803); computed on the GPU (dli_mesh_line_segments)."""
        ne = check(lib().dli_di2_line_segments(self._h, None, None, None, None))
        seg = np.empty((ne, 2, 3))
        cc = np.empty(ne, np.int32)
        mn, mx = np.empty(3), np.empty(3)
        check(lib().dli_di2_line_segments(self._h, seg.ctypes.data, cc.ctypes.data, mn.ctypes.data, mx.ctypes.data))
        s = seg.tolist()
        return [(s[i][0], s[i][1], int(cc[i])) for i in range(ne)], mn.tolist(), mx.tolist()

    def get_edges(self):
        """host-side list of the finite edges [[x, y, f], [x, y, f]] (no GPU, no flags)."""
        n = lib().dli_di2_get_line_segments(self._h, None)
        a = np.empty((n, 2, 3))
        lib().dli_di2_get_line_segments(self._h, a.ctypes.data)
        return a.tolist()

    def export(self):
        """(coords (nv, 2), f (nv,), triangles (nf, 3) ccw, neighbors (nf, 3)) of the current triangulation."""
        nv, nf = lib().dli_di2_num_vertices(self._h), lib().dli_di2_num_faces(self._h)
        c, f = np.empty((nv, 2)), np.empty(nv)
        t, nb = np.empty((nf, 3), np.int32), np.empty((nf, 3), np.int32)
        check(lib().dli_di2_export(self._h, c.ctypes.data, f.ctypes.data, t.ctypes.data if nf else None,
                                   nb.ctypes.data if nf else None))
        return c, f, t, nb

    def sync_counts(self):
        """(full re-flattens, in-place patches) of the device mesh so far: after insert / insert_largest_error_point
        the next interp only re-sends the stars of the inserted points (dli_mesh_update)."""
        a, b = C.c_int64(), C.c_int64()
        lib().dli_di2_sync_counts(self._h, C.byref(a), C.byref(b))
        return a.value, b.value

    @property
    def n_vertices(self):
        return lib().dli_di2_num_vertices(self._h)

    @property
    def n_faces(self):
        return lib().dli_di2_num_faces(self._h)

    @property
    def dimension(self):
        return lib().dli_di2_dimension(self._h)
